// arrow_ops.cpp — streaming operators over host Arrow batches (push / pull):
// the C-ABI form of `ExecutionPlan::execute() -> BoxStream<RecordBatch>`
// (crates/engine/src/physical_plan.rs:10-17) for
//   FilterExec / ProjectionExec   (operators/filter.rs:39-68, projection.rs:48-72)
//   HashJoinExec                  (operators/hash_join.rs:100-213)
//
// Host batches arrive as Arrow C Data Interface struct arrays (what arrow-rs
// `arrow::ffi::to_ffi` exports).  The reference's upstream granularity is
// 1024-row batches (parquet_scan.rs:54); they are coalesced in a pinned staging
// chunk (igl_config.coalesce_rows, default 4 Mi rows), uploaded with
// cudaMemcpyAsync, run through the device operators and copied back into pinned
// buffers that are handed out as Arrow arrays whose release callback recycles
// them.  Row multisets, schema and error behaviour follow the reference; batch
// boundaries follow the coalescing (one output batch per chunk; a FilterExec
// chunk that keeps no row emits nothing, filter.rs:59-63).
#include <algorithm>
#include <chrono>
#include <cstring>
#include <deque>
#include <memory>
#include <thread>

#include "internal.h"
#include "kernels.h"

using namespace igl;

namespace {

// ---- process-wide pinned buffer pool (outlives any context: exported arrays
//      may be released after their operator and context are gone) -------------
struct PinnedPool {
    std::mutex mu;
    std::map<size_t, std::vector<void*>> free_lists;
    std::map<void*, size_t> live;
    static size_t bucket(size_t b) { size_t s = 4096; while (s < b) s <<= 1; return s; }
    void* get(size_t bytes) {
        const size_t b = bucket(bytes ? bytes : 1);
        {
            std::lock_guard<std::mutex> g(mu);
            auto it = free_lists.find(b);
            if (it != free_lists.end() && !it->second.empty()) {
                void* p = it->second.back(); it->second.pop_back(); live[p] = b; return p;
            }
        }
        void* p = nullptr;
        if (cudaHostAlloc(&p, b, cudaHostAllocDefault) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
        std::lock_guard<std::mutex> g(mu);
        live[p] = b;
        return p;
    }
    void put(void* p) {
        if (!p) return;
        std::lock_guard<std::mutex> g(mu);
        auto it = live.find(p);
        if (it == live.end()) return;
        const size_t b = it->second;
        live.erase(it);
        if (b > (256ull << 20)) { cudaFreeHost(p); return; }   // keep the pool modest
        free_lists[b].push_back(p);
    }
};
PinnedPool& pinned_pool() { static PinnedPool* p = new PinnedPool(); return *p; }

struct FieldMeta { std::string name; int type; bool nullable; };

// IGL_OP_TRACE=1: host wall-clock spent per step of the streaming operators, printed when an
// operator is destroyed (where does a chunk's time go: staging, waiting for the GPU, copies?).
struct OpTrace {
    bool on = getenv("IGL_OP_TRACE") && getenv("IGL_OP_TRACE")[0] == '1';
    double t[8] = {};
    long n[8] = {};
    static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
};
struct TraceScope {
    OpTrace& T; int i; double t0;
    TraceScope(OpTrace& t, int idx) : T(t), i(idx), t0(t.on ? OpTrace::now() : 0.0) {}
    ~TraceScope() { if (T.on) { T.t[i] += OpTrace::now() - t0; T.n[i]++; } }
};
enum { TR_STAGE = 0, TR_END = 1, TR_UPLOAD = 2, TR_BEGIN = 3, TR_WAIT_UP = 4, TR_PULL = 5, TR_BUILD = 6, TR_PUSH = 7 };

int type_of_format(const char* f) {
    if (!f) return 0;
    if (!strcmp(f, "b")) return IGL_T_BOOL;
    if (!strcmp(f, "i")) return IGL_T_I32;
    if (!strcmp(f, "l")) return IGL_T_I64;
    if (!strcmp(f, "g")) return IGL_T_F64;
    if (!strcmp(f, "u")) return IGL_T_UTF8;
    return 0;
}
const char* format_of_type(int t) {
    switch (t) { case IGL_T_BOOL: return "b"; case IGL_T_I32: return "i"; case IGL_T_I64: return "l";
                 case IGL_T_F64: return "g"; case IGL_T_UTF8: return "u"; }
    return "n";
}

int32_t import_schema(const ArrowSchema* s, std::vector<FieldMeta>& out) {
    if (!s || !s->format || strcmp(s->format, "+s")) return fail(IGL_ERR_INVALID_ARG, "schema must be a struct (\"+s\") of columns");
    for (int64_t i = 0; i < s->n_children; ++i) {
        const ArrowSchema* c = s->children[i];
        const int t = type_of_format(c->format);
        if (!t) return fail(IGL_ERR_UNSUPPORTED, "Arrow format \"%s\" of column %lld is not supported (Boolean, Int32, Int64, Float64, Utf8)",
                            c->format ? c->format : "?", (long long)i);
        out.push_back({c->name ? c->name : "", t, (c->flags & ARROW_FLAG_NULLABLE) != 0});
    }
    return IGL_OK;
}

void copy_bits(uint8_t* dst, int64_t dst_off, const uint8_t* src, int64_t src_off, int64_t n) {
    if (n <= 0) return;
    if ((dst_off & 7) == 0 && (src_off & 7) == 0) {
        memcpy(dst + (dst_off >> 3), src + (src_off >> 3), (size_t)(n >> 3));
        for (int64_t i = n & ~7ll; i < n; ++i) {
            const int b = (src[(src_off + i) >> 3] >> ((src_off + i) & 7)) & 1;
            uint8_t& d = dst[(dst_off + i) >> 3];
            d = (uint8_t)((d & ~(1u << ((dst_off + i) & 7))) | (b << ((dst_off + i) & 7)));
        }
        return;
    }
    for (int64_t i = 0; i < n; ++i) {
        const int b = (src[(src_off + i) >> 3] >> ((src_off + i) & 7)) & 1;
        uint8_t& d = dst[(dst_off + i) >> 3];
        d = (uint8_t)((d & ~(1u << ((dst_off + i) & 7))) | (b << ((dst_off + i) & 7)));
    }
}
int64_t count_zero_bits(const uint8_t* bm, int64_t n) {
    int64_t ones = 0;
    for (int64_t i = 0; i < (n >> 3); ++i) ones += __builtin_popcount(bm[i]);
    for (int64_t i = n & ~7ll; i < n; ++i) ones += (bm[i >> 3] >> (i & 7)) & 1;
    return n - ones;
}
size_t pad64(size_t b) { return (b + 63) / 64 * 64 + 64; }

// Host batches are pageable memory: every input byte is copied once into the pinned staging
// chunk, and one core moves ~10 GB/s while PCIe takes 50.  Copies of 4 MB and more are therefore
// split over up to 16 threads (IGL_STAGE_THREADS overrides; 1 = plain memcpy).
void staged_copy(uint8_t* dst, const uint8_t* src, size_t bytes) {
    static const unsigned max_threads = [] {
        if (const char* e = getenv("IGL_STAGE_THREADS")) return (unsigned)std::max(1, atoi(e));
        const unsigned hw = std::thread::hardware_concurrency();
        return std::min(16u, hw ? hw : 1u);
    }();
    const size_t kMinSlice = (size_t)2 << 20;
    unsigned nt = (unsigned)std::min<size_t>(max_threads, bytes / kMinSlice);
    if (nt < 2) { memcpy(dst, src, bytes); return; }
    const size_t slice = ((bytes + nt - 1) / nt + 63) & ~(size_t)63;
    std::vector<std::thread> th;
    th.reserve(nt - 1);
    for (unsigned t = 1; t < nt; ++t) {
        const size_t b = (size_t)t * slice;
        if (b >= bytes) break;
        th.emplace_back([=] { memcpy(dst + b, src + b, std::min(slice, bytes - b)); });
    }
    memcpy(dst, src, std::min(slice, bytes));
    for (auto& x : th) x.join();
}

// ---- one side's staging: pinned host chunk + device copy ---------------------
struct StagedCol {
    int type = 0; bool nullable = false;
    uint8_t* h_data = nullptr; size_t h_data_cap = 0;   // values / utf8 bytes (pinned)
    uint8_t* h_valid = nullptr;                          // bitmap for chunk_rows (pinned)
    int32_t* h_off = nullptr;                            // utf8 offsets (pinned)
    size_t utf8_used = 0;
    bool any_null = false;                               // a batch staged into this chunk carried nulls for this column
    void* d_data = nullptr; size_t d_data_cap = 0;
    uint8_t* d_valid = nullptr; int32_t* d_off = nullptr;
};

struct Staging {
    igl_ctx* ctx = nullptr;
    std::vector<FieldMeta> fields;
    std::vector<StagedCol> cols;
    std::vector<uint8_t> wanted;   // which columns are staged at all
    int64_t chunk_rows = 0, rows = 0;
    // join sides: a nullable column whose chunk holds no null goes to the device without a bitmap, so
    // the kernels for non-null data serve it (Arrow producers mark almost every field nullable)
    bool elide_all_valid = false;

    int32_t init(igl_ctx* c, const std::vector<FieldMeta>& f, int64_t chunk) {
        ctx = c; fields = f; chunk_rows = chunk; cols.resize(f.size()); wanted.assign(f.size(), 1);
        for (size_t i = 0; i < f.size(); ++i) { cols[i].type = f[i].type; cols[i].nullable = f[i].nullable; }
        return IGL_OK;
    }
    int32_t ensure_host(StagedCol& c, size_t data_bytes) {
        if (data_bytes > c.h_data_cap) {
            size_t cap = std::max<size_t>(data_bytes, c.h_data_cap * 2);
            uint8_t* p = (uint8_t*)pinned_pool().get(pad64(cap));
            if (!p) return fail(IGL_ERR_OOM, "pinned staging allocation of %zu bytes failed", cap);
            if (c.h_data) { memcpy(p, c.h_data, c.type == IGL_T_UTF8 ? c.utf8_used : c.h_data_cap); pinned_pool().put(c.h_data); }
            c.h_data = p; c.h_data_cap = cap;
        }
        if (c.nullable && !c.h_valid) {
            c.h_valid = (uint8_t*)pinned_pool().get(pad64((size_t)(chunk_rows + 7) / 8));
            if (!c.h_valid) return fail(IGL_ERR_OOM, "pinned staging allocation failed");
            memset(c.h_valid, 0, pad64((size_t)(chunk_rows + 7) / 8));
        }
        if (c.type == IGL_T_UTF8 && !c.h_off) {
            c.h_off = (int32_t*)pinned_pool().get(pad64((size_t)(chunk_rows + 1) * 4));
            if (!c.h_off) return fail(IGL_ERR_OOM, "pinned staging allocation failed");
            c.h_off[0] = 0;
        }
        return IGL_OK;
    }
    // Appends rows [from, from + n) of the batch.
    int32_t append(const ArrowArray* batch, int64_t from, int64_t n) {
        if ((size_t)batch->n_children != fields.size())
            return fail(IGL_ERR_INVALID_ARG, "batch has %lld columns, schema has %zu", (long long)batch->n_children, fields.size());
        for (size_t j = 0; j < fields.size(); ++j) {
            if (!wanted[j]) continue;
            StagedCol& c = cols[j];
            const ArrowArray* a = batch->children[j];
            const int64_t off = a->offset + batch->offset + from;
            const uint8_t* valid = (const uint8_t*)a->buffers[0];
            if (c.type == IGL_T_UTF8) {
                if (a->n_buffers < 3) return fail(IGL_ERR_INVALID_ARG, "Utf8 column %zu needs 3 buffers", j);
                const int32_t* so = (const int32_t*)a->buffers[1];
                const uint8_t* sb = (const uint8_t*)a->buffers[2];
                const int32_t b0 = so[off], b1 = so[off + n];
                IGL_TRY(ensure_host(c, c.utf8_used + (size_t)(b1 - b0)));
                if (b1 > b0) staged_copy(c.h_data + c.utf8_used, sb + b0, (size_t)(b1 - b0));
                const int32_t base = (int32_t)c.utf8_used - b0;
                for (int64_t i = 1; i <= n; ++i) c.h_off[rows + i] = so[off + i] + base;
                c.utf8_used += (size_t)(b1 - b0);
            } else if (c.type == IGL_T_BOOL) {
                IGL_TRY(ensure_host(c, (size_t)(rows + n + 7) / 8 + 8));
                copy_bits(c.h_data, rows, (const uint8_t*)a->buffers[1], off, n);
            } else {
                const int w = type_width(c.type);
                IGL_TRY(ensure_host(c, (size_t)(rows + n) * w));
                staged_copy(c.h_data + (size_t)rows * w, (const uint8_t*)a->buffers[1] + (size_t)off * w, (size_t)n * w);
            }
            if (c.nullable) {
                if (valid && a->null_count != 0) { copy_bits(c.h_valid, rows, valid, off, n); c.any_null = true; }
                else { // clear then set: the target bytes may hold stale bits from the previous chunk
                    for (int64_t i = 0; i < n; ++i) c.h_valid[(rows + i) >> 3] |= (uint8_t)(1u << ((rows + i) & 7));
                }
            } else if (valid && a->null_count > 0) {
                return fail(IGL_ERR_INVALID_ARG, "column %zu is declared non-nullable but the batch carries nulls", j);
            }
        }
        rows += n;
        return IGL_OK;
    }
    // Uploads the staged chunk; device pointers valid until the next upload.
    int32_t upload(cudaStream_t st, std::vector<igl_dev_col>& dev) {
        dev.assign(fields.size(), igl_dev_col{});
        for (size_t j = 0; j < fields.size(); ++j) {
            dev[j].type = fields[j].type;
            if (!wanted[j]) continue;
            StagedCol& c = cols[j];
            size_t bytes = c.type == IGL_T_UTF8 ? c.utf8_used : c.type == IGL_T_BOOL ? (size_t)(rows + 7) / 8 : (size_t)rows * type_width(c.type);
            if (pad64(bytes) > c.d_data_cap) {
                if (c.d_data) pool_free(ctx, c.d_data);
                c.d_data = nullptr; c.d_data_cap = 0;
                IGL_TRY(pool_alloc(ctx, pad64(bytes), &c.d_data));
                c.d_data_cap = pad64(bytes);
            }
            if (bytes) IGL_CUDA_OK(cudaMemcpyAsync(c.d_data, c.h_data, bytes, cudaMemcpyHostToDevice, st));
            dev[j].data = c.d_data;
            if (c.nullable && (c.any_null || !elide_all_valid)) {
                if (!c.d_valid) { void* p = nullptr; IGL_TRY(pool_alloc(ctx, pad64((size_t)(chunk_rows + 7) / 8), &p)); c.d_valid = (uint8_t*)p; }
                IGL_CUDA_OK(cudaMemcpyAsync(c.d_valid, c.h_valid, (size_t)(rows + 7) / 8, cudaMemcpyHostToDevice, st));
                dev[j].validity = c.d_valid;
            }
            if (c.type == IGL_T_UTF8) {
                if (!c.d_off) { void* p = nullptr; IGL_TRY(pool_alloc(ctx, pad64((size_t)(chunk_rows + 1) * 4), &p)); c.d_off = (int32_t*)p; }
                IGL_CUDA_OK(cudaMemcpyAsync(c.d_off, c.h_off, (size_t)(rows + 1) * 4, cudaMemcpyHostToDevice, st));
                dev[j].offsets = c.d_off;
            }
        }
        return IGL_OK;
    }
    void reset() {
        rows = 0;
        for (auto& c : cols) {
            c.utf8_used = 0;
            c.any_null = false;
            if (c.h_valid) memset(c.h_valid, 0, (size_t)(chunk_rows + 7) / 8);
        }
    }
    void destroy() {
        for (auto& c : cols) {
            pinned_pool().put(c.h_data); pinned_pool().put(c.h_valid); pinned_pool().put(c.h_off);
            pool_free(ctx, c.d_data); pool_free(ctx, c.d_valid); pool_free(ctx, c.d_off);
            c = StagedCol{};
        }
    }
};

// ---- output batches ------------------------------------------------------------
struct OutCol {
    int type = 0; void* data = nullptr; uint8_t* valid = nullptr; int32_t* off = nullptr; int64_t null_count = 0;
    bool alias = false;   // `data` belongs to another column of the same batch (two columns, one host buffer)
};
struct OutBatch {
    int64_t nrows = 0;
    std::vector<OutCol> cols;
    cudaEvent_t done = nullptr;        // join: recorded behind the batch's device->host copies (download stream)
    std::vector<void*> dev_free;       // device output blocks, returned to the pool once `done` has passed
};

struct ExportHolder {
    std::vector<void*> pinned;
    std::vector<ArrowArray> children;
    std::vector<ArrowArray*> child_ptrs;
    std::vector<std::vector<const void*>> buffers;
    const void* top_buffers[1] = {nullptr};
};
void release_child(ArrowArray* a) { a->release = nullptr; }
void release_top(ArrowArray* a) {
    ExportHolder* h = (ExportHolder*)a->private_data;
    for (auto& c : h->children) if (c.release) c.release(&c);
    for (void* p : h->pinned) pinned_pool().put(p);
    delete h;
    a->release = nullptr;
}
void export_batch(OutBatch&& b, ArrowArray* out) {
    ExportHolder* h = new ExportHolder();
    const size_t n = b.cols.size();
    h->children.resize(n); h->child_ptrs.resize(n); h->buffers.resize(n);
    for (size_t j = 0; j < n; ++j) {
        OutCol& c = b.cols[j];
        ArrowArray& a = h->children[j];
        memset(&a, 0, sizeof a);
        a.length = b.nrows; a.null_count = c.null_count; a.offset = 0;
        if (c.type == IGL_T_UTF8) h->buffers[j] = {c.null_count ? c.valid : nullptr, c.off, c.data};
        else h->buffers[j] = {c.null_count ? c.valid : nullptr, c.data};
        a.n_buffers = (int64_t)h->buffers[j].size();
        a.buffers = h->buffers[j].data();
        a.release = release_child;
        h->child_ptrs[j] = &a;
        if (!c.alias) h->pinned.push_back(c.data);
        h->pinned.push_back(c.valid); h->pinned.push_back(c.off);
    }
    memset(out, 0, sizeof *out);
    out->length = b.nrows; out->null_count = 0; out->offset = 0;
    out->n_buffers = 1; out->buffers = h->top_buffers;
    out->n_children = (int64_t)n; out->children = h->child_ptrs.data();
    out->release = release_top; out->private_data = h;
}

struct SchemaHolder {
    std::vector<ArrowSchema> children; std::vector<ArrowSchema*> ptrs; std::vector<std::string> names;
};
void release_schema_child(ArrowSchema* s) { s->release = nullptr; }
void release_schema_top(ArrowSchema* s) { delete (SchemaHolder*)s->private_data; s->release = nullptr; }
void export_schema(const std::vector<FieldMeta>& f, ArrowSchema* out) {
    SchemaHolder* h = new SchemaHolder();
    h->children.resize(f.size()); h->ptrs.resize(f.size()); h->names.resize(f.size());
    for (size_t j = 0; j < f.size(); ++j) {
        h->names[j] = f[j].name;
        ArrowSchema& c = h->children[j];
        memset(&c, 0, sizeof c);
        c.format = format_of_type(f[j].type); c.name = h->names[j].c_str();
        c.flags = f[j].nullable ? ARROW_FLAG_NULLABLE : 0; c.release = release_schema_child;
        h->ptrs[j] = &c;
    }
    memset(out, 0, sizeof *out);
    out->format = "+s"; out->name = ""; out->n_children = (int64_t)f.size(); out->children = h->ptrs.data();
    out->release = release_schema_top; out->private_data = h;
}

}  // namespace

// ---- the operator --------------------------------------------------------------
struct igl_op {
    igl_ctx* ctx = nullptr;
    bool is_join = false;
    std::vector<FieldMeta> out_fields;
    std::deque<OutBatch> ready;
    bool input_done = false;
    // filter / projection
    Staging in;
    // Filter / projection chunks are double-buffered: when a chunk is full its upload and kernels are
    // queued (fp_begin) and the caller goes on staging the next chunk into `in` while the GPU works on
    // `in_alt`; the chunk is finished (count read back, outputs copied down, batch queued: fp_end) when
    // the next one is full or the input ends.
    Staging in_alt;
    struct FpPending {
        bool active = false;
        int64_t n = 0;
        std::vector<igl_dev_col> dev, dout;
        std::vector<void*> to_free;
    } pend;
    igl_expr* pred = nullptr;
    std::vector<igl_expr*> proj_exprs;       // device expressions (owned)
    igl_fp* fp = nullptr;
    bool has_pred = false, need_rowid = false;
    struct OutSpec { bool utf8_pass; int index; };   // index: fp output # or input column #
    std::vector<OutSpec> out_spec;
    int n_fp_out = 0;
    // join
    Staging build;
    igl_join* join = nullptr;
    igl_expr *lkey = nullptr, *rkey = nullptr;   // the first pair (aliases of lkeys[0] / rkeys[0])
    std::vector<igl_expr*> lkeys, rkeys;         // owned
    bool build_done = false, left_join = false;
    int n_left = 0, n_right = 0;
    // Inner equi-join on integer keys of one type, neither side nullable: every output row's
    // right key equals its left key bit for bit, so the right key column is not copied back
    // over PCIe — the exported batch's two key columns share one host buffer.
    int alias_right_key = -1, alias_left_key = -1;
    // Probe chunks are pipelined over three streams: chunk i's upload (s_up), its probe (the context's
    // stream) and the download of its output (s_down) overlap the host-side staging of chunk i + 1 and
    // each other across chunks; igl_op_pull hands a batch out once its download has finished.
    cudaStream_t s_up = nullptr, s_down = nullptr;
    cudaEvent_t ev_up = nullptr;
    OpTrace trace;
    struct JoinPending {
        bool active = false;
        int64_t n = 0, cap = 0;
        std::vector<igl_dev_col> dev, dout;
        std::vector<void*> to_free;
    } jpend;
    std::vector<void*> build_dev;            // grown device columns of the whole build side
    std::vector<size_t> build_cap;
    std::vector<uint8_t*> build_valid;
    std::vector<uint8_t> build_any_null;     // a build batch carried nulls in this column (else the bitmap is dropped at finish_build)
    // Utf8 columns ride along by row id: the device join sees an Int32 row-id column in a string
    // column's place (the probe chunk's row number / the build side's row number) and the strings are
    // gathered by the ids it emits (utf8.cu), as filter_record_batch's compaction does for Filter.
    std::vector<uint8_t> l_utf8, r_utf8;     // per column: 1 = Utf8
    bool any_utf8 = false;
    void* d_iota = nullptr;                  // 0 .. max(chunk_rows, build rows) - 1
    int64_t iota_rows = 0;
    std::vector<void*> build_str;            // per Utf8 build column: all its bytes on the device
    std::vector<size_t> build_str_cap, build_str_used;
    std::vector<void*> build_off;            // per Utf8 build column: build_rows + 1 offsets
    int64_t build_rows = 0, build_rows_cap = 0;
    // Utf8 join keys (hash_join.rs:117-122 joins on the printed string): the device join is given a
    // 64-bit hash of the string as an Int64 key column on each side (NULL where the string is NULL, so
    // NULL == NULL as in the reference), and every joined pair is then compared byte for byte through
    // the row ids both sides carry.  Differing pairs (a hash collision) are filtered out of an Inner /
    // Right join's output on the device; a Left / Full join reports the collision instead.
    struct Utf8Key { int lcol, rcol; };
    std::vector<Utf8Key> utf8_keys;
    std::vector<void*> build_hash;           // per string key: the build side's hash column
    igl_fp* collide_fp = nullptr;            // output columns + keep flag -> output columns where keep != 0
    int utf8_hash_bits = 64;                 // IGL_UTF8_KEY_HASH_BITS: fewer bits force collisions (tests)
    unsigned long long* h_mismatch = nullptr;   // pinned
    unsigned long long* d_mismatch = nullptr;
};

namespace {

int32_t d2h_col(igl_ctx* ctx, cudaStream_t st, OutCol& oc, const void* d_data, const uint8_t* d_valid, int64_t n, bool nullable) {
    const size_t bytes = oc.type == IGL_T_BOOL ? (size_t)(n + 7) / 8 : (size_t)n * type_width(oc.type);
    oc.data = pinned_pool().get(pad64(bytes));
    if (!oc.data) return fail(IGL_ERR_OOM, "pinned output allocation of %zu bytes failed", bytes);
    if (bytes) IGL_CUDA_OK(cudaMemcpyAsync(oc.data, d_data, bytes, cudaMemcpyDeviceToHost, st));
    if (nullable && d_valid) {
        oc.valid = (uint8_t*)pinned_pool().get(pad64((size_t)(n + 7) / 8));
        if (!oc.valid) return fail(IGL_ERR_OOM, "pinned output allocation failed");
        IGL_CUDA_OK(cudaMemcpyAsync(oc.valid, d_valid, (size_t)(n + 7) / 8, cudaMemcpyDeviceToHost, st));
    }
    (void)ctx;
    return IGL_OK;
}

// Runs one staged chunk of a filter / projection operator.
// Queues the staged chunk: upload, output buffers, kernel launches.  Does not wait.
int32_t fp_begin(igl_op* op) {
    Staging& S = op->in;
    const int64_t n = S.rows;
    if (n == 0) return IGL_OK;
    igl_ctx* ctx = op->ctx;
    cudaStream_t st = ctx->stream;
    auto& PD = op->pend;
    PD.dev.clear(); PD.dout.clear(); PD.to_free.clear();
    IGL_TRY(S.upload(st, PD.dev));
    const int nfo = op->n_fp_out + (op->need_rowid ? 1 : 0);
    PD.dout.assign(nfo, igl_dev_col{});
    auto cleanup = [&]() { for (void* p : PD.to_free) pool_free(ctx, p); PD.to_free.clear(); };
    if (op->fp) {
        for (int j = 0; j < nfo; ++j) {
            int32_t t = 0, nl = 0;
            igl_fp_output_type(op->fp, j, &t, &nl);
            PD.dout[j].type = t;
            const size_t bytes = t == IGL_T_BOOL ? (size_t)(n + 7) / 8 : (size_t)n * type_width(t);
            void* p = nullptr;
            int32_t rc = pool_alloc(ctx, pad64(bytes), &p);
            if (rc < 0) { cleanup(); return rc; }
            PD.to_free.push_back(p); PD.dout[j].data = p;
            if (nl) {
                rc = pool_alloc(ctx, pad64((size_t)(n + 7) / 8), &p);
                if (rc < 0) { cleanup(); return rc; }
                PD.to_free.push_back(p); PD.dout[j].validity = (uint8_t*)p;
            }
        }
        int32_t rc = igl_fp_launch(op->fp, PD.dev.data(), n, PD.dout.data(), st);
        if (rc < 0) { cleanup(); return rc; }
    }
    PD.n = n;
    PD.active = true;
    std::swap(op->in, op->in_alt);   // the queued chunk now lives in in_alt; `in` is the empty one
    return IGL_OK;
}

// Finishes the queued chunk: waits for it, copies the outputs down, queues the batch.
int32_t fp_end(igl_op* op) {
    auto& PD = op->pend;
    if (!PD.active) return IGL_OK;
    PD.active = false;
    Staging& S = op->in_alt;
    const int64_t n = PD.n;
    igl_ctx* ctx = op->ctx;
    cudaStream_t st = ctx->stream;
    std::vector<igl_dev_col>& dev = PD.dev;
    std::vector<igl_dev_col>& dout = PD.dout;
    std::vector<void*>& to_free = PD.to_free;
    auto cleanup = [&]() { for (void* p : to_free) pool_free(ctx, p); to_free.clear(); };
    int64_t n_out = n;
    if (op->fp) {
        int32_t rc = igl_fp_finish(op->fp, &n_out, st);
        if (rc < 0) { cudaStreamSynchronize(st); S.reset(); cleanup(); return rc; }
    }
    S.reset();
    if (n_out == 0 && op->has_pred) { cleanup(); return IGL_OK; }   // filter.rs:59-63: empty results are dropped
    OutBatch ob;
    ob.nrows = n_out;
    ob.cols.resize(op->out_spec.size());
    const int32_t* d_rowid = op->need_rowid ? (const int32_t*)dout[op->n_fp_out].data : nullptr;
    std::vector<std::pair<size_t, int32_t*>> utf8_pending;   // (output index, device offsets)
    std::vector<uint8_t*> utf8_valid_bytes(op->out_spec.size(), nullptr);
    int32_t rc = IGL_OK;
    for (size_t j = 0; j < op->out_spec.size() && rc >= 0; ++j) {
        OutCol& oc = ob.cols[j];
        const auto& sp = op->out_spec[j];
        oc.type = op->out_fields[j].type;
        if (!sp.utf8_pass) {
            rc = d2h_col(ctx, st, oc, dout[sp.index].data, dout[sp.index].validity, n_out, op->out_fields[j].nullable);
        } else {
            // K2s: offsets (lengths + scan) now, bytes once the total is known
            const igl_dev_col& src = dev[sp.index];
            void* d_off = nullptr; void* scratch = nullptr; void* vbytes = nullptr;
            const size_t sbytes = 256 + ((size_t)n_out / 2048 + 8) * 8;
            rc = pool_alloc(ctx, pad64((size_t)(n_out + 1) * 4), &d_off);
            if (rc >= 0) { to_free.push_back(d_off); rc = pool_alloc(ctx, sbytes, &scratch); }
            if (rc >= 0) { to_free.push_back(scratch); if (src.validity) { rc = pool_alloc(ctx, pad64((size_t)n_out + 1), &vbytes); if (rc >= 0) to_free.push_back(vbytes); } }
            if (rc >= 0) rc = utf8_gather_offsets(ctx, src.offsets, src.validity, d_rowid, n_out, (int32_t*)d_off, (uint8_t*)vbytes, scratch, sbytes, st);
            if (rc >= 0) {
                oc.off = (int32_t*)pinned_pool().get(pad64((size_t)(n_out + 1) * 4));
                if (!oc.off) rc = fail(IGL_ERR_OOM, "pinned output allocation failed");
            }
            if (rc >= 0 && cudaMemcpyAsync(oc.off, d_off, (size_t)(n_out + 1) * 4, cudaMemcpyDeviceToHost, st) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "offset copy failed");
            utf8_pending.push_back({j, (int32_t*)d_off});
            utf8_valid_bytes[j] = (uint8_t*)vbytes;
        }
    }
    if (rc >= 0 && cudaStreamSynchronize(st) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "stream synchronise failed: %s", cudaGetErrorString(cudaGetLastError()));
    for (auto& pr : utf8_pending) {
        if (rc < 0) break;
        const size_t j = pr.first;
        OutCol& oc = ob.cols[j];
        const igl_dev_col& src = dev[op->out_spec[j].index];
        const size_t total = (size_t)oc.off[n_out];
        void* d_bytes = nullptr;
        rc = pool_alloc(ctx, pad64(total), &d_bytes);
        if (rc < 0) break;
        to_free.push_back(d_bytes);
        rc = utf8_gather_bytes(ctx, src.offsets, (const uint8_t*)src.data, d_rowid, n_out, pr.second, (uint8_t*)d_bytes, st);
        if (rc < 0) break;
        oc.data = pinned_pool().get(pad64(total));
        if (!oc.data) { rc = fail(IGL_ERR_OOM, "pinned output allocation failed"); break; }
        if (total && cudaMemcpyAsync(oc.data, d_bytes, total, cudaMemcpyDeviceToHost, st) != cudaSuccess) { rc = fail(IGL_ERR_CUDA, "string copy failed"); break; }
        if (utf8_valid_bytes[j]) {   // validity gathered as bytes -> bitmap on the device, then down
            void* d_bits = nullptr;
            rc = pool_alloc(ctx, pad64((size_t)(n_out + 7) / 8 + 4), &d_bits);
            if (rc < 0) break;
            to_free.push_back(d_bits);
            if (launch_pack_bits(utf8_valid_bytes[j], (uint32_t*)d_bits, nullptr, n_out, n_out, ctx->sm_count, st) != cudaSuccess) { rc = fail(IGL_ERR_CUDA, "validity pack failed"); break; }
            ctx->launches++;
            oc.valid = (uint8_t*)pinned_pool().get(pad64((size_t)(n_out + 7) / 8 + 4));
            if (!oc.valid) { rc = fail(IGL_ERR_OOM, "pinned output allocation failed"); break; }
            if (cudaMemcpyAsync(oc.valid, d_bits, (size_t)(n_out + 7) / 8, cudaMemcpyDeviceToHost, st) != cudaSuccess) { rc = fail(IGL_ERR_CUDA, "validity copy failed"); break; }
        }
    }
    if (rc >= 0 && !utf8_pending.empty() && cudaStreamSynchronize(st) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "stream synchronise failed");
    cleanup();
    if (rc < 0) {
        for (auto& c : ob.cols) { pinned_pool().put(c.data); pinned_pool().put(c.valid); pinned_pool().put(c.off); }
        return rc;
    }
    for (auto& c : ob.cols) c.null_count = c.valid ? count_zero_bits(c.valid, n_out) : 0;
    op->ready.push_back(std::move(ob));
    return IGL_OK;
}

// A chunk is full (or the input ended): finish the chunk queued before, queue this one.
int32_t fp_flush(igl_op* op) {
    IGL_TRY(fp_end(op));
    return fp_begin(op);
}

// 0, 1, 2, ... on the device: the row-id column that stands in for string columns inside the join.
int32_t ensure_iota(igl_op* op, int64_t rows) {
    if (rows <= op->iota_rows && op->d_iota) return IGL_OK;
    if (rows >= ((int64_t)1 << 31)) return fail(IGL_ERR_UNSUPPORTED, "a join side with Utf8 columns is limited to 2^31 rows");
    igl_ctx* ctx = op->ctx;
    IGL_CUDA_OK(cudaStreamSynchronize(ctx->stream));   // nothing still reads the old one
    if (op->d_iota) pool_free(ctx, op->d_iota);
    op->d_iota = nullptr; op->iota_rows = 0;
    const int64_t cap = rows + rows / 4 + 1024;
    IGL_TRY(pool_alloc(ctx, pad64((size_t)cap * 4), &op->d_iota));
    IGL_TRY(launch_iota_i32(ctx, (int32_t*)op->d_iota, cap, ctx->stream));
    op->iota_rows = cap;
    return IGL_OK;
}

// Room for `n` more rows on the device-resident build side (every column grows by doubling; one
// wait per growth step, after all the copies have been queued).
int32_t join_build_reserve(igl_op* op, int64_t n) {
    Staging& S = op->build;
    igl_ctx* ctx = op->ctx;
    cudaStream_t st = ctx->stream;
    if (op->build_rows + n <= op->build_rows_cap) return IGL_OK;
    const int64_t cap = std::max<int64_t>(op->build_rows + n, op->build_rows_cap * 2);
    std::vector<void*> old_blocks;
    for (int c = 0; c < op->n_right; ++c) {
        const bool bits = S.fields[c].type == IGL_T_BOOL;   // bit-packed: grows bytewise, like a validity bitmap
        const int w = type_width(S.fields[c].type);
        void* p = nullptr;
        IGL_TRY(pool_alloc(ctx, pad64(bits ? (size_t)(cap + 7) / 8 + 8 : (size_t)cap * w), &p));
        if (bits) IGL_CUDA_OK(cudaMemsetAsync(p, 0, pad64((size_t)(cap + 7) / 8 + 8), st));
        if (op->build_dev[c]) {
            IGL_CUDA_OK(cudaMemcpyAsync(p, op->build_dev[c], bits ? (size_t)(op->build_rows + 7) / 8 : (size_t)op->build_rows * w, cudaMemcpyDeviceToDevice, st));
            old_blocks.push_back(op->build_dev[c]);
        }
        op->build_dev[c] = p;
        if (S.fields[c].nullable) {
            void* v = nullptr;
            IGL_TRY(pool_alloc(ctx, pad64((size_t)(cap + 7) / 8 + 8), &v));
            IGL_CUDA_OK(cudaMemsetAsync(v, 0, pad64((size_t)(cap + 7) / 8 + 8), st));
            if (op->build_valid[c]) {
                IGL_CUDA_OK(cudaMemcpyAsync(v, op->build_valid[c], (size_t)(op->build_rows + 7) / 8, cudaMemcpyDeviceToDevice, st));
                old_blocks.push_back(op->build_valid[c]);
            }
            op->build_valid[c] = (uint8_t*)v;
        }
    }
    for (int c = 0; c < op->n_right; ++c) {
        if (!op->r_utf8[c]) continue;
        void* p = nullptr;
        IGL_TRY(pool_alloc(ctx, pad64((size_t)(cap + 1) * 4), &p));
        if (op->build_off[c]) {
            IGL_CUDA_OK(cudaMemcpyAsync(p, op->build_off[c], (size_t)(op->build_rows + 1) * 4, cudaMemcpyDeviceToDevice, st));
            old_blocks.push_back(op->build_off[c]);
        }
        op->build_off[c] = p;
    }
    if (!old_blocks.empty()) IGL_CUDA_OK(cudaStreamSynchronize(st));
    for (void* q : old_blocks) pool_free(ctx, q);
    op->build_rows_cap = cap;
    return IGL_OK;
}

// Appends one staged chunk of a Utf8 build column: its bytes behind the column's bytes so far, its
// offsets shifted by that amount behind the offsets so far.
int32_t join_build_append_utf8(igl_op* op, int c, StagedCol& sc, int64_t n, cudaStream_t st) {
    igl_ctx* ctx = op->ctx;
    const size_t used = op->build_str_used[c], add = sc.utf8_used;
    if (used + add >= ((size_t)1 << 31)) return fail(IGL_ERR_UNSUPPORTED, "a Utf8 build column is limited to 2 GiB of string bytes");
    if (used + add > op->build_str_cap[c]) {
        const size_t cap = std::max(used + add, op->build_str_cap[c] * 2);
        void* p = nullptr;
        IGL_TRY(pool_alloc(ctx, pad64(cap), &p));
        if (op->build_str[c]) {
            IGL_CUDA_OK(cudaMemcpyAsync(p, op->build_str[c], used, cudaMemcpyDeviceToDevice, st));
            IGL_CUDA_OK(cudaStreamSynchronize(st));
            pool_free(ctx, op->build_str[c]);
        }
        op->build_str[c] = p; op->build_str_cap[c] = cap;
    }
    if (add) IGL_CUDA_OK(cudaMemcpyAsync((uint8_t*)op->build_str[c] + used, sc.h_data, add, cudaMemcpyHostToDevice, st));
    for (int64_t i = 0; i <= n; ++i) sc.h_off[i] += (int32_t)used;   // chunk-relative -> column-relative (undone below)
    IGL_CUDA_OK(cudaMemcpyAsync((int32_t*)op->build_off[c] + op->build_rows, sc.h_off, (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    for (int64_t i = 0; i <= n; ++i) sc.h_off[i] -= (int32_t)used;
    op->build_str_used[c] = used + add;
    return IGL_OK;
}

// Appends the staged build chunk to the device-resident build side.
int32_t join_build_flush(igl_op* op) {
    Staging& S = op->build;
    const int64_t n = S.rows;
    if (n == 0) return IGL_OK;
    igl_ctx* ctx = op->ctx;
    cudaStream_t st = ctx->stream;
    IGL_TRY(join_build_reserve(op, n));
    if ((op->build_rows & 7) != 0) {
        // validity bitmaps and Boolean values are appended bytewise: keep chunk boundaries byte aligned
        for (int c = 0; c < op->n_right; ++c)
            if (S.fields[c].nullable || S.fields[c].type == IGL_T_BOOL) return fail(IGL_ERR_UNSUPPORTED, "nullable or Boolean build columns need coalesce_rows to be a multiple of 8");
    }
    for (int c = 0; c < op->n_right; ++c) {
        const int w = type_width(S.fields[c].type);
        if (S.fields[c].type == IGL_T_UTF8)
            IGL_TRY(join_build_append_utf8(op, c, S.cols[c], n, st));
        else if (S.fields[c].type == IGL_T_BOOL)
            IGL_CUDA_OK(cudaMemcpyAsync((uint8_t*)op->build_dev[c] + (op->build_rows >> 3), S.cols[c].h_data, (size_t)(n + 7) / 8, cudaMemcpyHostToDevice, st));
        else
        IGL_CUDA_OK(cudaMemcpyAsync((uint8_t*)op->build_dev[c] + (size_t)op->build_rows * w, S.cols[c].h_data, (size_t)n * w, cudaMemcpyHostToDevice, st));
        if (S.fields[c].nullable)
            IGL_CUDA_OK(cudaMemcpyAsync(op->build_valid[c] + (op->build_rows >> 3), S.cols[c].h_valid, (size_t)(n + 7) / 8, cudaMemcpyHostToDevice, st));
    }
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    op->build_rows += n;
    for (int c = 0; c < op->n_right; ++c) if (S.cols[c].any_null) op->build_any_null[c] = 1;
    S.reset();
    return IGL_OK;
}

// One Utf8 output column of a finished probe chunk: the join emitted row ids in its place (probe chunk
// rows for a left column, build rows for a right one; the null-extended rows of a Left join carry an
// unset validity bit); the strings are gathered by them — lengths, scan, bytes (utf8.cu) — and copied
// down on the download stream.
int32_t join_gather_utf8(igl_op* op, int c, int64_t n_out, OutCol& oc, std::vector<void*>& dev_free) {
    igl_ctx* ctx = op->ctx;
    cudaStream_t st = ctx->stream;
    auto& PD = op->jpend;
    const bool is_left = c < op->n_left;
    const int sc = is_left ? c : c - op->n_left;
    int32_t* rowid = (int32_t*)PD.dout[c].data;
    const int32_t* src_off = is_left ? PD.dev[c].offsets : (const int32_t*)op->build_off[sc];
    const uint8_t* src_bytes = is_left ? (const uint8_t*)PD.dev[c].data : (const uint8_t*)op->build_str[sc];
    const uint8_t* src_valid = is_left ? PD.dev[c].validity : op->build_valid[sc];
    const bool null_rows = !is_left && op->left_join;
    if (null_rows) IGL_TRY(launch_mask_rowids(ctx, rowid, PD.dout[c].validity, n_out, st));
    const bool want_valid = src_valid != nullptr || null_rows;
    void *d_off = nullptr, *scratch = nullptr, *vbytes = nullptr, *d_bytes = nullptr, *d_bits = nullptr;
    const size_t sbytes = 256 + ((size_t)n_out / 2048 + 8) * 8 + 64;
    IGL_TRY(pool_alloc(ctx, pad64((size_t)(n_out + 1) * 4), &d_off)); dev_free.push_back(d_off);
    IGL_TRY(pool_alloc(ctx, sbytes, &scratch)); dev_free.push_back(scratch);
    if (want_valid) { IGL_TRY(pool_alloc(ctx, pad64((size_t)n_out + 1), &vbytes)); dev_free.push_back(vbytes); }
    IGL_TRY(utf8_gather_offsets(ctx, src_off, src_valid, rowid, n_out, (int32_t*)d_off, (uint8_t*)vbytes, scratch, sbytes, st));
    oc.off = (int32_t*)pinned_pool().get(pad64((size_t)(n_out + 1) * 4));
    if (!oc.off) return fail(IGL_ERR_OOM, "pinned output allocation failed");
    IGL_CUDA_OK(cudaMemcpyAsync(oc.off, d_off, (size_t)(n_out + 1) * 4, cudaMemcpyDeviceToHost, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));   // the byte total decides the size of what follows
    const size_t total = (size_t)oc.off[n_out];
    IGL_TRY(pool_alloc(ctx, pad64(total), &d_bytes)); dev_free.push_back(d_bytes);
    IGL_TRY(utf8_gather_bytes(ctx, src_off, src_bytes, rowid, n_out, (const int32_t*)d_off, (uint8_t*)d_bytes, st));
    if (want_valid) {
        IGL_TRY(pool_alloc(ctx, pad64((size_t)(n_out + 7) / 8 + 4), &d_bits)); dev_free.push_back(d_bits);
        if (launch_pack_bits((const uint8_t*)vbytes, (uint32_t*)d_bits, nullptr, n_out, n_out, ctx->sm_count, st) != cudaSuccess) return fail(IGL_ERR_CUDA, "validity pack failed");
        ctx->launches++;
    }
    IGL_CUDA_OK(cudaStreamSynchronize(st));   // the download stream may start
    oc.data = pinned_pool().get(pad64(total));
    if (!oc.data) return fail(IGL_ERR_OOM, "pinned output allocation failed");
    if (total) IGL_CUDA_OK(cudaMemcpyAsync(oc.data, d_bytes, total, cudaMemcpyDeviceToHost, op->s_down));
    if (want_valid) {
        oc.valid = (uint8_t*)pinned_pool().get(pad64((size_t)(n_out + 7) / 8 + 4));
        if (!oc.valid) return fail(IGL_ERR_OOM, "pinned output allocation failed");
        IGL_CUDA_OK(cudaMemcpyAsync(oc.valid, d_bits, (size_t)(n_out + 7) / 8, cudaMemcpyDeviceToHost, op->s_down));
    }
    return IGL_OK;
}

// String keys: every joined pair's strings are compared through the row ids the join emitted; pairs
// that differ (their 64-bit hashes collided) are filtered out of the chunk's output columns.
int32_t join_verify_utf8_keys(igl_op* op, int64_t* n_out) {
    igl_ctx* ctx = op->ctx;
    cudaStream_t st = ctx->stream;
    auto& PD = op->jpend;
    const int64_t n = *n_out;
    const int nc = op->n_left + op->n_right;
    void* keep = nullptr;
    IGL_TRY(pool_alloc(ctx, pad64((size_t)n * 4), &keep));
    PD.to_free.push_back(keep);
    IGL_TRY(launch_fill_i32(ctx, (int32_t*)keep, n, 1, st));
    IGL_CUDA_OK(cudaMemsetAsync(op->d_mismatch, 0, sizeof(unsigned long long), st));
    for (auto& k : op->utf8_keys) {
        const int oc = op->n_left + k.rcol;
        IGL_TRY(utf8_verify_keys(ctx, (const int32_t*)PD.dout[k.lcol].data, (const int32_t*)PD.dout[oc].data,
                                 op->left_join ? PD.dout[oc].validity : nullptr, n,
                                 PD.dev[k.lcol].offsets, (const uint8_t*)PD.dev[k.lcol].data, PD.dev[k.lcol].validity,
                                 (const int32_t*)op->build_off[k.rcol], (const uint8_t*)op->build_str[k.rcol], op->build_valid[k.rcol],
                                 (int32_t*)keep, op->d_mismatch, st));
    }
    IGL_CUDA_OK(cudaMemcpyAsync(op->h_mismatch, op->d_mismatch, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    if (*op->h_mismatch == 0) return IGL_OK;
    if (op->left_join)
        return fail(IGL_ERR_UNSUPPORTED, "Left/Full join on a Utf8 key: %llu joined pairs have equal 64-bit hashes and different strings; "
                    "whether their left rows have a true match is not tracked on the device", *op->h_mismatch);
    // the same columns once more, without the differing pairs
    std::vector<igl_dev_col> mid(nc + 1), fin(nc);
    for (int c = 0; c < nc; ++c) {
        mid[c] = PD.dout[c];
        const bool is_left = c < op->n_left;
        const bool wrote_valid = is_left ? PD.dev[c].validity != nullptr : op->build_valid[c - op->n_left] != nullptr;
        if (!wrote_valid || (is_left ? op->l_utf8[c] : op->r_utf8[c - op->n_left])) mid[c].validity = nullptr;
        fin[c] = igl_dev_col{mid[c].type, 0, nullptr, nullptr, nullptr};
        void* p = nullptr;
        IGL_TRY(pool_alloc(ctx, pad64(mid[c].type == IGL_T_BOOL ? (size_t)(n + 7) / 8 + 8 : (size_t)n * type_width(mid[c].type)), &p));
        PD.to_free.push_back(p); fin[c].data = p;
        IGL_TRY(pool_alloc(ctx, pad64((size_t)(n + 7) / 8 + 8), &p));
        PD.to_free.push_back(p); fin[c].validity = (uint8_t*)p;
    }
    mid[nc] = igl_dev_col{IGL_T_I32, 0, keep, nullptr, nullptr};
    int64_t kept = 0;
    IGL_TRY(igl_fp_run(op->collide_fp, mid.data(), n, fin.data(), &kept, (void*)st));
    for (int c = 0; c < nc; ++c) PD.dout[c] = fin[c];
    *n_out = kept;
    return IGL_OK;
}

// Finishes the probe chunk queued before: waits for its kernels, queues the download of its output
// on the download stream and appends the batch to `ready` (igl_op_pull waits for the download).
int32_t join_probe_end_chunk(igl_op* op) {
    auto& PD = op->jpend;
    if (!PD.active) return IGL_OK;
    PD.active = false;
    igl_ctx* ctx = op->ctx;
    cudaStream_t st = ctx->stream;
    auto cleanup = [&]() { for (void* p : PD.to_free) pool_free(ctx, p); PD.to_free.clear(); };
    int64_t n_out = 0;
    int32_t rc = join_probe_end(op->join, &n_out, st);   // the compute stream is drained here
    if (rc < 0) { cudaStreamSynchronize(st); cleanup(); return rc; }
    if (n_out == 0) { cleanup(); return IGL_OK; }
    const int nc = op->n_left + op->n_right;
    if (!op->utf8_keys.empty()) {
        rc = join_verify_utf8_keys(op, &n_out);
        if (rc < 0) { cudaStreamSynchronize(st); cleanup(); return rc; }
        if (n_out == 0) { cleanup(); return IGL_OK; }
    }
    OutBatch ob;
    ob.nrows = n_out;
    ob.cols.resize(nc);
    for (int c = 0; c < nc && rc >= 0; ++c) {
        ob.cols[c].type = op->out_fields[c].type;
        if (op->out_fields[c].type == IGL_T_UTF8) { rc = join_gather_utf8(op, c, n_out, ob.cols[c], ob.dev_free); continue; }
        if (op->alias_right_key >= 0 && c == op->n_left + op->alias_right_key && PD.dev[op->alias_left_key].validity == nullptr &&
            op->build_valid[op->alias_right_key] == nullptr) {   // no NULL keys on either side in this chunk
            ob.cols[c].data = ob.cols[op->alias_left_key].data;   // left columns come first: already set
            ob.cols[c].alias = true;
            continue;
        }
        // igl_join_probe writes a validity bitmap for a left column that has one and
        // for a right column that can be null (source nullable, or null-extended)
        const bool is_left = c < op->n_left;
        const bool wrote_valid = is_left ? PD.dev[c].validity != nullptr
                                         : (op->left_join || op->build_valid[c - op->n_left] != nullptr);
        rc = d2h_col(ctx, op->s_down, ob.cols[c], PD.dout[c].data, wrote_valid ? PD.dout[c].validity : nullptr, n_out, true);
    }
    if (rc >= 0 && (cudaEventCreateWithFlags(&ob.done, cudaEventDisableTiming) != cudaSuccess || cudaEventRecord(ob.done, op->s_down) != cudaSuccess))
        rc = fail(IGL_ERR_CUDA, "event for the output download failed");
    if (rc < 0) {
        cudaStreamSynchronize(op->s_down);
        cudaStreamSynchronize(st);
        cleanup();
        for (void* p : ob.dev_free) pool_free(ctx, p);
        if (ob.done) cudaEventDestroy(ob.done);
        for (auto& c : ob.cols) { if (!c.alias) pinned_pool().put(c.data); pinned_pool().put(c.valid); pinned_pool().put(c.off); }
        return rc;
    }
    ob.dev_free.insert(ob.dev_free.end(), PD.to_free.begin(), PD.to_free.end());   // the device output blocks live until the download has passed
    PD.to_free.clear();
    op->ready.push_back(std::move(ob));
    return IGL_OK;
}

// Queues the probe of one chunk whose columns are (being) uploaded into `dev`; does not wait for it.
int32_t join_probe_begin_chunk(igl_op* op, int64_t n, std::vector<igl_dev_col>&& dev) {
    igl_ctx* ctx = op->ctx;
    cudaStream_t st = ctx->stream;
    auto& PD = op->jpend;
    PD.dev = std::move(dev);
    PD.dout.clear(); PD.to_free.clear();
    {
        HostTrace tr("op.begin.event");
        IGL_CUDA_OK(cudaEventRecord(op->ev_up, op->s_up));
        IGL_CUDA_OK(cudaStreamWaitEvent(st, op->ev_up, 0));
    }
    HostTrace tr_alloc("op.begin.alloc+probe");
    int64_t cap = n;
    auto cleanup = [&]() { for (void* p : PD.to_free) pool_free(ctx, p); PD.to_free.clear(); };
    const int nu = (int)op->utf8_keys.size();
    std::vector<igl_dev_col> jdev = PD.dev;   // what the device join reads: row ids where the strings are,
    for (int c = 0; c < op->n_left; ++c)      // and behind them one hash column per string key
        if (op->l_utf8[c]) jdev[c] = igl_dev_col{IGL_T_I32, 0, op->d_iota, nullptr, nullptr};
    for (int i = 0; i < nu; ++i) {
        const int lc = op->utf8_keys[i].lcol;
        void* p = nullptr;
        int32_t rc = pool_alloc(ctx, pad64((size_t)std::max<int64_t>(n, 1) * 8), &p);
        if (rc < 0) { cleanup(); return rc; }
        PD.to_free.push_back(p);
        rc = utf8_hash_keys(ctx, PD.dev[lc].offsets, (const uint8_t*)PD.dev[lc].data, n, op->utf8_hash_bits, (int64_t*)p, st);
        if (rc < 0) { cleanup(); return rc; }
        jdev.push_back(igl_dev_col{IGL_T_I64, 0, p, PD.dev[lc].validity, nullptr});
    }
    if (igl_join_has_duplicates(op->join)) {
        int32_t rc = igl_join_probe_count(op->join, jdev.data(), n, &cap, st);
        if (rc < 0) { cleanup(); return rc; }
    }
    const int nc = op->n_left + op->n_right;
    PD.dout.assign(nc, igl_dev_col{});
    for (int c = 0; c < nc; ++c) {
        PD.dout[c].type = op->out_fields[c].type == IGL_T_UTF8 ? IGL_T_I32 : op->out_fields[c].type;
        void* p = nullptr;
        const size_t cap1 = (size_t)std::max<int64_t>(cap, 1);
        int32_t rc = pool_alloc(ctx, pad64(PD.dout[c].type == IGL_T_BOOL ? (cap1 + 7) / 8 + 8 : cap1 * type_width(PD.dout[c].type)), &p);
        if (rc < 0) { cleanup(); return rc; }
        PD.to_free.push_back(p); PD.dout[c].data = p;
        rc = pool_alloc(ctx, pad64((size_t)(cap + 7) / 8 + 8), &p);
        if (rc < 0) { cleanup(); return rc; }
        PD.to_free.push_back(p); PD.dout[c].validity = (uint8_t*)p;
    }
    // the device join's own column order: left columns, left hash columns, right columns, right hash columns
    std::vector<igl_dev_col> jout(PD.dout.begin(), PD.dout.begin() + op->n_left);
    for (int side = 0; side < 2; ++side) {
        for (int i = 0; i < nu; ++i) {
            void* p = nullptr;
            int32_t rc2 = pool_alloc(ctx, pad64((size_t)std::max<int64_t>(cap, 1) * 8), &p);
            if (rc2 < 0) { cleanup(); return rc2; }
            PD.to_free.push_back(p);
            void* v = nullptr;
            rc2 = pool_alloc(ctx, pad64((size_t)(cap + 7) / 8 + 8), &v);
            if (rc2 < 0) { cleanup(); return rc2; }
            PD.to_free.push_back(v);
            jout.push_back(igl_dev_col{IGL_T_I64, 0, p, (uint8_t*)v, nullptr});
        }
        if (side == 0) jout.insert(jout.end(), PD.dout.begin() + op->n_left, PD.dout.end());
    }
    int32_t rc;
    {
        HostTrace tr("op.begin.probe");
        rc = join_probe_begin(op->join, jdev.data(), n, jout.data(), cap, st);
    }
    if (rc < 0) { cudaStreamSynchronize(st); cleanup(); return rc; }
    PD.n = n; PD.cap = cap;
    PD.active = true;
    return IGL_OK;
}

// A staged chunk is full (or the input ended): finish the chunk queued before, queue this one.
int32_t join_probe_flush(igl_op* op) {
    { TraceScope ts(op->trace, TR_END); IGL_TRY(join_probe_end_chunk(op)); }
    Staging& S = op->in;
    const int64_t n = S.rows;
    if (n == 0) return IGL_OK;
    if (op->any_utf8) IGL_TRY(ensure_iota(op, std::max<int64_t>(n, S.chunk_rows)));
    std::vector<igl_dev_col> dev;
    { TraceScope ts(op->trace, TR_UPLOAD); IGL_TRY(S.upload(op->s_up, dev)); }
    { TraceScope ts(op->trace, TR_BEGIN); IGL_TRY(join_probe_begin_chunk(op, n, std::move(dev))); }
    std::swap(op->in, op->in_alt);   // the queued chunk's pinned and device buffers now sit in in_alt
    op->in.reset();                  // the chunk that used these buffers was finished above
    return IGL_OK;
}

// Page-locked host buffers need no staging copy: the DMA engine reads them where they are.
bool is_pinned_host(const void* p) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// True when rows of `batch` can be uploaded straight from the caller's buffers: fixed-width
// columns without nulls in page-locked memory, and enough rows to be worth a chunk of their own.
bool batch_is_direct(const Staging& S, const ArrowArray* batch) {
    static const bool enabled = !(getenv("IGL_OP_DIRECT") && getenv("IGL_OP_DIRECT")[0] == '0');
    if (!enabled || S.rows != 0 || batch->length < (64 << 10) || (size_t)batch->n_children != S.fields.size()) return false;
    for (size_t j = 0; j < S.fields.size(); ++j) {
        if (!S.wanted[j]) continue;
        const int t = S.fields[j].type;
        const ArrowArray* a = batch->children[j];
        if (type_width(t) == 0 || a->null_count != 0 || a->n_buffers < 2 || !is_pinned_host(a->buffers[1])) return false;
    }
    return true;
}

// Uploads rows [from, from + n) of a direct batch into the device buffers of `S` on stream `up`.
int32_t direct_upload(Staging& S, const ArrowArray* batch, int64_t from, int64_t n, cudaStream_t up, std::vector<igl_dev_col>& dev) {
    igl_ctx* ctx = S.ctx;
    dev.assign(S.fields.size(), igl_dev_col{});
    for (size_t j = 0; j < S.fields.size(); ++j) {
        dev[j].type = S.fields[j].type;
        if (!S.wanted[j]) continue;
        StagedCol& c = S.cols[j];
        const ArrowArray* a = batch->children[j];
        const int w = type_width(c.type);
        const size_t bytes = (size_t)n * w;
        if (pad64(bytes) > c.d_data_cap) {
            if (c.d_data) pool_free(ctx, c.d_data);
            c.d_data = nullptr; c.d_data_cap = 0;
            IGL_TRY(pool_alloc(ctx, pad64(std::max(bytes, (size_t)S.chunk_rows * w)), &c.d_data));
            c.d_data_cap = pad64(std::max(bytes, (size_t)S.chunk_rows * w));
        }
        IGL_CUDA_OK(cudaMemcpyAsync(c.d_data, (const uint8_t*)a->buffers[1] + (size_t)(a->offset + batch->offset + from) * w, bytes,
                                    cudaMemcpyHostToDevice, up));
        dev[j].data = c.d_data;
        if (c.nullable && !S.elide_all_valid) {   // declared nullable, no nulls in this batch: an all-ones bitmap
            if (!c.d_valid) { void* p = nullptr; IGL_TRY(pool_alloc(ctx, pad64((size_t)(S.chunk_rows + 7) / 8), &p)); c.d_valid = (uint8_t*)p; }
            IGL_CUDA_OK(cudaMemsetAsync(c.d_valid, 0xFF, (size_t)(n + 7) / 8, up));
            dev[j].validity = c.d_valid;
        }
    }
    return IGL_OK;
}

// Probe rows straight from the caller's page-locked batch, chunk by chunk.  The batch is only
// borrowed for this call, so each chunk's upload is awaited before the next step — while it
// travels, the previous chunk's output is on its way down and its successor's probe runs.
int32_t join_probe_direct(igl_op* op, const ArrowArray* batch) {
    int64_t done = 0;
    while (done < batch->length) {
        const int64_t n = std::min<int64_t>(batch->length - done, op->in.chunk_rows);
        // the upload goes first: it travels while the previous chunk's probe is awaited (that chunk
        // reads the other pair of device buffers) and its output download is queued
        std::vector<igl_dev_col> dev;
        { TraceScope ts(op->trace, TR_UPLOAD); IGL_TRY(direct_upload(op->in, batch, done, n, op->s_up, dev)); }
        { TraceScope ts(op->trace, TR_END); IGL_TRY(join_probe_end_chunk(op)); }
        { TraceScope ts(op->trace, TR_BEGIN); IGL_TRY(join_probe_begin_chunk(op, n, std::move(dev))); }
        { TraceScope ts(op->trace, TR_WAIT_UP); IGL_CUDA_OK(cudaEventSynchronize(op->ev_up)); }   // the caller's buffers are free again
        std::swap(op->in, op->in_alt);
        done += n;
    }
    return IGL_OK;
}

int32_t push_rows(igl_op* op, Staging& S, const ArrowArray* batch, int32_t (*flush)(igl_op*)) {
    if (!batch) return fail(IGL_ERR_INVALID_ARG, "batch is NULL");
    int64_t done = 0;
    while (done < batch->length) {
        const int64_t take = std::min<int64_t>(batch->length - done, S.chunk_rows - S.rows);
        { TraceScope ts(op->trace, TR_STAGE); IGL_TRY(S.append(batch, done, take)); }
        done += take;
        if (S.rows == S.chunk_rows) IGL_TRY(flush(op));
    }
    return IGL_OK;
}

bool is_bare_col(const uint8_t* bc, size_t len, int* col) {
    if (len == 3 && bc[0] == IGL_OP_COL) { *col = bc[1] | (bc[2] << 8); return true; }
    return false;
}

}  // namespace

extern "C" {

int32_t igl_op_filter_project_create(igl_ctx* ctx, const struct ArrowSchema* in_schema, const uint8_t* pred, size_t pred_len,
                                     const uint8_t* const* proj, const size_t* proj_len, const char* const* proj_names,
                                     int32_t nproj, igl_op** out) {
    if (!ctx || !out) return fail(IGL_ERR_INVALID_ARG, "igl_op_filter_project_create: NULL argument");
    *out = nullptr;
    std::vector<FieldMeta> fields;
    IGL_TRY(import_schema(in_schema, fields));
    if (fields.empty()) return fail(IGL_ERR_INVALID_ARG, "input schema has no columns");
    auto op = std::make_unique<igl_op>();
    op->ctx = ctx;
    DeviceGuard g(ctx->device);
    op->in.init(ctx, fields, ctx->cfg.coalesce_rows);
    const int n_in = (int)fields.size();
    std::vector<int32_t> types(n_in); std::vector<uint8_t> nulls(n_in);
    for (int i = 0; i < n_in; ++i) { types[i] = fields[i].type; nulls[i] = fields[i].nullable; }
    auto destroy_partial = [&]() { igl_op_destroy(op.release()); };
    if (pred && pred_len) {
        int32_t rc = igl_expr_compile(ctx, pred, pred_len, types.data(), nulls.data(), n_in, &op->pred);
        if (rc < 0) { destroy_partial(); return rc; }
        op->has_pred = true;
    }
    std::vector<uint8_t> skip(n_in, 0);
    std::vector<uint8_t> wanted(n_in, 0);
    std::vector<bool> used(n_in, false);
    if (nproj == 0) {   // FilterExec: schema = input schema (filter.rs:70-72)
        for (int c = 0; c < n_in; ++c) {
            op->out_fields.push_back(fields[c]);
            if (fields[c].type == IGL_T_UTF8) { skip[c] = 1; op->out_spec.push_back({true, c}); }
            else op->out_spec.push_back({false, op->n_fp_out++});
            wanted[c] = 1;
        }
    } else {
        for (int j = 0; j < nproj; ++j) {
            int c = -1;
            const char* nm = proj_names && proj_names[j] ? proj_names[j] : "";
            if (is_bare_col(proj[j], proj_len[j], &c) && c < n_in && fields[c].type == IGL_T_UTF8) {
                op->out_spec.push_back({true, c});
                op->out_fields.push_back({nm, IGL_T_UTF8, fields[c].nullable});
                wanted[c] = 1;
                continue;
            }
            igl_expr* e = nullptr;
            int32_t rc = igl_expr_compile(ctx, proj[j], proj_len[j], types.data(), nulls.data(), n_in, &e);
            if (rc < 0) { destroy_partial(); return rc; }
            op->proj_exprs.push_back(e);
            op->out_spec.push_back({false, op->n_fp_out++});
            op->out_fields.push_back({nm, e->type, e->nullable});
            for (auto& nd : e->nodes) if (nd.op == IGL_OP_COL) wanted[nd.col] = 1;
        }
    }
    if (op->pred) for (auto& nd : op->pred->nodes) if (nd.op == IGL_OP_COL) wanted[nd.col] = 1;
    op->in.wanted = wanted;
    op->in_alt = op->in;   // second (still empty) staging buffer of the same shape
    bool any_utf8 = false;
    for (auto& sp : op->out_spec) any_utf8 |= sp.utf8_pass;
    op->need_rowid = any_utf8 && op->has_pred;
    if (op->n_fp_out > 0 || op->has_pred) {
        std::vector<const igl_expr*> pe(op->proj_exprs.begin(), op->proj_exprs.end());
        // a filter whose only outputs are strings still needs the row ids: give it a plan with no value outputs
        int32_t rc;
        if (nproj == 0) rc = fp_create_ex(ctx, types.data(), nulls.data(), n_in, op->pred, nullptr, 0, skip.data(), op->need_rowid, &op->fp);
        else if (!pe.empty()) rc = fp_create_ex(ctx, types.data(), nulls.data(), n_in, op->pred, pe.data(), (int32_t)pe.size(), nullptr, op->need_rowid, &op->fp);
        else {   // projection of string columns only, under a predicate
            std::fill(skip.begin(), skip.end(), 1);
            rc = fp_create_ex(ctx, types.data(), nulls.data(), n_in, op->pred, nullptr, 0, skip.data(), op->need_rowid, &op->fp);
        }
        if (rc < 0) { destroy_partial(); return rc; }
    }
    *out = op.release();
    return IGL_OK;
}

int32_t igl_op_hash_join_create(igl_ctx* ctx, int32_t join_type, const struct ArrowSchema* left_schema,
                                const struct ArrowSchema* right_schema, const uint8_t* left_key, size_t left_key_len,
                                const uint8_t* right_key, size_t right_key_len, igl_op** out) {
    if (!left_key || !right_key) return fail(IGL_ERR_INVALID_ARG, "igl_op_hash_join_create: NULL argument");
    return igl_op_hash_join_create_ex(ctx, join_type, left_schema, right_schema, &left_key, &left_key_len, &right_key, &right_key_len, 1, out);
}

int32_t igl_op_hash_join_create_ex(igl_ctx* ctx, int32_t join_type, const struct ArrowSchema* left_schema,
                                   const struct ArrowSchema* right_schema, const uint8_t* const* left_keys, const size_t* left_key_lens,
                                   const uint8_t* const* right_keys, const size_t* right_key_lens, int32_t nkeys, igl_op** out) {
    if (!ctx || !out || !left_keys || !right_keys || !left_key_lens || !right_key_lens || nkeys < 1)
        return fail(IGL_ERR_INVALID_ARG, "igl_op_hash_join_create: NULL argument");
    *out = nullptr;
    std::vector<FieldMeta> lf, rf;
    IGL_TRY(import_schema(left_schema, lf));
    IGL_TRY(import_schema(right_schema, rf));
    auto op = std::make_unique<igl_op>();
    op->ctx = ctx; op->is_join = true;
    DeviceGuard g(ctx->device);
    op->n_left = (int)lf.size(); op->n_right = (int)rf.size();
    op->in.init(ctx, lf, ctx->cfg.coalesce_rows);
    op->build.init(ctx, rf, ctx->cfg.coalesce_rows);
    std::vector<int32_t> lt, rt; std::vector<uint8_t> ln, rn;
    for (auto& f : lf) { lt.push_back(f.type); ln.push_back(f.nullable); }
    for (auto& f : rf) { rt.push_back(f.type); rn.push_back(f.nullable); }
    auto destroy_partial = [&]() { igl_op_destroy(op.release()); };
    // Utf8 columns: the device join gets an Int32 row-id column in their place (never NULL itself; the
    // string's own validity is applied when the strings are gathered)
    op->l_utf8.assign(op->n_left, 0); op->r_utf8.assign(op->n_right, 0);
    for (int c = 0; c < op->n_left; ++c) if (lt[c] == IGL_T_UTF8) { op->l_utf8[c] = 1; op->any_utf8 = true; lt[c] = IGL_T_I32; ln[c] = 0; }
    for (int c = 0; c < op->n_right; ++c) if (rt[c] == IGL_T_UTF8) { op->r_utf8[c] = 1; op->any_utf8 = true; rt[c] = IGL_T_I32; rn[c] = 0; }
    int32_t rc = IGL_OK;
    // string keys first, against the real schema: a bare Utf8 column on both sides of a pair
    std::vector<int> key_utf8(nkeys, -1);
    {
        std::vector<int32_t> lt0, rt0; std::vector<uint8_t> ln0, rn0;
        for (auto& f : lf) { lt0.push_back(f.type); ln0.push_back(f.nullable); }
        for (auto& f : rf) { rt0.push_back(f.type); rn0.push_back(f.nullable); }
        for (int k = 0; k < nkeys && rc >= 0; ++k) {
            igl_expr *le = nullptr, *re = nullptr;
            rc = igl_expr_compile(ctx, left_keys[k], left_key_lens[k], lt0.data(), ln0.data(), op->n_left, &le);
            if (rc >= 0) rc = igl_expr_compile(ctx, right_keys[k], right_key_lens[k], rt0.data(), rn0.data(), op->n_right, &re);
            if (rc >= 0 && (le->type == IGL_T_UTF8) != (re->type == IGL_T_UTF8))
                rc = fail(IGL_ERR_UNSUPPORTED, "join key %d pairs a Utf8 key with a key of another type (such a pair never matches: hash_join.rs:120 prints the type); not supported on the GPU", k);
            if (rc >= 0 && le->type == IGL_T_UTF8) {
                if (!le->is_plain_col() || !re->is_plain_col()) rc = fail(IGL_ERR_UNSUPPORTED, "Utf8 join key %d must be a bare column", k);
                else { key_utf8[k] = (int)op->utf8_keys.size(); op->utf8_keys.push_back({le->nodes[0].col, re->nodes[0].col}); }
            }
            igl_expr_destroy(le); igl_expr_destroy(re);
        }
        if (rc < 0) { destroy_partial(); return rc; }
    }
    const int nu = (int)op->utf8_keys.size();
    for (int i = 0; i < nu; ++i) {   // the hash columns follow each side's own columns
        lt.push_back(IGL_T_I64); ln.push_back(lf[op->utf8_keys[i].lcol].nullable);
        rt.push_back(IGL_T_I64); rn.push_back(rf[op->utf8_keys[i].rcol].nullable);
    }
    if (const char* e = getenv("IGL_UTF8_KEY_HASH_BITS")) op->utf8_hash_bits = std::min(64, std::max(1, atoi(e)));
    for (int k = 0; k < nkeys && rc >= 0; ++k) {   // one pair per `join.on` entry (physical_planner.rs:106-120)
        igl_expr *le = nullptr, *re = nullptr;
        if (key_utf8[k] >= 0) {
            const int lc = op->n_left + key_utf8[k], rc2 = op->n_right + key_utf8[k];
            const uint8_t lb[3] = {IGL_OP_COL, (uint8_t)(lc & 0xFF), (uint8_t)(lc >> 8)}, rb[3] = {IGL_OP_COL, (uint8_t)(rc2 & 0xFF), (uint8_t)(rc2 >> 8)};
            rc = igl_expr_compile(ctx, lb, 3, lt.data(), ln.data(), op->n_left + nu, &le);
            if (rc >= 0) rc = igl_expr_compile(ctx, rb, 3, rt.data(), rn.data(), op->n_right + nu, &re);
        } else {
            rc = igl_expr_compile(ctx, left_keys[k], left_key_lens[k], lt.data(), ln.data(), op->n_left + nu, &le);
            if (rc >= 0) rc = igl_expr_compile(ctx, right_keys[k], right_key_lens[k], rt.data(), rn.data(), op->n_right + nu, &re);
        }
        if (le) op->lkeys.push_back(le);
        if (re) op->rkeys.push_back(re);
    }
    if (rc >= 0) {
        op->lkey = op->lkeys[0]; op->rkey = op->rkeys[0];
        rc = igl_join_create_ex(ctx, join_type, lt.data(), ln.data(), op->n_left + nu, rt.data(), rn.data(), op->n_right + nu,
                                op->lkeys.data(), op->rkeys.data(), nkeys, &op->join);
    }
    if (rc >= 0 && nu > 0) {
        // what removes a colliding pair from the output: the output columns as the device join writes them
        // (row ids where the strings are) plus one Int32 keep flag; keeps rows with flag != 0
        std::vector<int32_t> ot(lt.begin(), lt.begin() + op->n_left);
        ot.insert(ot.end(), rt.begin(), rt.begin() + op->n_right);
        std::vector<uint8_t> on(ln.begin(), ln.begin() + op->n_left);
        on.insert(on.end(), rn.begin(), rn.begin() + op->n_right);
        const int kc = (int)ot.size();
        ot.push_back(IGL_T_I32); on.push_back(0);
        std::vector<uint8_t> skip(ot.size(), 0);
        skip[kc] = 1;
        const uint8_t bc[9] = {IGL_OP_COL, (uint8_t)(kc & 0xFF), (uint8_t)(kc >> 8), IGL_OP_LIT_I32, 0, 0, 0, 0, IGL_OP_NE};
        igl_expr* pred = nullptr;
        rc = igl_expr_compile(ctx, bc, sizeof bc, ot.data(), on.data(), (int32_t)ot.size(), &pred);
        if (rc >= 0) rc = fp_create_ex(ctx, ot.data(), on.data(), (int32_t)ot.size(), pred, nullptr, 0, skip.data(), false, &op->collide_fp);
        igl_expr_destroy(pred);
        if (rc >= 0 && cudaHostAlloc((void**)&op->h_mismatch, sizeof(unsigned long long), cudaHostAllocDefault) != cudaSuccess) rc = fail(IGL_ERR_OOM, "pinned allocation failed");
        if (rc >= 0) { void* p = nullptr; rc = pool_alloc(ctx, 256, &p); op->d_mismatch = (unsigned long long*)p; }
        op->build_hash.assign(nu, nullptr);
    }
    if (rc < 0) { destroy_partial(); return rc; }
    // output schema: hash_join.rs:35-76 (left fields, right fields, clashes renamed right_<name>)
    op->left_join = join_type == IGL_JOIN_LEFT || join_type == IGL_JOIN_FULL;
    for (auto& f : lf) op->out_fields.push_back(f);
    for (auto& f : rf) {
        FieldMeta o = f;
        for (auto& l : lf) if (l.name == f.name) { o.name = "right_" + f.name; break; }
        // The reference keeps the declared nullability (hash_join.rs:57-61), so its
        // null-extended rows of a Left/Full join fail RecordBatch::try_new when the
        // right field was declared non-nullable.  Here such fields become nullable.
        if (op->left_join) o.nullable = true;
        op->out_fields.push_back(o);
    }
    if (!op->left_join && nkeys == 1 && nu == 0 && op->lkey->is_plain_col() && op->rkey->is_plain_col()) {
        const int lc = op->lkey->nodes[0].col, rc2 = op->rkey->nodes[0].col;
        if (lt[lc] == rt[rc2] && (lt[lc] == IGL_T_I32 || lt[lc] == IGL_T_I64)) { op->alias_left_key = lc; op->alias_right_key = rc2; }
    }
    op->build_dev.assign(op->n_right, nullptr);
    op->build_cap.assign(op->n_right, 0);
    op->build_valid.assign(op->n_right, nullptr);
    op->build_any_null.assign(op->n_right, 0);
    op->build_str.assign(op->n_right, nullptr); op->build_off.assign(op->n_right, nullptr);
    op->build_str_cap.assign(op->n_right, 0); op->build_str_used.assign(op->n_right, 0);
    op->in.elide_all_valid = true;
    op->in_alt = op->in;   // second probe chunk buffer of the same shape
    if (cudaStreamCreateWithFlags(&op->s_up, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&op->s_down, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&op->ev_up, cudaEventDisableTiming) != cudaSuccess) {
        destroy_partial();
        return fail(IGL_ERR_CUDA, "igl_op_hash_join_create: stream / event creation failed");
    }
    *out = op.release();
    return IGL_OK;
}

int32_t igl_op_schema(igl_op* op, struct ArrowSchema* out) {
    if (!op || !out) return fail(IGL_ERR_INVALID_ARG, "igl_op_schema: NULL argument");
    export_schema(op->out_fields, out);
    return IGL_OK;
}

int32_t igl_op_push_build(igl_op* op, const struct ArrowArray* batch) {
    if (!op || !op->is_join) return fail(IGL_ERR_INVALID_ARG, "igl_op_push_build: not a join operator");
    if (op->build_done) return fail(IGL_ERR_INVALID_ARG, "igl_op_push_build: build side already finished");
    TraceScope ts_build(op->trace, TR_BUILD);
    DeviceGuard g(op->ctx->device);
    if (batch && batch_is_direct(op->build, batch) && (op->build_rows & 7) == 0) {
        // page-locked batch: straight from the caller's buffers to the end of the device columns
        Staging& S = op->build;
        const int64_t n = batch->length;
        IGL_TRY(join_build_reserve(op, n));
        cudaStream_t st = op->ctx->stream;
        for (int c = 0; c < op->n_right; ++c) {
            const ArrowArray* a = batch->children[c];
            const int w = type_width(S.fields[c].type);
            IGL_CUDA_OK(cudaMemcpyAsync((uint8_t*)op->build_dev[c] + (size_t)op->build_rows * w,
                                        (const uint8_t*)a->buffers[1] + (size_t)(a->offset + batch->offset) * w, (size_t)n * w, cudaMemcpyHostToDevice, st));
            if (S.fields[c].nullable)   // no nulls in a direct batch: all ones (a partial last byte is completed by the next append)
                IGL_CUDA_OK(cudaMemsetAsync(op->build_valid[c] + (op->build_rows >> 3), 0xFF, (size_t)(n + 7) / 8, st));
        }
        IGL_CUDA_OK(cudaStreamSynchronize(st));   // the batch is only borrowed for this call
        op->build_rows += n;
        return IGL_OK;
    }
    return push_rows(op, op->build, batch, join_build_flush);
}

int32_t igl_op_finish_build(igl_op* op) {
    if (!op || !op->is_join) return fail(IGL_ERR_INVALID_ARG, "igl_op_finish_build: not a join operator");
    if (op->build_done) return IGL_OK;
    DeviceGuard g(op->ctx->device);
    IGL_TRY(join_build_flush(op));
    for (int c = 0; c < op->n_right; ++c)   // a nullable column that never carried a null: no bitmap, non-null kernels
        if (op->build_valid[c] && !op->build_any_null[c]) { pool_free(op->ctx, op->build_valid[c]); op->build_valid[c] = nullptr; }
    if (op->any_utf8) IGL_TRY(ensure_iota(op, std::max<int64_t>(op->build_rows, op->in.chunk_rows)));
    std::vector<igl_dev_col> cols(op->n_right);
    for (int c = 0; c < op->n_right; ++c) {
        cols[c].type = op->build.fields[c].type;
        if (op->r_utf8[c]) {   // the join carries the build row number; the strings follow it after the probe
            cols[c] = igl_dev_col{IGL_T_I32, 0, op->d_iota, nullptr, nullptr};
            if (!op->build_off[c]) {   // an empty build side still needs a valid offsets array
                void* p = nullptr; IGL_TRY(pool_alloc(op->ctx, 256, &p)); op->build_off[c] = p;
                IGL_CUDA_OK(cudaMemsetAsync(p, 0, 256, op->ctx->stream));
            }
            continue;
        }
        if (!op->build_dev[c]) { void* p = nullptr; IGL_TRY(pool_alloc(op->ctx, 256, &p)); op->build_dev[c] = p; }
        cols[c].data = op->build_dev[c];
        cols[c].validity = op->build_valid[c];
    }
    for (size_t i = 0; i < op->utf8_keys.size(); ++i) {   // string keys: the build side's hash columns, kept with the table
        const int c = op->utf8_keys[i].rcol;
        if (!op->build_hash[i]) IGL_TRY(pool_alloc(op->ctx, pad64((size_t)std::max<int64_t>(op->build_rows, 1) * 8), &op->build_hash[i]));
        IGL_TRY(utf8_hash_keys(op->ctx, (const int32_t*)op->build_off[c], (const uint8_t*)op->build_str[c], op->build_rows, op->utf8_hash_bits,
                               (int64_t*)op->build_hash[i], op->ctx->stream));
        cols.push_back(igl_dev_col{IGL_T_I64, 0, op->build_hash[i], op->build_valid[c], nullptr});
    }
    IGL_TRY(igl_join_build(op->join, cols.data(), op->build_rows, op->ctx->stream));
    op->build_done = true;
    return IGL_OK;
}

int32_t igl_op_push(igl_op* op, const struct ArrowArray* batch) {
    if (!op) return fail(IGL_ERR_INVALID_ARG, "igl_op_push: NULL operator");
    TraceScope ts_push(op->trace, TR_PUSH);
    if (op->input_done) return fail(IGL_ERR_INVALID_ARG, "igl_op_push: input already finished");
    DeviceGuard g(op->ctx->device);
    if (op->is_join) {
        if (!op->build_done) return fail(IGL_ERR_INVALID_ARG, "igl_op_push: finish the build side first (hash_join.rs:100-128 drains the right child before probing)");
        if (batch && batch_is_direct(op->in, batch)) return join_probe_direct(op, batch);
        return push_rows(op, op->in, batch, join_probe_flush);
    }
    return push_rows(op, op->in, batch, fp_flush);
}

int32_t igl_op_finish_input(igl_op* op) {
    if (!op) return fail(IGL_ERR_INVALID_ARG, "igl_op_finish_input: NULL operator");
    if (op->input_done) return IGL_OK;
    DeviceGuard g(op->ctx->device);
    if (op->is_join) {
        if (!op->build_done) IGL_TRY(igl_op_finish_build(op));
        IGL_TRY(join_probe_flush(op));
        IGL_TRY(join_probe_end_chunk(op));
    } else {
        IGL_TRY(fp_flush(op));
        IGL_TRY(fp_end(op));
    }
    op->input_done = true;
    return IGL_OK;
}

int32_t igl_op_pull(igl_op* op, struct ArrowArray* out) {
    if (!op || !out) return fail(IGL_ERR_INVALID_ARG, "igl_op_pull: NULL argument");
    TraceScope ts_pull(op->trace, TR_PULL);
    if (op->ready.empty()) return op->input_done ? IGL_EOS : IGL_AGAIN;
    OutBatch& f = op->ready.front();
    if (f.done) {
        // still travelling: while input is flowing the caller should push on (the copy overlaps the
        // next chunk's upload); once the input has ended there is nothing better to do than wait
        DeviceGuard g(op->ctx->device);
        if (!op->input_done && cudaEventQuery(f.done) == cudaErrorNotReady) return IGL_AGAIN;
        (void)cudaGetLastError();
        if (cudaEventSynchronize(f.done) != cudaSuccess) return fail(IGL_ERR_CUDA, "output download failed: %s", cudaGetErrorString(cudaGetLastError()));
        cudaEventDestroy(f.done);
        f.done = nullptr;
        for (void* p : f.dev_free) pool_free(op->ctx, p);
        f.dev_free.clear();
        for (auto& c : f.cols) c.null_count = c.valid ? count_zero_bits(c.valid, f.nrows) : 0;
    }
    OutBatch b = std::move(op->ready.front());
    op->ready.pop_front();
    export_batch(std::move(b), out);
    return IGL_OK;
}

void igl_op_destroy(igl_op* op) {
    if (!op) return;
    if (op->trace.on) {
        static const char* nm[8] = {"stage", "end_chunk", "upload", "begin_chunk", "wait_upload", "pull", "push_build", "push"};
        fprintf(stderr, "[igl_op trace]");
        for (int i = 0; i < 8; ++i) fprintf(stderr, " %s=%.1fms/%ld", nm[i], op->trace.t[i] * 1e3, op->trace.n[i]);
        fprintf(stderr, "\n");
    }
    DeviceGuard g(op->ctx->device);
    if (op->s_down) cudaStreamSynchronize(op->s_down);
    if (op->s_up) cudaStreamSynchronize(op->s_up);
    if (op->jpend.active) {
        int64_t dummy = 0;
        join_probe_end(op->join, &dummy, op->ctx->stream);
        for (void* p : op->jpend.to_free) pool_free(op->ctx, p);
    }
    for (auto& b : op->ready) {
        if (b.done) cudaEventDestroy(b.done);
        for (void* p : b.dev_free) pool_free(op->ctx, p);
        for (auto& c : b.cols) { if (!c.alias) pinned_pool().put(c.data); pinned_pool().put(c.valid); pinned_pool().put(c.off); }
    }
    if (op->s_up) cudaStreamDestroy(op->s_up);
    if (op->s_down) cudaStreamDestroy(op->s_down);
    if (op->ev_up) cudaEventDestroy(op->ev_up);
    if (op->pend.active) { cudaStreamSynchronize(op->ctx->stream); for (void* p : op->pend.to_free) pool_free(op->ctx, p); }
    op->in.destroy();
    op->in_alt.destroy();
    op->build.destroy();
    if (op->fp) igl_fp_destroy(op->fp);
    if (op->join) igl_join_destroy(op->join);
    igl_expr_destroy(op->pred);
    for (auto* e : op->proj_exprs) igl_expr_destroy(e);
    for (auto* e : op->lkeys) igl_expr_destroy(e);
    for (auto* e : op->rkeys) igl_expr_destroy(e);
    for (void* p : op->build_dev) pool_free(op->ctx, p);
    for (void* p : op->build_str) pool_free(op->ctx, p);
    for (void* p : op->build_off) pool_free(op->ctx, p);
    for (void* p : op->build_hash) pool_free(op->ctx, p);
    if (op->collide_fp) igl_fp_destroy(op->collide_fp);
    if (op->h_mismatch) cudaFreeHost(op->h_mismatch);
    pool_free(op->ctx, op->d_mismatch);
    pool_free(op->ctx, op->d_iota);
    for (uint8_t* p : op->build_valid) pool_free(op->ctx, p);
    delete op;
}

}  // extern "C"
