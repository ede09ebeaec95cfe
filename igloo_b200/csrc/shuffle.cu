// shuffle.cu — the hash join across the GPUs of one node, behind the C ABI
// (igl_comm_init, igl_join_shuffle_*): the data plane of the "worker fragment
// exchange" the reference only declares.
//
// Reference anchors: FragmentType::Shuffle is never constructed
// (crates/coordinator/src/fragment.rs:8-13); DistributedPlanner pins the join fragment to
// the coordinator and feeds it from scan fragments (distributed_planner.rs:65-92);
// GetDataForTask ("data fetch request for shuffle") returns empty bytes
// (crates/worker/src/service.rs:26-32).  A worker is one process (worker/src/main.rs:14-52);
// here it is one process per GPU, each calling igl_join_shuffle_run with its shard of both
// sides.
//
// One join, per rank (all device work on the caller's stream):
//   1. histogram of both sides over bins = (destination GPU, partition of the join hash)
//   2. ONE ncclAllGather of a small message (row counts, histograms, IPC handles of the
//      staging buffers): the only collective on the control path.  Every rank then derives
//      the same plan on its host (igl_shuffle_plan).
//   3. both sides are radix-partitioned LOCALLY into staging buffers, bin by bin
//      (partition.cu: HBM speed, long runs — nothing crosses NVLink yet)
//   4. a 4-byte ncclAllReduce: every rank's staging is complete
//   5. build side: the runs (partition p, source s) destined for this GPU are pulled from the
//      peers' staging buffers into one local slab (seg_copy_kernel), partition by partition,
//      and the table is built from it
//   6. probe side: join_probe_flat_kernel<.., SEG> walks the same kind of run list and stages
//      each 128-row chunk with a bulk async copy STRAIGHT FROM THE PEER'S STAGING BUFFER into
//      shared memory — the exchange of the probe side and the probe are one kernel, the
//      NVLink transfer of a chunk overlaps the lookups of the chunks before it, and no probe
//      row is written to a receive buffer.  Joins the flat kernel does not serve (Left/Full,
//      duplicate build keys) pull the probe side into a local slab first and take the
//      general probe.
//   mode IGL_EXCHANGE_NCCL replaces 4-6's peer reads by grouped ncclSend/ncclRecv of each
//   destination's contiguous staging range into local receive slabs (the all-to-all-v the
//   north_star names); the same run lists then point into those slabs.
//
// dest = low 32 bits of mix64(key) scaled to the rank count; partitions use the TOP bits, the
// bits the local table is indexed with, so rows arrive in the order the local join wants.
#include "internal.h"
#include "kernels.h"
#include "device_utils.cuh"

#include <dlfcn.h>

#include <algorithm>
#include <cstring>
#include <map>
#include <memory>

namespace igl {

// ---- NCCL, resolved at run time (like NVRTC): the library has no link-time dependency on it ----
typedef struct ncclComm* ncclComm_t;
struct NcclUniqueId { char internal[128]; };
enum { kNcclSuccess = 0, kNcclInt8 = 0, kNcclUint8 = 1, kNcclInt32 = 2, kNcclSum = 0 };

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    const char* (*GetLastError)(ncclComm_t) = nullptr;
};

static std::mutex g_nccl_mu;
static NcclApi g_nccl;

static int32_t nccl_load() {
    std::lock_guard<std::mutex> g(g_nccl_mu);
    if (g_nccl.handle) return IGL_OK;
    const char* names[] = {getenv("IGL_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names) {
        if (!n || !*n) continue;
        h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return fail(IGL_ERR_NCCL, "NCCL is not available: dlopen(libnccl.so.2) failed (%s); set IGL_NCCL_LIB", dlerror());
    NcclApi a;
    a.handle = h;
#define IGL_NCCL_SYM(field, sym)                                                              \
    *(void**)(&a.field) = dlsym(h, sym);                                                      \
    if (!a.field) return fail(IGL_ERR_NCCL, "NCCL library lacks the symbol %s", sym)
    IGL_NCCL_SYM(GetUniqueId, "ncclGetUniqueId");
    IGL_NCCL_SYM(CommInitRank, "ncclCommInitRank");
    IGL_NCCL_SYM(CommDestroy, "ncclCommDestroy");
    IGL_NCCL_SYM(AllGather, "ncclAllGather");
    IGL_NCCL_SYM(AllReduce, "ncclAllReduce");
    IGL_NCCL_SYM(Send, "ncclSend");
    IGL_NCCL_SYM(Recv, "ncclRecv");
    IGL_NCCL_SYM(GroupStart, "ncclGroupStart");
    IGL_NCCL_SYM(GroupEnd, "ncclGroupEnd");
    IGL_NCCL_SYM(GetErrorString, "ncclGetErrorString");
#undef IGL_NCCL_SYM
    *(void**)(&a.GetLastError) = dlsym(h, "ncclGetLastError");   // optional (NCCL >= 2.13)
    g_nccl = a;
    return IGL_OK;
}

struct Comm {
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0;
};

#define IGL_NCCL_OK(call)                                                                          \
    do {                                                                                           \
        int r__ = (call);                                                                          \
        if (r__ != kNcclSuccess)                                                                   \
            return igl::fail(IGL_ERR_NCCL, "%s failed: %s (%s:%d)", #call, g_nccl.GetErrorString(r__), __FILE__, __LINE__); \
    } while (0)

void comm_release(igl_ctx* ctx) {
    if (!ctx || !ctx->comm) return;
    Comm* c = (Comm*)ctx->comm;
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    delete c;
    ctx->comm = nullptr;
}

// ---- the plan: pure host arithmetic on the gathered histograms ---------------------------------
// A source lays its staging buffer out bin by bin (destination-major), every bin starting on
// a multiple of kSegAlignRows rows so that 128-row chunks of a run can be fetched with
// 16-byte-granular bulk copies.
static inline int64_t align_rows(int64_t n) { return (n + kSegAlignRows - 1) / kSegAlignRows * kSegAlignRows; }

}  // namespace igl

using namespace igl;

extern "C" {

int32_t igl_shuffle_plan(int32_t nranks, int32_t rank, int32_t log2_parts, const int64_t* counts, int64_t* src_offsets,
                         int32_t* seg_src, int64_t* seg_row0, int64_t* seg_rows, int64_t* total_rows) {
    return igl_shuffle_plan_ex(nranks, rank, log2_parts, 0, counts, src_offsets, seg_src, seg_row0, seg_rows, total_rows);
}

int32_t igl_shuffle_plan_ex(int32_t nranks, int32_t rank, int32_t log2_parts, int32_t broadcast, const int64_t* counts,
                            int64_t* src_offsets, int32_t* seg_src, int64_t* seg_row0, int64_t* seg_rows, int64_t* total_rows) {
    if (nranks < 1 || nranks > 16 || rank < 0 || rank >= nranks || log2_parts < 0 || broadcast < 0 || broadcast > 1 ||
        ((int64_t)(nranks + broadcast) << log2_parts) > 1024 || !counts || !src_offsets)
        return fail(IGL_ERR_INVALID_ARG, "igl_shuffle_plan: bad arguments");
    const int parts = 1 << log2_parts, bins = (nranks + broadcast) << log2_parts;
    for (int s = 0; s < nranks; ++s) {
        int64_t off = 0;
        for (int b = 0; b < bins; ++b) {
            const int64_t c = counts[(size_t)s * bins + b];
            if (c < 0) return fail(IGL_ERR_INVALID_ARG, "igl_shuffle_plan: negative count");
            src_offsets[(size_t)s * bins + b] = off;
            off += align_rows(c);
        }
    }
    int64_t total = 0;
    const int per_part = nranks * (1 + broadcast);
    for (int p = 0; p < parts; ++p)
        for (int k = 0; k < 1 + broadcast; ++k)        // the rank's own bins, then the "every rank reads this" bins
            for (int s = 0; s < nranks; ++s) {
                const size_t b = (size_t)s * bins + ((size_t)(k ? nranks : rank) << log2_parts) + p;
                const int i = p * per_part + k * nranks + s;
                if (seg_src) seg_src[i] = s;
                if (seg_row0) seg_row0[i] = src_offsets[b];
                if (seg_rows) seg_rows[i] = counts[b];
                total += counts[b];
            }
    if (total_rows) *total_rows = total;
    return IGL_OK;
}

}  // extern "C"

namespace igl {

// ---- pulling runs into one local buffer ----------------------------------------------------------
struct CopySeg {
    uint64_t src_row0;
    uint64_t dst_row0;
    uint32_t rows;
    uint32_t blk0;     // first block of this run in the virtual block space
    uint32_t src;
    uint32_t pad;
};
constexpr int kCopyBlockRows = 2048;
constexpr int kCopyThreads = 256;
struct CopyParams {
    const uint8_t* src[kSegMaxSrc][16];
    uint8_t* dst[16];
    uint8_t shift[16];
    int ncols, nsegs;
    const CopySeg* segs;
    uint32_t nblocks;
};

// Every CTA takes 2048-row blocks of the run list; a run may sit in a peer GPU's memory, so
// these loads are the NVLink transfer (8 independent 8-byte loads per thread and column in
// flight).  Sources start on 16-byte boundaries; destinations are packed, hence per-element.
__global__ void __launch_bounds__(kCopyThreads) seg_copy_kernel(const __grid_constant__ CopyParams P) {
    for (uint32_t blk = blockIdx.x; blk < P.nblocks; blk += gridDim.x) {
        int lo = 0, hi = P.nsegs - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (__ldg(&P.segs[mid].blk0) <= blk) lo = mid; else hi = mid - 1;
        }
        const CopySeg sg = P.segs[lo];
        const uint32_t r0 = (blk - sg.blk0) * kCopyBlockRows;
        const int n = (int)min((uint32_t)kCopyBlockRows, sg.rows - r0);
        for (int c = 0; c < P.ncols; ++c) {
            if (P.shift[c] == 3) {
                const uint64_t* s = (const uint64_t*)P.src[sg.src][c] + sg.src_row0 + r0;
                uint64_t* d = (uint64_t*)P.dst[c] + sg.dst_row0 + r0;
                uint64_t v[kCopyBlockRows / kCopyThreads];
#pragma unroll
                for (int u = 0; u < kCopyBlockRows / kCopyThreads; ++u) { const int i = u * kCopyThreads + threadIdx.x; if (i < n) v[u] = s[i]; }
#pragma unroll
                for (int u = 0; u < kCopyBlockRows / kCopyThreads; ++u) { const int i = u * kCopyThreads + threadIdx.x; if (i < n) d[i] = v[u]; }
            } else {
                const uint32_t* s = (const uint32_t*)P.src[sg.src][c] + sg.src_row0 + r0;
                uint32_t* d = (uint32_t*)P.dst[c] + sg.dst_row0 + r0;
                uint32_t v[kCopyBlockRows / kCopyThreads];
#pragma unroll
                for (int u = 0; u < kCopyBlockRows / kCopyThreads; ++u) { const int i = u * kCopyThreads + threadIdx.x; if (i < n) v[u] = s[i]; }
#pragma unroll
                for (int u = 0; u < kCopyBlockRows / kCopyThreads; ++u) { const int i = u * kCopyThreads + threadIdx.x; if (i < n) d[i] = v[u]; }
            }
        }
    }
}

// ---- a few CTAs acting as copy engines ------------------------------------------------------------
// Moves contiguous byte ranges (peer staging -> local receive slab) with bulk async copies only: one
// thread per CTA keeps kDmaAhead 16 KB loads in flight into a shared-memory ring and writes each
// piece out with a bulk store as it lands.  No registers, no issue slots to speak of: the kernel
// takes the shared memory of the few SMs it sits on and leaves the rest of the GPU to whatever
// else is running (the table build).  TMA multicast loads here read only this GPU's memory, since
// on B200 a TMA multicast load from another GPU's memory hung the GPU — these are plain bulk copies.
constexpr uint32_t kDmaPiece = 16u << 10;
constexpr int kDmaSlots = 12, kDmaAhead = 8;
constexpr int kDmaMaxRuns = 64;
struct DmaParams {
    const uint8_t* src[kDmaMaxRuns];
    uint8_t* dst[kDmaMaxRuns];
    uint64_t bytes[kDmaMaxRuns];
    uint32_t piece0[kDmaMaxRuns + 1];
    int nruns;
};

__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

__global__ void __launch_bounds__(32) dma_copy_kernel(const __grid_constant__ DmaParams P) {
    extern __shared__ __align__(128) uint8_t ring[];
    __shared__ uint64_t s_full[kDmaSlots];
    if (threadIdx.x != 0) return;
    for (int k = 0; k < kDmaSlots; ++k) mbar_init(&s_full[k], 1);
    mbar_fence_init();
    const uint32_t npieces = P.piece0[P.nruns];
    const uint32_t mine = blockIdx.x < npieces ? (npieces - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;   // pieces blockIdx.x, + gridDim.x, ...
    int run = 0;
    auto load = [&](uint32_t t) {   // t-th piece of this CTA -> slot t % kDmaSlots (pieces are visited in rising order: `run` only moves forward)
        const uint32_t piece = blockIdx.x + t * gridDim.x;
        while (P.piece0[run + 1] <= piece) ++run;
        const uint64_t off = (uint64_t)(piece - P.piece0[run]) * kDmaPiece;
        const uint32_t n = (uint32_t)min((uint64_t)kDmaPiece, P.bytes[run] - off);
        uint64_t* bar = &s_full[t % kDmaSlots];
        mbar_arrive_expect_tx(bar, n);
        bulk_g2s(ring + (size_t)(t % kDmaSlots) * kDmaPiece, P.src[run] + off, n, bar);
    };
    for (uint32_t t = 0; t < mine && t < (uint32_t)kDmaAhead; ++t) load(t);
    int srun = 0;
    for (uint32_t t = 0; t < mine; ++t) {
        const uint32_t piece = blockIdx.x + t * gridDim.x;
        while (P.piece0[srun + 1] <= piece) ++srun;
        const uint64_t off = (uint64_t)(piece - P.piece0[srun]) * kDmaPiece;
        const uint32_t n = (uint32_t)min((uint64_t)kDmaPiece, P.bytes[srun] - off);
        mbar_wait(&s_full[t % kDmaSlots], (t / kDmaSlots) & 1u);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        bulk_s2g(P.dst[srun] + off, ring + (size_t)(t % kDmaSlots) * kDmaPiece, n);
        if (t + kDmaAhead < mine) {
            // the slot of piece t + kDmaAhead was last read by the store of piece t + kDmaAhead - kDmaSlots:
            // kDmaSlots - kDmaAhead stores have been committed after that one
            asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kDmaSlots - kDmaAhead) : "memory");
            load(t + kDmaAhead);
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

constexpr int kMsgBins = 1024;
constexpr int kShufMaxCols = 16;

// What every rank tells every other rank at the start of a join (one ncclAllGather).
struct ShuffleMsg {
    int64_t n_rows[2];                      // [0] right / build, [1] left / probe: rows of this rank's shard
    uint64_t gen[2];                        // bumped whenever a side's staging buffers were reallocated
    uint8_t handles[2][kShufMaxCols][64];   // CUDA IPC handles of the staging buffers (p2p mode)
    unsigned long long hist[2][kMsgBins];   // rows per (destination, partition) at the finest partition count
};

struct ShuffleSide {
    std::vector<int> types;
    std::vector<uint8_t> shift;
    int ncols = 0, key = 0;
    std::vector<void*> staging;             // this rank's rows, bin by bin
    int64_t staging_rows = 0;
    uint64_t gen = 0;
    std::vector<void*> retired;             // replaced staging blocks: peers may still be reading them until the next all-gather
    struct Peer { uint64_t gen = 0; bool open = false; std::vector<void*> ptrs; } peer[16];
    std::vector<void*> slab;                // pulled runs, one contiguous column set (build side; probe side in the general path)
    int64_t slab_rows = 0;
    std::vector<void*> recv;                // nccl mode: receive slabs
    int64_t recv_rows = 0;
    int row_bytes() const { int b = 0; for (int t : types) b += t == IGL_T_I32 ? 4 : 8; return b; }
};

}  // namespace igl

struct igl_join_shuffle {
    igl_ctx* ctx = nullptr;
    igl_join* join = nullptr;
    int mode = IGL_EXCHANGE_P2P;
    int join_type = 0;
    bool left_join = false;
    ShuffleSide side[2];                    // [0] right / build, [1] left / probe
    ShuffleMsg* h_send = nullptr;           // pinned
    ShuffleMsg* h_recv = nullptr;           // pinned, [nranks]
    ShuffleMsg* d_send = nullptr;
    ShuffleMsg* d_recv = nullptr;
    unsigned long long* d_plan = nullptr;   // [2][kMsgBins] scatter offsets, then [2][kMsgBins] cursors
    unsigned long long* h_plan = nullptr;   // pinned [2][kMsgBins]
    void* d_segs = nullptr;                 // run lists (device)
    void* h_segs = nullptr;                 // pinned
    size_t segs_bytes = 0;
    int* d_flag = nullptr;
    // hot keys (skewed probe sides): device set, sample and candidate buffers
    unsigned long long* d_hot = nullptr;       // [kHotSlots] open-addressing set of canonical keys
    unsigned long long* d_samples = nullptr;   // [kHotSamples] then [2 * kHotCand] candidates to send, then [nranks][2 * kHotCand] received
    unsigned long long* h_samples = nullptr;   // pinned mirror of d_samples
    std::vector<void*> out;                 // library-owned output columns (+ validity for a Left/Full join)
    std::vector<void*> out_valid;
    int64_t out_cap = 0;
    int64_t out_rows = 0;
    cudaEvent_t ev[7] = {};
    cudaEvent_t ev_plan = nullptr, ev_wait = nullptr, ev_l[3] = {};   // second stream: probe-side partition start / end / barrier end
    cudaStream_t stream2 = nullptr, stream3 = nullptr, stream_hi = nullptr;   // stream_hi: highest priority (the build side's barrier, pull and table build)
    cudaEvent_t ev_copy = nullptr, ev_segs = nullptr;
    cudaStream_t peer_stream[kShufMaxRanks] = {};   // copy-engine prefetch: one stream per peer
    cudaEvent_t peer_ev[kShufMaxRanks] = {};
    double last_build_ms = -1.0;            // measured build phase of the previous run: sizes the probe-side prefetch
    bool timed = false;
    int64_t stats[12] = {};
    int log2_parts = 0;
};

namespace {

constexpr int kHotSlots = 512;      // device set size (a power of two)
constexpr int kHotSamples = 4096;   // probe keys sampled per rank
constexpr int kHotCand = 128;       // (key, count) candidates each rank contributes; also the most keys kept

// Finds the keys that carry an outsized share of the probe rows: every rank samples its probe keys,
// counts them on the host, contributes its most frequent ones; the lists are all-gathered and merged
// the same way on every rank (keys whose share of all samples is at least 1/1000, at most kHotCand
// of them), and the set is uploaded to js->d_hot.  *nhot = 0: nothing stands out.
int32_t find_hot_keys(igl_join_shuffle* js, const igl_dev_col& key, int64_t nrows, cudaStream_t st, int32_t* nhot) {
    igl_ctx* ctx = js->ctx;
    Comm* comm = (Comm*)ctx->comm;
    const int W = comm->nranks;
    *nhot = 0;
    const int ns = (int)std::min<int64_t>(kHotSamples, nrows);
    unsigned long long* d_cand = js->d_samples + kHotSamples;
    unsigned long long* d_all = d_cand + 2 * kHotCand;
    unsigned long long* h_cand = js->h_samples + kHotSamples;
    unsigned long long* h_all = h_cand + 2 * kHotCand;
    if (ns > 0) {
        IGL_TRY(sample_keys_async(ctx, &key, nrows, ns, js->d_samples, st));
        IGL_CUDA_OK(cudaMemcpyAsync(js->h_samples, js->d_samples, sizeof(unsigned long long) * ns, cudaMemcpyDeviceToHost, st));
        IGL_CUDA_OK(cudaStreamSynchronize(st));
    }
    std::sort(js->h_samples, js->h_samples + ns);
    std::vector<std::pair<unsigned long long, unsigned long long>> runs;   // (count, key)
    for (int i = 0; i < ns;) {
        int j = i;
        while (j < ns && js->h_samples[j] == js->h_samples[i]) ++j;
        if (j - i >= 4) runs.push_back({(unsigned long long)(j - i), js->h_samples[i]});
        i = j;
    }
    std::sort(runs.begin(), runs.end(), [](const auto& a, const auto& b) { return a.first != b.first ? a.first > b.first : a.second < b.second; });
    for (int i = 0; i < kHotCand; ++i) {
        h_cand[2 * i] = i < (int)runs.size() ? runs[i].second : 0;
        h_cand[2 * i + 1] = i < (int)runs.size() ? runs[i].first : 0;   // count 0 = no candidate
    }
    IGL_CUDA_OK(cudaMemcpyAsync(d_cand, h_cand, sizeof(unsigned long long) * 2 * kHotCand, cudaMemcpyHostToDevice, st));
    IGL_NCCL_OK(g_nccl.AllGather(d_cand, d_all, sizeof(unsigned long long) * 2 * kHotCand, kNcclUint8, comm->comm, st));
    IGL_CUDA_OK(cudaMemcpyAsync(h_all, d_all, sizeof(unsigned long long) * 2 * kHotCand * W, cudaMemcpyDeviceToHost, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    std::map<unsigned long long, unsigned long long> merged;
    for (int i = 0; i < W * kHotCand; ++i)
        if (h_all[2 * i + 1]) merged[h_all[2 * i]] += h_all[2 * i + 1];
    const unsigned long long floor_count = std::max<unsigned long long>(8, (unsigned long long)W * kHotSamples / 1000);
    std::vector<std::pair<unsigned long long, unsigned long long>> hotv;
    for (auto& kv : merged)
        if (kv.second >= floor_count && kv.first != 0xFFFFFFFFFFFFFFFFull) hotv.push_back({kv.second, kv.first});
    std::sort(hotv.begin(), hotv.end(), [](const auto& a, const auto& b) { return a.first != b.first ? a.first > b.first : a.second < b.second; });
    if ((int)hotv.size() > kHotCand) hotv.resize(kHotCand);
    if (hotv.empty()) return IGL_OK;
    std::vector<unsigned long long> table(kHotSlots, 0xFFFFFFFFFFFFFFFFull);
    for (auto& hk : hotv) {
        uint32_t sl = (uint32_t)(mix64(hk.second) >> 24) & (kHotSlots - 1);
        while (table[sl] != 0xFFFFFFFFFFFFFFFFull) sl = (sl + 1) & (kHotSlots - 1);
        table[sl] = hk.second;
    }
    memcpy(js->h_samples, table.data(), sizeof(unsigned long long) * kHotSlots);   // pinned staging for the upload
    IGL_CUDA_OK(cudaMemcpyAsync(js->d_hot, js->h_samples, sizeof(unsigned long long) * kHotSlots, cudaMemcpyHostToDevice, st));
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    *nhot = (int32_t)hotv.size();
    return IGL_OK;
}

int floor_log2(int v) { int l = 0; while ((2 << l) <= v) ++l; return l; }

int32_t ensure_cols(igl_ctx* ctx, std::vector<void*>& cols, int64_t& cap_rows, const ShuffleSide& S, int64_t need_rows) {
    if (need_rows <= cap_rows && !cols.empty()) return IGL_OK;
    for (void* p : cols) pool_free(ctx, p);
    cols.assign(S.ncols, nullptr);
    const int64_t rows = need_rows + need_rows / 16 + 1024;
    for (int c = 0; c < S.ncols; ++c) IGL_TRY(pool_alloc(ctx, ((size_t)rows << S.shift[c]) + 256, &cols[c]));
    cap_rows = rows;
    return IGL_OK;
}

int32_t launch_seg_copy(igl_ctx* ctx, const ShuffleSide& S, const uint8_t* const src[][16], int nsrc, const CopySeg* d_segs, int nsegs,
                        uint32_t nblocks, void* const* dst, cudaStream_t st) {
    if (nsegs == 0 || nblocks == 0) return IGL_OK;
    CopyParams P{};
    for (int s = 0; s < nsrc; ++s)
        for (int c = 0; c < S.ncols; ++c) P.src[s][c] = src[s][c];
    for (int c = 0; c < S.ncols; ++c) { P.dst[c] = (uint8_t*)dst[c]; P.shift[c] = S.shift[c]; }
    P.ncols = S.ncols; P.nsegs = nsegs; P.segs = d_segs; P.nblocks = nblocks;
    const int grid = (int)std::min<uint32_t>(nblocks, (uint32_t)ctx->sm_count * 8);
    seg_copy_kernel<<<grid, kCopyThreads, 0, st>>>(P);
    IGL_CUDA_OK(cudaGetLastError());
    ctx->launches++;
    return IGL_OK;
}

}  // namespace

extern "C" {

int32_t igl_comm_unique_id(uint8_t id_out[128]) {
    if (!id_out) return fail(IGL_ERR_INVALID_ARG, "igl_comm_unique_id: NULL argument");
    IGL_TRY(nccl_load());
    NcclUniqueId id;
    IGL_NCCL_OK(g_nccl.GetUniqueId(&id));
    memcpy(id_out, id.internal, 128);
    return IGL_OK;
}

int32_t igl_comm_init(igl_ctx* ctx, int32_t nranks, int32_t rank, const uint8_t id[128]) {
    if (!ctx || !id || nranks < 1 || nranks > kShufMaxRanks || rank < 0 || rank >= nranks)
        return fail(IGL_ERR_INVALID_ARG, "igl_comm_init: bad arguments (1 <= nranks <= %d)", kShufMaxRanks);
    if (ctx->comm) return fail(IGL_ERR_INVALID_ARG, "igl_comm_init: this context already has a communicator");
    IGL_TRY(nccl_load());
    DeviceGuard g(ctx->device);
    NcclUniqueId uid;
    memcpy(uid.internal, id, 128);
    auto c = std::make_unique<Comm>();
    c->nranks = nranks; c->rank = rank;
    IGL_NCCL_OK(g_nccl.CommInitRank(&c->comm, nranks, uid, rank));
    ctx->comm = c.release();
    return IGL_OK;
}

int32_t igl_comm_rank(const igl_ctx* ctx, int32_t* rank, int32_t* nranks) {
    if (!ctx || !ctx->comm) return fail(IGL_ERR_INVALID_ARG, "igl_comm_rank: no communicator (igl_comm_init)");
    const Comm* c = (const Comm*)ctx->comm;
    if (rank) *rank = c->rank;
    if (nranks) *nranks = c->nranks;
    return IGL_OK;
}

int32_t igl_join_shuffle_create(igl_ctx* ctx, int32_t join_type, const int32_t* left_types, const uint8_t* left_nullable,
                                int32_t n_left, const int32_t* right_types, const uint8_t* right_nullable, int32_t n_right,
                                const igl_expr* left_key, const igl_expr* right_key, int32_t mode, igl_join_shuffle** out) {
    if (!ctx || !out || !left_types || !right_types || !left_key || !right_key)
        return fail(IGL_ERR_INVALID_ARG, "igl_join_shuffle_create: bad arguments");
    *out = nullptr;
    if (!ctx->comm) return fail(IGL_ERR_INVALID_ARG, "igl_join_shuffle_create: call igl_comm_init on this context first");
    if (mode != IGL_EXCHANGE_P2P && mode != IGL_EXCHANGE_NCCL) return fail(IGL_ERR_INVALID_ARG, "unknown exchange mode %d", mode);
    if (n_left < 1 || n_right < 1 || n_left > kShufMaxCols || n_right > kShufMaxCols)
        return fail(IGL_ERR_UNSUPPORTED, "1 to %d columns per join side", kShufMaxCols);
    for (int c = 0; c < n_left; ++c) if (left_nullable && left_nullable[c]) return fail(IGL_ERR_UNSUPPORTED, "nullable columns are not repartitioned across GPUs (left column %d)", c);
    for (int c = 0; c < n_right; ++c) if (right_nullable && right_nullable[c]) return fail(IGL_ERR_UNSUPPORTED, "nullable columns are not repartitioned across GPUs (right column %d)", c);
    auto js = std::make_unique<igl_join_shuffle>();
    js->ctx = ctx; js->mode = mode; js->join_type = join_type;
    js->left_join = join_type == IGL_JOIN_LEFT || join_type == IGL_JOIN_FULL;
    IGL_TRY(igl_join_create(ctx, join_type, left_types, left_nullable, n_left, right_types, right_nullable, n_right, left_key, right_key, &js->join));
    auto fill = [](ShuffleSide& S, const int32_t* t, int n, int key) {
        S.types.assign(t, t + n); S.ncols = n; S.key = key;
        S.shift.resize(n);
        for (int c = 0; c < n; ++c) S.shift[c] = t[c] == IGL_T_I32 ? 2 : 3;
    };
    fill(js->side[0], right_types, n_right, right_key->nodes[0].col);
    fill(js->side[1], left_types, n_left, left_key->nodes[0].col);
    DeviceGuard g(ctx->device);
    const int W = ((Comm*)ctx->comm)->nranks;
    cudaError_t e = cudaHostAlloc((void**)&js->h_send, sizeof(ShuffleMsg), cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&js->h_recv, sizeof(ShuffleMsg) * W, cudaHostAllocDefault);
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&js->h_plan, sizeof(unsigned long long) * 2 * kMsgBins, cudaHostAllocDefault);
    if (e != cudaSuccess) { igl_join_shuffle_destroy(js.release()); return fail(IGL_ERR_OOM, "pinned allocation failed: %s", cudaGetErrorString(e)); }
    memset(js->h_send, 0, sizeof(ShuffleMsg));
    void* p = nullptr;
    int32_t rc = pool_alloc(ctx, sizeof(ShuffleMsg), &p); js->d_send = (ShuffleMsg*)p;
    if (rc >= 0) { rc = pool_alloc(ctx, sizeof(ShuffleMsg) * W, &p); js->d_recv = (ShuffleMsg*)p; }
    if (rc >= 0) { rc = pool_alloc(ctx, sizeof(unsigned long long) * 4 * kMsgBins, &p); js->d_plan = (unsigned long long*)p; }
    if (rc >= 0) { rc = pool_alloc(ctx, 256, &p); js->d_flag = (int*)p; }
    const size_t hot_words = (size_t)kHotSamples + 2 * kHotCand * (1 + W);
    if (rc >= 0) { rc = pool_alloc(ctx, sizeof(unsigned long long) * kHotSlots, &p); js->d_hot = (unsigned long long*)p; }
    if (rc >= 0) { rc = pool_alloc(ctx, sizeof(unsigned long long) * hot_words, &p); js->d_samples = (unsigned long long*)p; }
    if (rc >= 0 && cudaHostAlloc((void**)&js->h_samples, sizeof(unsigned long long) * hot_words, cudaHostAllocDefault) != cudaSuccess) rc = fail(IGL_ERR_OOM, "pinned allocation failed");
    if (rc >= 0 && cudaMemset(js->d_flag, 0, 256) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "cudaMemset failed");
    for (auto& ev : js->ev) if (rc >= 0 && cudaEventCreate(&ev) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "cudaEventCreate failed");
    for (auto& ev : js->ev_l) if (rc >= 0 && cudaEventCreate(&ev) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "cudaEventCreate failed");
    if (rc >= 0 && (cudaEventCreate(&js->ev_plan) != cudaSuccess || cudaEventCreate(&js->ev_wait) != cudaSuccess)) rc = fail(IGL_ERR_CUDA, "cudaEventCreate failed");
    if (rc >= 0 && cudaStreamCreateWithFlags(&js->stream2, cudaStreamNonBlocking) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "cudaStreamCreate failed");
    if (rc >= 0) {   // the prefetch stream and the build-chain stream: highest priority, their CTAs go first when an SM has room
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        if (cudaStreamCreateWithPriority(&js->stream3, cudaStreamNonBlocking, hi) != cudaSuccess ||
            cudaStreamCreateWithPriority(&js->stream_hi, cudaStreamNonBlocking, hi) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "cudaStreamCreate failed");
    }
    if (rc >= 0 && cudaEventCreateWithFlags(&js->ev_copy, cudaEventDisableTiming) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "cudaEventCreate failed");
    if (rc >= 0 && cudaEventCreateWithFlags(&js->ev_segs, cudaEventDisableTiming) != cudaSuccess) rc = fail(IGL_ERR_CUDA, "cudaEventCreate failed");
    if (rc < 0) { igl_join_shuffle_destroy(js.release()); return rc; }
    *out = js.release();
    return IGL_OK;
}

int32_t igl_join_shuffle_run(igl_join_shuffle* js, const igl_dev_col* right, int64_t n_right, const igl_dev_col* left,
                             int64_t n_left, int64_t* out_rows, void* stream) {
    if (!js || !right || !left || n_right < 0 || n_left < 0) return fail(IGL_ERR_INVALID_ARG, "igl_join_shuffle_run: bad arguments");
    igl_ctx* ctx = js->ctx;
    Comm* comm = (Comm*)ctx->comm;
    if (!comm) return fail(IGL_ERR_INVALID_ARG, "igl_join_shuffle_run: the context has no communicator");
    DeviceGuard g(ctx->device);
    cudaStream_t st = pick_stream(ctx, stream);
    const int W = comm->nranks, me = comm->rank;
    const igl_dev_col* in[2] = {right, left};
    const int64_t n_in[2] = {n_right, n_left};
    for (int sd = 0; sd < 2; ++sd) {
        const ShuffleSide& S = js->side[sd];
        for (int c = 0; c < S.ncols; ++c) {
            if (in[sd][c].type != S.types[c]) return fail(IGL_ERR_INVALID_ARG, "%s column %d has type %d, the join expects %d", sd ? "left" : "right", c, in[sd][c].type, S.types[c]);
            if (in[sd][c].validity) return fail(IGL_ERR_UNSUPPORTED, "nullable columns are not repartitioned across GPUs (%s column %d)", sd ? "left" : "right", c);
        }
        if (!radix_partition_supported(in[sd], S.ncols)) return fail(IGL_ERR_UNSUPPORTED, "%s side cannot be radix-partitioned (fixed-width columns, at most 46 bytes per row)", sd ? "left" : "right");
        if (n_in[sd] >= (1ll << 32)) return fail(IGL_ERR_UNSUPPORTED, "at most 2^32-1 rows per rank and side");
    }
    cudaEventRecord(js->ev[0], st);
    // ---- staging buffers: this rank's rows, bin by bin (grow-only; peers re-map them when `gen` changes)
    for (int sd = 0; sd < 2; ++sd) {
        ShuffleSide& S = js->side[sd];
        const int64_t need = n_in[sd] + (int64_t)kSegAlignRows * kMsgBins + 64;
        if (need > S.staging_rows || S.staging.empty()) {
            for (void* p : S.staging) S.retired.push_back(p);   // a slow peer may still be reading last join's runs
            S.staging.assign(S.ncols, nullptr);
            const int64_t rows = need + need / 16;
            for (int c = 0; c < S.ncols; ++c) IGL_TRY(pool_alloc(ctx, ((size_t)rows << S.shift[c]) + 256, &S.staging[c]));
            S.staging_rows = rows;
            S.gen++;
            if (js->mode == IGL_EXCHANGE_P2P && W > 1)
                for (int c = 0; c < S.ncols; ++c) IGL_TRY(igl_ipc_export(ctx, S.staging[c], js->h_send->handles[sd][c]));
        }
        js->h_send->n_rows[sd] = n_in[sd];
        js->h_send->gen[sd] = S.gen;
    }
    // ---- 1. histograms, 2. the all-gather.  nd[side] destinations (the build side gets one more,
    // "every rank reads this", when hot keys are in play), lp_h partitions in the histogram; the
    // join's own partition count is chosen once the totals are known and the counts are folded.
    int nd[2] = {W, W};
    HotKeys hot[2];
    int lp_h = std::min(10, floor_log2(kMsgBins / W));
    auto exchange_counts = [&]() -> int32_t {
        IGL_CUDA_OK(cudaMemcpyAsync(js->d_send, js->h_send, offsetof(ShuffleMsg, hist), cudaMemcpyHostToDevice, st));
        for (int sd = 0; sd < 2; ++sd) {
            IGL_CUDA_OK(cudaMemsetAsync(js->d_send->hist[sd], 0, sizeof(unsigned long long) * kMsgBins, st));
            IGL_TRY(radix_histogram_async(ctx, &in[sd][js->side[sd].key], n_in[sd], nd[sd], W, lp_h, hot[sd], js->d_send->hist[sd], st));
        }
        IGL_NCCL_OK(g_nccl.AllGather(js->d_send, js->d_recv, sizeof(ShuffleMsg), kNcclUint8, comm->comm, st));
        IGL_CUDA_OK(cudaMemcpyAsync(js->h_recv, js->d_recv, sizeof(ShuffleMsg) * W, cudaMemcpyDeviceToHost, st));
        return IGL_OK;
    };
    IGL_TRY(exchange_counts());
    cudaEventRecord(js->ev[1], st);
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    // every rank has entered this join, so nobody reads last join's staging any more
    for (int sd = 0; sd < 2; ++sd) { for (void* p : js->side[sd].retired) pool_free(ctx, p); js->side[sd].retired.clear(); }
    int64_t total_rows[2] = {0, 0};
    for (int s = 0; s < W; ++s) for (int sd = 0; sd < 2; ++sd) total_rows[sd] += js->h_recv[s].n_rows[sd];
    // ---- skew: if one GPU would receive far more probe rows than its share, find the keys that
    // cause it (a sample of every rank's probe keys, merged on every rank alike), keep their probe
    // rows where they are and send their build rows to every rank; then count again.
    js->stats[10] = 0;
    {
        const char* he = getenv("IGL_SHUFFLE_HOTKEYS");
        const bool allowed = W > 1 && js->mode == IGL_EXCHANGE_P2P && !(he && he[0] == '0');   // the NCCL mode ships destination ranges only
        int64_t per_dest[kShufMaxRanks] = {}, most = 0;
        for (int s = 0; s < W; ++s)
            for (int b = 0; b < (W << lp_h); ++b) per_dest[b >> lp_h] += (int64_t)js->h_recv[s].hist[1][b];
        for (int d = 0; d < W; ++d) most = std::max(most, per_dest[d]);
        if (allowed && total_rows[1] >= (1 << 20) && (double)most > 1.15 * (double)total_rows[1] / W && ((W + 1) << 1) <= kMsgBins) {
            int32_t nhot = 0;
            IGL_TRY(find_hot_keys(js, in[1][js->side[1].key], n_in[1], st, &nhot));
            if (nhot > 0) {
                nd[0] = W + 1;
                lp_h = std::min(10, floor_log2(kMsgBins / (W + 1)));
                hot[0] = HotKeys{js->d_hot, kHotSlots - 1, W};    // build rows of hot keys: the broadcast destination
                hot[1] = HotKeys{js->d_hot, kHotSlots - 1, me};   // probe rows of hot keys: stay here
                IGL_TRY(exchange_counts());
                IGL_CUDA_OK(cudaStreamSynchronize(st));
                js->stats[10] = nhot;
            }
        }
    }
    // ---- the plan (identical arithmetic on every rank)
    int lp = std::min(lp_h, (int)igl_join_partition_bits(ctx, std::max<int64_t>(1, total_rows[0] / W)));
    js->log2_parts = lp;
    const int parts = 1 << lp, fold = lp_h - lp;
    const int bins = W << lp;                         // probe side (and build side without hot keys)
    const int binsS[2] = {nd[0] << lp, nd[1] << lp};
    std::vector<int64_t> cnt[2], off[2];
    for (int sd = 0; sd < 2; ++sd) {
        cnt[sd].assign((size_t)W * binsS[sd], 0);
        off[sd].assign((size_t)W * binsS[sd], 0);
        for (int s = 0; s < W; ++s) {
            int64_t* c = &cnt[sd][(size_t)s * binsS[sd]];
            for (int b = 0; b < (nd[sd] << lp_h); ++b) {
                const int d = b >> lp_h, ph = b & ((1 << lp_h) - 1);
                c[(d << lp) | (ph >> fold)] += (int64_t)js->h_recv[s].hist[sd][b];
            }
        }
    }
    const int nsegS[2] = {W * parts * (nd[0] > W ? 2 : 1), W * parts};
    std::vector<int32_t> seg_src[2];
    std::vector<int64_t> seg_row0[2], seg_rows[2];
    int64_t recv_total[2] = {0, 0};
    for (int sd = 0; sd < 2; ++sd) {
        seg_src[sd].assign(nsegS[sd], 0); seg_row0[sd].assign(nsegS[sd], 0); seg_rows[sd].assign(nsegS[sd], 0);
        IGL_TRY(igl_shuffle_plan_ex(W, me, lp, nd[sd] - W, cnt[sd].data(), off[sd].data(), seg_src[sd].data(), seg_row0[sd].data(),
                                    seg_rows[sd].data(), &recv_total[sd]));
    }
    for (int sd = 0; sd < 2; ++sd)
        for (int b = 0; b < binsS[sd]; ++b) js->h_plan[(size_t)sd * kMsgBins + b] = (unsigned long long)off[sd][(size_t)me * binsS[sd] + b];
    IGL_CUDA_OK(cudaMemcpyAsync(js->d_plan, js->h_plan, sizeof(unsigned long long) * 2 * kMsgBins, cudaMemcpyHostToDevice, st));
    IGL_CUDA_OK(cudaMemsetAsync(js->d_plan + 2 * kMsgBins, 0, sizeof(unsigned long long) * 2 * kMsgBins, st));
    // ---- 3. local radix partition of both sides into the staging buffers.  p2p mode: the probe side
    // (the long one) is partitioned on a second stream, so it overlaps the build side's barrier,
    // pull and table build; the probe kernel waits for it (and for every peer's) below.
    const bool overlap = js->mode == IGL_EXCHANGE_P2P && !(getenv("IGL_SHUFFLE_OVERLAP") && getenv("IGL_SHUFFLE_OVERLAP")[0] == '0');
    cudaStream_t st_l = overlap ? js->stream2 : st;
    // The probe side's partition kernel, launched as one persistent grid, holds every SM until it is done, and
    // the build side's pull and table build wait behind it although they need other resources (NVLink, L2
    // atomics).  Yielding form: one CTA per tile for the probe side's partitioning, and the build side's chain
    // on a stream of the highest priority — its CTAs get onto the SMs between tiles.  Measured (C3): the two then
    // do run side by side and finish no earlier — 16.43 against 16.48 ms on 2 GPUs, 5.78 against 5.59 ms on 8 —
    // so it stays off unless IGL_SHUFFLE_YIELD=1.
    const bool yield = overlap && W > 1 && getenv("IGL_SHUFFLE_YIELD") && getenv("IGL_SHUFFLE_YIELD")[0] == '1';
    cudaStream_t st_b = yield ? js->stream_hi : st;
    if (overlap) {
        cudaEventRecord(js->ev_plan, st);
        IGL_CUDA_OK(cudaStreamWaitEvent(st_l, js->ev_plan, 0));
        cudaEventRecord(js->ev_l[0], st_l);
    }
    for (int sd = 0; sd < 2; ++sd)
        IGL_TRY(radix_scatter_bins_async(ctx, in[sd], js->side[sd].ncols, js->side[sd].key, n_in[sd], nd[sd], W, lp, hot[sd],
                                         js->side[sd].staging.data(), js->d_plan + (size_t)sd * kMsgBins,
                                         js->d_plan + (size_t)(2 + sd) * kMsgBins, sd == 1 ? st_l : st, sd == 1 && yield));
    if (overlap) cudaEventRecord(js->ev_l[1], st_l);
    cudaEventRecord(js->ev[2], st);
    // ---- where the runs are read from
    const uint8_t* src_cols[2][kSegMaxSrc][16] = {};
    int64_t src_base[2][kSegMaxSrc] = {};   // row of src_cols[..][s] inside source s's staging buffer (0 when it IS the staging buffer)
    int64_t remote_bytes = 0, pref_bytes = 0;
    int pref_parts = 0;                     // probe-side partitions [0, pref_parts) of every peer are prefetched into local memory
    bool prefetching = false;
    auto range = [&](int sd, int src, int dst, int np, int64_t& first, int64_t& len) {   // rows of src's staging that hold dst's partitions [0, np)
        const int64_t* o = &off[sd][(size_t)src * binsS[sd]];
        const int64_t* c = &cnt[sd][(size_t)src * binsS[sd]];
        first = o[(size_t)dst << lp];
        const int last = (dst << lp) + np - 1;
        len = o[last] + align_rows(c[last]) - first;
    };
    int64_t pf_first[kShufMaxRanks] = {}, pf_len[kShufMaxRanks] = {}, pf_base[kShufMaxRanks] = {};
    int64_t rbase[2][kShufMaxRanks + 1] = {};
    if (js->mode == IGL_EXCHANGE_P2P) {
        for (int sd = 0; sd < 2; ++sd) {
            ShuffleSide& S = js->side[sd];
            for (int s = 0; s < W; ++s) {
                if (s == me) { for (int c = 0; c < S.ncols; ++c) src_cols[sd][s][c] = (const uint8_t*)S.staging[c]; continue; }
                ShuffleSide::Peer& P = S.peer[s];
                const uint64_t pg = js->h_recv[s].gen[sd];
                if (!P.open || P.gen != pg) {   // the peer (re)allocated its staging: map the new blocks
                    if (P.open) for (void* q : P.ptrs) igl_ipc_close(ctx, q);
                    P.ptrs.assign(S.ncols, nullptr);
                    P.open = false;
                    for (int c = 0; c < S.ncols; ++c) IGL_TRY(igl_ipc_open(ctx, js->h_recv[s].handles[sd][c], &P.ptrs[c]));
                    P.open = true; P.gen = pg;
                }
                for (int c = 0; c < S.ncols; ++c) src_cols[sd][s][c] = (const uint8_t*)P.ptrs[c];
            }
        }
        // ---- NVLink is idle while the table is built: a copy kernel on its own stream uses that time to
        // bring the probe rows of the first `pref_parts` partitions over from every peer (each peer's range
        // is one contiguous run per column); the probe kernel then finds those runs in local memory and
        // fetches only the rest across the link itself.  Only a probe that the link would hold back is
        // worth it, and only as much as fits under the build (sized from the previous run's build time).
        // Measured at C3 on 8 GPUs (5.6-5.7 ms per join without), every variant slower: one copy-engine
        // stream moved the 0.71 GB at only ~245 GB/s and the probe waited for it (8.44 ms); one copy-engine
        // stream per peer 6.97 ms (8.87 ms at 60 % of the partitions); a copy kernel over all SMs took the
        // build kernel's SMs and memory system (build 1.1 -> 3.4 ms; 7.80 ms); 24 CTAs working as copy engines
        // (dma_copy_kernel, 20 % of the partitions) 6.33 ms: the probe kernel gained 0.19 ms (2.74 -> 2.55: its
        // time does not fall in proportion to the rows made local), the build lost 0.2-0.9 ms.  Off unless
        // IGL_SHUFFLE_PREFETCH names a mode: "ce" = one copy-engine stream per peer, "sm" = copy kernel,
        // "dma" = dma_copy_kernel; IGL_SHUFFLE_PREFETCH_FRAC sets the share of partitions.
        const char* pe = getenv("IGL_SHUFFLE_PREFETCH");
        const bool pref_ce = pe && pe[0] == 'c', pref_sm = pe && pe[0] == 's', pref_dma = pe && pe[0] == 'd';
        if (overlap && W > 1 && (pref_ce || pref_sm || (pref_dma && (W - 1) * js->side[1].ncols <= kDmaMaxRuns))) {
            ShuffleSide& S = js->side[1];
            const char* fe = getenv("IGL_SHUFFLE_PREFETCH_FRAC");
            double frac = fe ? atof(fe) : -1.0;
            int64_t remote_rows = 0;
            for (int s = 0; s < W; ++s)
                if (s != me) for (int p = 0; p < parts; ++p) remote_rows += cnt[1][(size_t)s * bins + ((size_t)me << lp) + p];
            if (frac < 0) {
                // a probe kernel streams about 3.3 TB/s of (input + output) bytes from local memory and
                // about 650 GB/s across NVLink
                const double remote_bytes_l = (double)remote_rows * S.row_bytes();
                const double local_ms = (double)recv_total[1] * (2.0 * S.row_bytes() + js->side[0].row_bytes()) / 3.3e9;
                const double link_ms = remote_bytes_l / 650e6;
                frac = 0.0;
                if (remote_rows > 0 && link_ms > 0.9 * local_ms) {
                    const double build_ms = js->last_build_ms > 0 ? js->last_build_ms : (double)recv_total[0] * 6.5e-8;
                    const double budget = 0.8 * build_ms * 600e6;
                    const double want = remote_bytes_l - 0.85 * local_ms * 650e6;
                    frac = std::max(0.0, std::min(1.0, std::min(budget, want) / remote_bytes_l));
                }
            }
            pref_parts = std::min(parts, (int)(frac * parts));
            if (pref_parts > 0 && remote_rows > 0) {
                int64_t rb = 0;
                for (int s = 0; s < W; ++s) {
                    if (s == me) continue;
                    range(1, s, me, pref_parts, pf_first[s], pf_len[s]);
                    pf_base[s] = rb; rb += pf_len[s];
                }
                IGL_TRY(ensure_cols(ctx, S.recv, S.recv_rows, S, rb + 64));
                for (int s = 0; s < W; ++s) {
                    if (s == me || pf_len[s] == 0) continue;
                    for (int c = 0; c < S.ncols; ++c) src_cols[1][W + s][c] = (const uint8_t*)S.recv[c] + ((size_t)pf_base[s] << S.shift[c]);
                    src_base[1][W + s] = pf_first[s];
                    pref_bytes += pf_len[s] * S.row_bytes();
                }
                prefetching = true;
            } else pref_parts = 0;
        }
    } else {
        // NCCL mode: every source's range for this rank lands in a local receive slab, source by source
        for (int sd = 0; sd < 2; ++sd) {
            ShuffleSide& S = js->side[sd];
            for (int s = 0; s < W; ++s) { int64_t f, l; range(sd, s, me, parts, f, l); rbase[sd][s + 1] = rbase[sd][s] + l; src_base[sd][s] = f; }
            IGL_TRY(ensure_cols(ctx, S.recv, S.recv_rows, S, rbase[sd][W] + 64));
            for (int s = 0; s < W; ++s)
                for (int c = 0; c < S.ncols; ++c) src_cols[sd][s][c] = (const uint8_t*)S.recv[c] + ((size_t)rbase[sd][s] << S.shift[c]);
        }
    }
    // ---- run lists (one upload): what the build pull, the probe and the prefetch walk
    const int nseg_max = 2 * W * parts;   // per side (the build side lists broadcast runs too when hot keys are in play)
    const size_t seg_bytes = (size_t)nseg_max * (sizeof(CopySeg) * 2 + sizeof(ProbeSeg)) + sizeof(CopySeg) * kShufMaxRanks + 256;
    if (seg_bytes > js->segs_bytes) {
        if (js->d_segs) pool_free(ctx, js->d_segs);
        if (js->h_segs) cudaFreeHost(js->h_segs);
        js->d_segs = js->h_segs = nullptr; js->segs_bytes = 0;
        IGL_TRY(pool_alloc(ctx, seg_bytes, &js->d_segs));
        IGL_CUDA_OK(cudaHostAlloc(&js->h_segs, seg_bytes, cudaHostAllocDefault));
        js->segs_bytes = seg_bytes;
    }
    CopySeg* h_copy[2] = {(CopySeg*)js->h_segs, (CopySeg*)js->h_segs + nseg_max};
    ProbeSeg* h_probe = (ProbeSeg*)((CopySeg*)js->h_segs + 2 * nseg_max);
    CopySeg* h_pref = (CopySeg*)(h_probe + nseg_max);
    int ncopy[2] = {0, 0}, nprobe = 0, npref = 0;
    uint32_t copy_blocks[2] = {0, 0}, pref_blocks = 0;
    long long probe_chunks = 0;
    for (int sd = 0; sd < 2; ++sd) {
        int64_t dst = 0;
        for (int i = 0; i < nsegS[sd]; ++i) {
            const int64_t rows = seg_rows[sd][i];
            if (rows == 0) continue;
            int s = seg_src[sd][i];
            if (s != me) remote_bytes += rows * js->side[sd].row_bytes();
            if (sd == 1 && prefetching && s != me && i / W < pref_parts) s += W;   // this run has a local copy
            const int64_t row0 = seg_row0[sd][i] - src_base[sd][s];
            CopySeg& cs = h_copy[sd][ncopy[sd]++];
            cs.src_row0 = (uint64_t)row0; cs.dst_row0 = (uint64_t)dst; cs.rows = (uint32_t)rows; cs.blk0 = copy_blocks[sd]; cs.src = (uint32_t)s; cs.pad = 0;
            copy_blocks[sd] += (uint32_t)((rows + kCopyBlockRows - 1) / kCopyBlockRows);
            dst += rows;
            if (sd == 1) {
                ProbeSeg& ps = h_probe[nprobe++];
                ps.chunk0 = (uint32_t)probe_chunks; ps.rows = (uint32_t)rows;
                ps.row0_src = (uint64_t)row0 | ((uint64_t)s << 56);
                ps.out0 = cs.dst_row0; ps.pad = 0;
                probe_chunks += (rows + kSegChunkRows - 1) / kSegChunkRows;
            }
        }
    }
    if (prefetching)
        for (int s = 0; s < W; ++s) {
            if (s == me || pf_len[s] == 0) continue;
            CopySeg& cs = h_pref[npref++];
            cs.src_row0 = (uint64_t)pf_first[s]; cs.dst_row0 = (uint64_t)pf_base[s]; cs.rows = (uint32_t)pf_len[s]; cs.blk0 = pref_blocks; cs.src = (uint32_t)s; cs.pad = 0;
            pref_blocks += (uint32_t)((pf_len[s] + kCopyBlockRows - 1) / kCopyBlockRows);
        }
    IGL_CUDA_OK(cudaMemcpyAsync(js->d_segs, js->h_segs, seg_bytes - 256, cudaMemcpyHostToDevice, st));
    const CopySeg* d_copy[2] = {(const CopySeg*)js->d_segs, (const CopySeg*)js->d_segs + nseg_max};
    const ProbeSeg* d_probe = (const ProbeSeg*)((const CopySeg*)js->d_segs + 2 * nseg_max);
    const CopySeg* d_pref = (const CopySeg*)(d_probe + nseg_max);
    cudaEventRecord(js->ev_segs, st);
    // ---- 4. every rank's staging is complete (p2p), or the rows travel now (nccl)
    if (js->mode == IGL_EXCHANGE_P2P) {
        // the build side's barrier (on this stream) first, then the probe side's (on its own stream when
        // overlapped) — the same order on every rank
        if (yield) IGL_CUDA_OK(cudaStreamWaitEvent(st_b, js->ev_segs, 0));
        IGL_NCCL_OK(g_nccl.AllReduce(js->d_flag, js->d_flag, 1, kNcclInt32, kNcclSum, comm->comm, st_b));
        if (overlap) {
            IGL_NCCL_OK(g_nccl.AllReduce(js->d_flag + 8, js->d_flag + 8, 1, kNcclInt32, kNcclSum, comm->comm, st_l));
            cudaEventRecord(js->ev_l[2], st_l);
        }
        if (prefetching) {
            const char* pm = getenv("IGL_SHUFFLE_PREFETCH");
            if (pm && pm[0] == 'c') {   // copy engines, one stream per peer so the engines work side by side
                ShuffleSide& S = js->side[1];
                for (int s = 0; s < W; ++s) {
                    if (s == me || pf_len[s] == 0) continue;
                    if (!js->peer_stream[s]) {
                        IGL_CUDA_OK(cudaStreamCreateWithFlags(&js->peer_stream[s], cudaStreamNonBlocking));
                        IGL_CUDA_OK(cudaEventCreateWithFlags(&js->peer_ev[s], cudaEventDisableTiming));
                    }
                    IGL_CUDA_OK(cudaStreamWaitEvent(js->peer_stream[s], js->ev_l[2], 0));
                    for (int c = 0; c < S.ncols; ++c)
                        IGL_CUDA_OK(cudaMemcpyAsync((uint8_t*)S.recv[c] + ((size_t)pf_base[s] << S.shift[c]),
                                                    src_cols[1][s][c] + ((size_t)pf_first[s] << S.shift[c]), (size_t)pf_len[s] << S.shift[c],
                                                    cudaMemcpyDefault, js->peer_stream[s]));
                    IGL_CUDA_OK(cudaEventRecord(js->peer_ev[s], js->peer_stream[s]));
                    IGL_CUDA_OK(cudaStreamWaitEvent(js->stream3, js->peer_ev[s], 0));
                }
            } else if (pm && pm[0] == 'd') {   // a few CTAs as copy engines (dma_copy_kernel)
                ShuffleSide& S = js->side[1];
                DmaParams D{};
                uint32_t piece = 0;
                for (int s = 0; s < W; ++s) {
                    if (s == me || pf_len[s] == 0) continue;
                    for (int c = 0; c < S.ncols; ++c) {
                        D.src[D.nruns] = src_cols[1][s][c] + ((size_t)pf_first[s] << S.shift[c]);
                        D.dst[D.nruns] = (uint8_t*)S.recv[c] + ((size_t)pf_base[s] << S.shift[c]);
                        D.bytes[D.nruns] = (uint64_t)pf_len[s] << S.shift[c];
                        D.piece0[D.nruns] = piece;
                        piece += (uint32_t)((D.bytes[D.nruns] + kDmaPiece - 1) / kDmaPiece);
                        ++D.nruns;
                    }
                }
                D.piece0[D.nruns] = piece;
                const char* ce = getenv("IGL_SHUFFLE_PREFETCH_CTAS");
                const int ctas = std::max(1, std::min(ctx->sm_count, ce ? atoi(ce) : 24));
                static bool attr_set = false;
                if (!attr_set) {
                    IGL_CUDA_OK(cudaFuncSetAttribute((const void*)dma_copy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kDmaSlots * kDmaPiece)));
                    attr_set = true;
                }
                IGL_CUDA_OK(cudaStreamWaitEvent(js->stream3, js->ev_l[2], 0));
                if (piece) {
                    dma_copy_kernel<<<ctas, 32, kDmaSlots * kDmaPiece, js->stream3>>>(D);
                    IGL_CUDA_OK(cudaGetLastError());
                    ctx->launches++;
                }
            } else {
                IGL_CUDA_OK(cudaStreamWaitEvent(js->stream3, js->ev_l[2], 0));
                IGL_CUDA_OK(cudaStreamWaitEvent(js->stream3, js->ev_segs, 0));
                IGL_TRY(launch_seg_copy(ctx, js->side[1], src_cols[1], W, d_pref, npref, pref_blocks, js->side[1].recv.data(), js->stream3));
            }
            IGL_CUDA_OK(cudaEventRecord(js->ev_copy, js->stream3));
        }
    } else {
        // grouped send/recv of each destination's contiguous staging range (padding included)
        for (int sd = 0; sd < 2; ++sd) {
            ShuffleSide& S = js->side[sd];
            IGL_NCCL_OK(g_nccl.GroupStart());
            for (int c = 0; c < S.ncols; ++c)
                for (int p = 0; p < W; ++p) {
                    int64_t f, l;
                    range(sd, me, p, parts, f, l);
                    if (l) IGL_NCCL_OK(g_nccl.Send((const uint8_t*)S.staging[c] + ((size_t)f << S.shift[c]), (size_t)l << S.shift[c], kNcclUint8, p, comm->comm, st));
                    range(sd, p, me, parts, f, l);
                    if (l) IGL_NCCL_OK(g_nccl.Recv((uint8_t*)S.recv[c] + ((size_t)rbase[sd][p] << S.shift[c]), (size_t)l << S.shift[c], kNcclUint8, p, comm->comm, st));
                }
            IGL_NCCL_OK(g_nccl.GroupEnd());
        }
    }
    cudaEventRecord(js->ev[3], st_b);
    // ---- 5. build side: pull, build
    ShuffleSide& R = js->side[0];
    ShuffleSide& L = js->side[1];
    IGL_TRY(ensure_cols(ctx, R.slab, R.slab_rows, R, recv_total[0] + 64));
    IGL_TRY(launch_seg_copy(ctx, R, src_cols[0], W, d_copy[0], ncopy[0], copy_blocks[0], R.slab.data(), st_b));
    std::vector<igl_dev_col> rcols(R.ncols), lcols(L.ncols);
    for (int c = 0; c < R.ncols; ++c) rcols[c] = igl_dev_col{R.types[c], 0, R.slab[c], nullptr, nullptr};
    IGL_TRY(igl_join_set_prepartitioned(js->join, lp));   // by how many hash bits the slab is ordered
    IGL_TRY(igl_join_build(js->join, rcols.data(), recv_total[0], st_b));
    cudaEventRecord(js->ev[4], st_b);
    if (yield) IGL_CUDA_OK(cudaStreamWaitEvent(st, js->ev[4], 0));
    // ---- 6. probe side
    if (overlap) IGL_CUDA_OK(cudaStreamWaitEvent(st, js->ev_l[2], 0));
    if (prefetching) IGL_CUDA_OK(cudaStreamWaitEvent(st, js->ev_copy, 0));
    cudaEventRecord(js->ev_wait, st);
    const int nsrc = prefetching ? 2 * W : W;
    const int ncols_out = L.ncols + R.ncols;
    auto ensure_out = [&](int64_t rows) -> int32_t {
        if (rows <= js->out_cap && !js->out.empty()) return IGL_OK;
        for (void* p : js->out) pool_free(ctx, p);
        for (void* p : js->out_valid) pool_free(ctx, p);
        js->out.assign(ncols_out, nullptr); js->out_valid.assign(ncols_out, nullptr);
        const int64_t cap = rows + rows / 16 + 1024;
        for (int c = 0; c < ncols_out; ++c) {
            const int sh = c < L.ncols ? L.shift[c] : R.shift[c - L.ncols];
            IGL_TRY(pool_alloc(ctx, ((size_t)cap << sh) + 256, &js->out[c]));
            if (js->left_join && c >= L.ncols) IGL_TRY(pool_alloc(ctx, (size_t)(cap + 7) / 8 + 256, &js->out_valid[c]));
        }
        js->out_cap = cap;
        return IGL_OK;
    };
    auto out_cols = [&](std::vector<igl_dev_col>& oc) {
        oc.resize(ncols_out);
        for (int c = 0; c < ncols_out; ++c)
            oc[c] = igl_dev_col{c < L.ncols ? L.types[c] : R.types[c - L.ncols], 0, js->out[c], (uint8_t*)js->out_valid[c], nullptr};
    };
    std::vector<igl_dev_col> oc;
    int64_t n_out = 0;
    const char* segenv = getenv("IGL_SHUFFLE_FUSED_PROBE");
    if (!js->left_join && !(segenv && segenv[0] == '0') && join_segments_eligible(js->join)) {
        SegSource SS{};
        for (int s = 0; s < nsrc; ++s) for (int c = 0; c < L.ncols; ++c) SS.cols[s][c] = src_cols[1][s][c] ? src_cols[1][s][c] : src_cols[1][me][c];
        SS.nsrc = nsrc; SS.dev_segs = d_probe; SS.nsegs = nprobe; SS.nchunks = probe_chunks; SS.total_rows = recv_total[1];
        IGL_TRY(ensure_out(std::max<int64_t>(recv_total[1], 1)));
        out_cols(oc);
        IGL_TRY(join_probe_segments(js->join, SS, oc.data(), js->out_cap, &n_out, st));
        js->stats[4] = 1;
    } else {
        IGL_TRY(ensure_cols(ctx, L.slab, L.slab_rows, L, recv_total[1] + 64));
        IGL_TRY(launch_seg_copy(ctx, L, src_cols[1], nsrc, d_copy[1], ncopy[1], copy_blocks[1], L.slab.data(), st));
        for (int c = 0; c < L.ncols; ++c) lcols[c] = igl_dev_col{L.types[c], 0, L.slab[c], nullptr, nullptr};
        int64_t need = recv_total[1];
        if (igl_join_has_duplicates(js->join)) IGL_TRY(igl_join_probe_count(js->join, lcols.data(), recv_total[1], &need, st));
        IGL_TRY(ensure_out(std::max<int64_t>(need, 1)));
        out_cols(oc);
        IGL_TRY(igl_join_probe(js->join, lcols.data(), recv_total[1], oc.data(), js->out_cap, &n_out, st));
        js->stats[4] = 0;
    }
    cudaEventRecord(js->ev[5], st);
    // every rank is done reading its peers' staging buffers: after this they may be overwritten,
    // freed or handed to another join without asking anybody
    if (js->mode == IGL_EXCHANGE_P2P && W > 1)
        IGL_NCCL_OK(g_nccl.AllReduce(js->d_flag, js->d_flag, 1, kNcclInt32, kNcclSum, comm->comm, st));
    cudaEventRecord(js->ev[6], st);
    IGL_CUDA_OK(cudaStreamSynchronize(st));
    js->timed = true;
    js->out_rows = n_out;
    js->stats[0] = recv_total[0]; js->stats[1] = recv_total[1]; js->stats[2] = remote_bytes; js->stats[3] = lp;
    js->stats[5] = n_out;
    {
        float bms = 0.f;
        if (cudaEventElapsedTime(&bms, js->ev[3], js->ev[4]) == cudaSuccess) js->last_build_ms = bms;
    }
    js->stats[2] = remote_bytes;
    js->stats[8] = pref_bytes; js->stats[9] = pref_parts;
    js->stats[6] = js->stats[7] = -1;   // probe-side partitioning on the second stream (us), and how long the probe waited for it (us)
    if (overlap) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, js->ev_l[0], js->ev_l[1]) == cudaSuccess) js->stats[6] = (int64_t)(ms * 1e3);
        if (cudaEventElapsedTime(&ms, js->ev[4], js->ev_wait) == cudaSuccess) js->stats[7] = (int64_t)(ms * 1e3);
    }
    if (out_rows) *out_rows = n_out;
    return IGL_OK;
}

int32_t igl_join_shuffle_output(igl_join_shuffle* js, igl_dev_col* out, int64_t* rows) {
    if (!js || !out) return fail(IGL_ERR_INVALID_ARG, "igl_join_shuffle_output: NULL argument");
    if (js->out.empty()) return fail(IGL_ERR_INVALID_ARG, "igl_join_shuffle_output: no join has run yet");
    const ShuffleSide& L = js->side[1];
    const ShuffleSide& R = js->side[0];
    for (int c = 0; c < L.ncols + R.ncols; ++c)
        out[c] = igl_dev_col{c < L.ncols ? L.types[c] : R.types[c - L.ncols], 0, js->out[c], (uint8_t*)js->out_valid[c], nullptr};
    if (rows) *rows = js->out_rows;
    return IGL_OK;
}

int32_t igl_join_shuffle_phase_ms(igl_join_shuffle* js, double out[8]) {
    if (!js || !out) return fail(IGL_ERR_INVALID_ARG, "igl_join_shuffle_phase_ms: NULL argument");
    for (int i = 0; i < 8; ++i) out[i] = -1.0;
    if (!js->timed) return IGL_OK;
    DeviceGuard g(js->ctx->device);
    float ms = 0.f;
    for (int i = 0; i < 6; ++i)
        if (cudaEventElapsedTime(&ms, js->ev[i], js->ev[i + 1]) == cudaSuccess) out[i] = ms;
    if (cudaEventElapsedTime(&ms, js->ev_wait, js->ev[5]) == cudaSuccess) out[4] = ms;   // probe, without the wait for the probe side's partitioning
    if (cudaEventElapsedTime(&ms, js->ev[0], js->ev[6]) == cudaSuccess) out[6] = ms;
    double jp[4];
    if (igl_join_phase_ms(js->join, jp) == IGL_OK) out[7] = jp[3];   // the probe kernel alone
    return IGL_OK;
}

int32_t igl_join_shuffle_stats(igl_join_shuffle* js, int64_t out[12]) {
    if (!js || !out) return fail(IGL_ERR_INVALID_ARG, "igl_join_shuffle_stats: NULL argument");
    memcpy(out, js->stats, sizeof js->stats);
    return IGL_OK;
}

void igl_join_shuffle_destroy(igl_join_shuffle* js) {
    if (!js) return;
    igl_ctx* ctx = js->ctx;
    DeviceGuard g(ctx->device);
    Comm* comm = (Comm*)ctx->comm;
    // not a collective: igl_join_shuffle_run returns only after every rank has finished reading
    // its peers' staging buffers
    (void)comm;
    for (auto& S : js->side) {
        for (auto& P : S.peer) if (P.open) for (void* q : P.ptrs) if (q) cudaIpcCloseMemHandle(q);
        for (void* p : S.staging) pool_free(ctx, p);
        for (void* p : S.retired) pool_free(ctx, p);
        for (void* p : S.slab) pool_free(ctx, p);
        for (void* p : S.recv) pool_free(ctx, p);
    }
    for (void* p : js->out) pool_free(ctx, p);
    for (void* p : js->out_valid) pool_free(ctx, p);
    pool_free(ctx, js->d_hot); pool_free(ctx, js->d_samples);
    if (js->h_samples) cudaFreeHost(js->h_samples);
    pool_free(ctx, js->d_send); pool_free(ctx, js->d_recv); pool_free(ctx, js->d_plan); pool_free(ctx, js->d_segs); pool_free(ctx, js->d_flag);
    if (js->h_send) cudaFreeHost(js->h_send);
    if (js->h_recv) cudaFreeHost(js->h_recv);
    if (js->h_plan) cudaFreeHost(js->h_plan);
    if (js->h_segs) cudaFreeHost(js->h_segs);
    for (auto& ev : js->ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : js->ev_l) if (ev) cudaEventDestroy(ev);
    if (js->ev_plan) cudaEventDestroy(js->ev_plan);
    if (js->ev_wait) cudaEventDestroy(js->ev_wait);
    if (js->stream2) cudaStreamDestroy(js->stream2);
    if (js->stream3) cudaStreamDestroy(js->stream3);
    if (js->stream_hi) cudaStreamDestroy(js->stream_hi);
    for (auto& q : js->peer_stream) if (q) cudaStreamDestroy(q);
    for (auto& q : js->peer_ev) if (q) cudaEventDestroy(q);
    if (js->ev_copy) cudaEventDestroy(js->ev_copy);
    if (js->ev_segs) cudaEventDestroy(js->ev_segs);
    igl_join_destroy(js->join);
    delete js;
}

}  // extern "C"
