/*
 * igloo_cuda.h — C ABI of libigloo_cuda.so: B200 (sm_100a) replacements for
 * Igloo's Filter / Projection / HashJoin physical operators.
 *
 * This is the surface a Rust `igloo-cuda` crate binds with `extern "C"`
 * (INTEGRATION.md shows the binding).  The reference defines no FFI of its own;
 * each entry point cites the reference interface it stands in for, all under
 * /root/reference/crates/engine/src/.
 *
 * Conventions
 *   - every call returns int32_t: 0 = IGL_OK, > 0 = status, < 0 = error class;
 *     the message is igl_last_error() (thread-local).  Nothing throws or aborts
 *     across the ABI.  Rust maps errors to DataFusionError (External / Internal /
 *     NotImplemented) the way the operators return `DataFusionResult`
 *     (filter.rs:39, hash_join.rs:41).
 *   - handles are opaque and single-owner: one thread at a time per handle
 *     (`Send`, not `Sync`), any thread may call.  *_destroy(NULL) is a no-op.
 *   - "dev" pointers are CUDA device pointers on the context's device.  Column
 *     buffers must be 16-byte aligned and readable up to the next multiple of
 *     64 bytes (the padding Arrow recommends and igl_dev_alloc guarantees):
 *     tiles are staged with 16-byte-granular bulk copies.
 *   - there is no CPU fallback: without a CUDA device igl_ctx_create fails.
 */
#ifndef IGLOO_CUDA_H
#define IGLOO_CUDA_H

#include <stddef.h>
#include <stdint.h>

#include "igloo_expr.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- Arrow C Data Interface (verbatim ABI; what arrow-rs `arrow::ffi` and
 *      pyarrow `_export_to_c` / `_import_from_c` exchange).  Stands in for
 *      `arrow::record_batch::RecordBatch` at the operator boundary
 *      (physical_plan.rs:13). ---------------------------------------------- */
#ifndef ARROW_C_DATA_INTERFACE
#define ARROW_C_DATA_INTERFACE
#define ARROW_FLAG_DICTIONARY_ORDERED 1
#define ARROW_FLAG_NULLABLE 2
#define ARROW_FLAG_MAP_KEYS_SORTED 4
struct ArrowSchema {
    const char* format;
    const char* name;
    const char* metadata;
    int64_t flags;
    int64_t n_children;
    struct ArrowSchema** children;
    struct ArrowSchema* dictionary;
    void (*release)(struct ArrowSchema*);
    void* private_data;
};
struct ArrowArray {
    int64_t length;
    int64_t null_count;
    int64_t offset;
    int64_t n_buffers;
    int64_t n_children;
    const void** buffers;
    struct ArrowArray** children;
    struct ArrowArray* dictionary;
    void (*release)(struct ArrowArray*);
    void* private_data;
};
#endif

/* ---- status / error classes ---------------------------------------------- */
enum igl_status {
    IGL_OK = 0,
    IGL_AGAIN = 1, /* pull: nothing ready yet, push more or finish input */
    IGL_EOS = 2,   /* pull: stream finished                              */
    IGL_ERR_INVALID_ARG = -1,
    IGL_ERR_UNSUPPORTED = -2, /* DataFusionError::NotImplemented           */
    IGL_ERR_OOM = -3,
    IGL_ERR_CUDA = -4,
    IGL_ERR_NCCL = -5,
    IGL_ERR_DIV_BY_ZERO = -6,           /* ArrowError::DivideByZero         */
    IGL_ERR_PREDICATE_NOT_BOOLEAN = -7, /* filter.rs:50-55                  */
    IGL_ERR_ARITH_OVERFLOW = -8,        /* ArrowError::ArithmeticOverflow   */
    IGL_ERR_CAST = -9,                  /* ArrowError::CastError            */
    IGL_ERR_INTERNAL = -10
};

/* Thread-local message of the last failing call on this thread. */
const char* igl_last_error(void);
int32_t igl_version(void);

/* ---- context: device, streams, device buffer pool, pinned staging ring.
 *      New (the reference has no device); its knobs replace the literals
 *      batch size 1024 / channel depth 10 (parquet_scan.rs:44,54). --------- */
typedef struct igl_ctx igl_ctx;
typedef struct igl_config {
    int32_t device;           /* CUDA ordinal                                */
    int32_t flags;            /* reserved, 0                                 */
    int64_t coalesce_rows;    /* host batches are coalesced to this many rows
                                 per device chunk; 0 = default (4 Mi)        */
    double hash_load_factor;  /* join table fill; 0 = default (0.5)          */
    int64_t pinned_ring_bytes;/* staging ring size; 0 = default (256 MiB)    */
} igl_config;

int32_t igl_ctx_create(const igl_config* cfg, igl_ctx** out);
void igl_ctx_destroy(igl_ctx* ctx);
int32_t igl_ctx_device(const igl_ctx* ctx);
/* Number of kernels this context has launched (bench.py "gpu_launches"). */
int64_t igl_ctx_launch_count(const igl_ctx* ctx);

/* Device-resident batch buffer pool and pinned host memory (north_star:
 * "pinned-memory batch upload path", "device-resident batch buffer pool"). */
int32_t igl_dev_alloc(igl_ctx* ctx, size_t bytes, void** out);
int32_t igl_dev_free(igl_ctx* ctx, void* ptr);
int32_t igl_host_alloc_pinned(igl_ctx* ctx, size_t bytes, void** out);
int32_t igl_host_free_pinned(igl_ctx* ctx, void* ptr);
int32_t igl_memcpy_h2d(igl_ctx* ctx, void* dev_dst, const void* host_src, size_t bytes, void* stream);
int32_t igl_memcpy_d2h(igl_ctx* ctx, void* host_dst, const void* dev_src, size_t bytes, void* stream);
int32_t igl_stream_sync(igl_ctx* ctx, void* stream);

/* ---- expressions: `Arc<dyn PhysicalExpr>` (physical_planner.rs:60-64,86-90,
 *      107-116) as compiled bytecode.  col_types / col_nullable describe the
 *      child's schema the expression is lowered against. ------------------- */
typedef struct igl_expr igl_expr;
int32_t igl_expr_compile(igl_ctx* ctx, const uint8_t* bytecode, size_t len,
                         const int32_t* col_types, const uint8_t* col_nullable, int32_t ncols,
                         igl_expr** out);
/* PhysicalExpr::data_type / ::nullable (physical_planner.rs:69-70). */
int32_t igl_expr_type(const igl_expr* e, int32_t* type, int32_t* nullable);
void igl_expr_destroy(igl_expr* e);

/* ---- device-resident columns --------------------------------------------- */
typedef struct igl_dev_col {
    int32_t type;      /* igl_type                                            */
    int32_t reserved;
    void* data;        /* values; BOOL: Arrow bit-packed; UTF8: bytes         */
    uint8_t* validity; /* Arrow validity bitmap (LSB first) or NULL           */
    int32_t* offsets;  /* UTF8 only: length+1 offsets                         */
} igl_dev_col;

/* ---- fused Filter + Projection on device-resident columns.
 *      FilterExec::execute (filter.rs:39-68): pred != NULL, nproj == 0 keeps
 *        every input column, rows where the mask is valid AND true.
 *      ProjectionExec::execute (projection.rs:48-72): pred == NULL.
 *      Projection(Filter(x)) fused by the planner: both.
 *      Row order is preserved. -------------------------------------------- */
typedef struct igl_fp igl_fp;
int32_t igl_fp_create(igl_ctx* ctx, const int32_t* in_types, const uint8_t* in_nullable, int32_t n_in,
                      const igl_expr* pred, const igl_expr* const* proj, int32_t nproj, igl_fp** out);
int32_t igl_fp_num_outputs(const igl_fp* fp);
int32_t igl_fp_output_type(const igl_fp* fp, int32_t j, int32_t* type, int32_t* nullable);
/* Runs on `stream` (NULL = the context's stream) and returns after the stream
 * has drained.  out[j].data (and .validity when output j is nullable) must hold
 * `nrows` rows.  *out_rows receives the surviving row count.  A FilterExec that
 * keeps 0 rows reports 0: the caller drops the batch (filter.rs:59-63). */
int32_t igl_fp_run(igl_fp* fp, const igl_dev_col* in, int64_t nrows, igl_dev_col* out,
                   int64_t* out_rows, void* stream);
/* Same launch without the trailing synchronise / error read-back, for timing
 * with CUDA events: counts and error flags are fetched by igl_fp_finish. */
int32_t igl_fp_launch(igl_fp* fp, const igl_dev_col* in, int64_t nrows, igl_dev_col* out, void* stream);
int32_t igl_fp_finish(igl_fp* fp, int64_t* out_rows, void* stream);
void igl_fp_destroy(igl_fp* fp);
/* Run-time specialisation (no counterpart in the reference; DataFusion interprets its
 * PhysicalExpr trees).  The kernels above interpret the operator's program; once a plan has
 * been given IGL_FP_JIT_MIN_ROWS rows in total (environment, default 64 Mi; IGL_FP_JIT=0 never,
 * =1 always) its next launch first compiles the same device source with the program fixed as constants (NVRTC, sm_100a,
 * loaded through cudaLibraryLoadData, cached per process) and runs that instead — same
 * arithmetic, same results.  Without libnvrtc.so.12 the interpreter stays in use.
 * igl_fp_is_specialised: 1 when the plan's kernels are the specialised ones.
 * igl_fp_jit_compile_only / igl_jit_selftest: compile without loading (diagnostics). */
int32_t igl_fp_is_specialised(const igl_fp* fp);
/* Compile and load the plan's specialised kernels NOW (e.g. while the query is planned) instead of
 * inside the launch that crosses the row threshold; without NVRTC it does nothing and returns OK. */
int32_t igl_fp_prepare(igl_fp* fp);
int32_t igl_fp_jit_compile_only(igl_fp* fp, int64_t* cubin_bytes);
int32_t igl_jit_selftest(int64_t* cubin_bytes);
/* No counterpart in the reference.  igl_fp_run for callers that must not
 * allocate device memory while running (engines that reserve all of their HBM up
 * front, latency-critical streams that cannot absorb an allocation's implicit
 * synchronisation): the scan scratch is a fixed 88 KiB device array inside the
 * library instead of a pooled block, so no cudaMalloc / cudaFree / pool call is
 * made.  Limits: at most 4,194,304 rows per call (IGL_ERR_INVALID_ARG beyond);
 * plans with boolean or nullable outputs are rejected (IGL_ERR_UNSUPPORTED).
 * All calls on a GPU share that array: calls that overlap on the GPU race, so
 * serialise them.  The first call on each device resolves the array's address;
 * make it before any stream capture begins. */
int32_t igl_fp_run_static(igl_fp* fp, const igl_dev_col* in, int64_t nrows, igl_dev_col* out,
                          int64_t* out_rows, void* stream);

/* ---- hash join on device-resident columns.
 *      HashJoinExec::new (hash_join.rs:35-76): schema = left ++ right.
 *      build  (hash_join.rs:100-128): right side drained into the table.
 *      probe  (hash_join.rs:131-213): left rows looked up; Inner emits one row
 *        per (left row, matching right row); Left/Full add the null-extended
 *        row for an unmatched left row; Right == Inner and Full == Left because
 *        the reference never emits unmatched right rows (:189, :263-276).
 *      Key identity follows the printed form the reference hashes (:117-122):
 *      NULL == NULL, NaN == NaN, -0.0 != 0.0, differing key types never match.
 *      Output row order within one call is probe order (matches of one probe
 *      row in unspecified order); parity is on row multisets. -------------- */
enum igl_join_type { IGL_JOIN_INNER = 0, IGL_JOIN_LEFT = 1, IGL_JOIN_RIGHT = 2, IGL_JOIN_FULL = 3 };
typedef struct igl_join igl_join;
int32_t igl_join_create(igl_ctx* ctx, int32_t join_type,
                        const int32_t* left_types, const uint8_t* left_nullable, int32_t n_left,
                        const int32_t* right_types, const uint8_t* right_nullable, int32_t n_right,
                        const igl_expr* left_key, const igl_expr* right_key, igl_join** out);
/* HashJoinExec::new with everything it accepts (hash_join.rs:35-41: `left_keys` / `right_keys` are
 * vectors of expressions; physical_planner.rs:106-120 pushes one pair per `join.on` entry; the
 * key of a row is the printed form of ALL its key values, hash_join.rs:117-122): 1 to 4 key
 * pairs, each any expression of the side's columns (evaluated by the projection kernel; their
 * errors — division by zero, casts — surface from build / probe), and Boolean columns on either
 * side.  A pair whose types differ never matches.  Two Int32 keys that cannot be NULL are packed
 * into one exact 64-bit key; other key lists are hashed to 64 bits and the joined rows are then
 * filtered by the real, null-safe condition (plain-column keys only; a Left / Full join counts
 * the true matches per probe row and null-extends the rows left without one).
 * Utf8 columns are carried by the streaming operator (igl_op_hash_join_create), not here.
 * igl_join_probe_count on a hashed key list is an upper bound.  One plain key over fixed-width
 * columns is exactly igl_join_create. */
int32_t igl_join_create_ex(igl_ctx* ctx, int32_t join_type,
                           const int32_t* left_types, const uint8_t* left_nullable, int32_t n_left,
                           const int32_t* right_types, const uint8_t* right_nullable, int32_t n_right,
                           const igl_expr* const* left_keys, const igl_expr* const* right_keys, int32_t nkeys,
                           igl_join** out);
/* Build from the complete right side.  The columns are borrowed until
 * igl_join_destroy (right payload is gathered from them by row id). */
int32_t igl_join_build(igl_join* j, const igl_dev_col* right, int64_t nrows, void* stream);
/* Exact number of rows igl_join_probe will emit for this left chunk. */
int32_t igl_join_probe_count(igl_join* j, const igl_dev_col* left, int64_t nrows, int64_t* out_rows, void* stream);
/* out has n_left + n_right columns of capacity `out_capacity` rows; right-side
 * columns need .validity for Left/Full joins or when the source is nullable. */
int32_t igl_join_probe(igl_join* j, const igl_dev_col* left, int64_t nrows, igl_dev_col* out,
                       int64_t out_capacity, int64_t* out_rows, void* stream);
/* 1 when the build side held duplicate keys (chain walks), else 0. */
int32_t igl_join_has_duplicates(const igl_join* j);
/* Large build sides are reordered by the top bits of the join hash before the table is
 * built, and every probe chunk likewise, so each 16 MB region of the table is filled and
 * probed while it is L2-resident.  igl_join_partition_bits: log2 of the fan-out
 * igl_join_build would use for `build_rows` rows (0 = none).  A caller whose rows already
 * arrive in that order (igl_partition_scatter_radix) switches the join's own reordering
 * off with igl_join_set_prepartitioned(j, bits): bits = the number of leading hash bits the rows
 * are ordered by (1 = ordered, not told by how many; 0 = back to the join's own reordering).
 * A large unique build side is then ordered once more, down to 64 KB table sub-regions that one
 * CTA each builds in shared memory.  No counterpart in the reference. */
int32_t igl_join_partition_bits(igl_ctx* ctx, int64_t build_rows);
int32_t igl_join_set_prepartitioned(igl_join* j, int32_t on);
/* Device time (ms, CUDA events on the call's stream) of the last build and probe of `j`:
 * out[0] build-side partitioning (both levels when the table is built in shared memory), out[1]
 * the table build (init + global-atomic build, or the shared-memory build), out[2] probe-side
 * partitioning, out[3] the probe kernel (-1 where nothing was recorded).  Measurement aid. */
int32_t igl_join_phase_ms(igl_join* j, double out[4]);
void igl_join_destroy(igl_join* j);

/* ---- hash repartition for the multi-GPU join (the "worker fragment exchange":
 *      FragmentType::Shuffle is declared but never built in the reference,
 *      crates/coordinator/src/fragment.rs:8-13; GetDataForTask returns empty
 *      bytes, crates/worker/src/service.rs:26-32).  dest = top bits of
 *      mix64(key), disjoint from the bits the local table uses. ------------ */
int32_t igl_partition_histogram(igl_ctx* ctx, const igl_dev_col* key, int64_t nrows, int32_t nparts,
                                int64_t* dev_hist /* [nparts] */, void* stream);
/* Scatters every column of `cols` to dst[p * ncols + c] + dst_row_offset[p] rows;
 * dst pointers may be peer (NVLink-mapped) device pointers. */
int32_t igl_partition_scatter(igl_ctx* ctx, const igl_dev_col* cols, int32_t ncols, int32_t key_col,
                              int64_t nrows, int32_t nparts, void* const* dst /* host [nparts*ncols] */,
                              const int64_t* dst_row_offset /* host [nparts] */, void* stream);
/* The same exchange fused with the join's radix partitioning: bins = ndest << log2_parts,
 * bin = dest * 2^log2_parts + (top log2_parts bits of the join hash).  The histogram counts
 * rows per bin; the scatter sorts each 4096-row tile by bin in shared memory and stores runs
 * to dst[dest * ncols + c] + bin_row_offset[bin] rows, so a receiver that lays its slab out
 * partition by partition gets its rows in the order igl_join_build / igl_join_probe want.
 * Non-nullable fixed-width columns of at most 46 bytes per row (16-byte aligned; 24 bytes
 * per row otherwise): igl_partition_radix_supported tells. */
int32_t igl_partition_radix_supported(const igl_dev_col* cols, int32_t ncols);
int32_t igl_partition_histogram_radix(igl_ctx* ctx, const igl_dev_col* key, int64_t nrows, int32_t ndest,
                                      int32_t log2_parts, int64_t* dev_hist /* [ndest << log2_parts] */,
                                      void* stream);
int32_t igl_partition_scatter_radix(igl_ctx* ctx, const igl_dev_col* cols, int32_t ncols, int32_t key_col,
                                    int64_t nrows, int32_t ndest, int32_t log2_parts,
                                    void* const* dst /* host [ndest*ncols] */,
                                    const int64_t* bin_row_offset /* host [ndest << log2_parts] */, void* stream);
/* CUDA IPC plumbing so one process per GPU can map its peers' receive slabs. */
int32_t igl_ipc_export(igl_ctx* ctx, void* dev_ptr, uint8_t handle_out[64]);
int32_t igl_ipc_open(igl_ctx* ctx, const uint8_t handle[64], void** dev_ptr_out);
int32_t igl_ipc_close(igl_ctx* ctx, void* dev_ptr);

/* ---- the hash join across the GPUs of one node: one process per GPU, each
 *      holding a row-range shard of both sides.  Stands in for the exchange the reference only
 *      declares: FragmentType::Shuffle (crates/coordinator/src/fragment.rs:8-13), the join
 *      fragment fed by scan fragments (crates/coordinator/src/distributed_planner.rs:65-92,
 *      distributed_executor.rs:105-166), a worker process per executor
 *      (crates/worker/src/main.rs:14-52).
 *      igl_comm_unique_id / igl_comm_init: rank 0 makes an id, the host distributes it (the
 *        coordinator's registration reply, coordinator.proto:60-69, is the natural carrier),
 *        every rank joins; NCCL is resolved with dlopen (IGL_ERR_NCCL when absent or failing).
 *      igl_join_shuffle_run: collective — every rank calls it with its shard.  Both sides are
 *        radix-partitioned locally by (destination GPU, partition of the join hash); counts
 *        and buffer handles travel in one ncclAllGather; the build rows a GPU owns are pulled
 *        from its peers' staging buffers over NVLink and the table is built; the probe kernel
 *        then reads its probe rows straight out of the peers' staging buffers, tile by tile,
 *        while it probes (mode IGL_EXCHANGE_P2P), or both sides first travel by grouped
 *        ncclSend/ncclRecv (IGL_EXCHANGE_NCCL, the all-to-all-v baseline).  Equal keys meet
 *        on one GPU; the output stays sharded, in library-owned device columns that
 *        igl_join_shuffle_output lends until the next run or the destroy.
 *      Non-nullable Int32/Int64/Float64 columns of at most 46 bytes per row and side; column-
 *      reference keys.  Key semantics and join types as igl_join_*. ------------------------- */
enum igl_exchange_mode { IGL_EXCHANGE_P2P = 0, IGL_EXCHANGE_NCCL = 1 };
int32_t igl_comm_unique_id(uint8_t id_out[128]);
int32_t igl_comm_init(igl_ctx* ctx, int32_t nranks, int32_t rank, const uint8_t id[128]);
int32_t igl_comm_rank(const igl_ctx* ctx, int32_t* rank, int32_t* nranks);
typedef struct igl_join_shuffle igl_join_shuffle;
int32_t igl_join_shuffle_create(igl_ctx* ctx, int32_t join_type,
                                const int32_t* left_types, const uint8_t* left_nullable, int32_t n_left,
                                const int32_t* right_types, const uint8_t* right_nullable, int32_t n_right,
                                const igl_expr* left_key, const igl_expr* right_key, int32_t mode,
                                igl_join_shuffle** out);
int32_t igl_join_shuffle_run(igl_join_shuffle* js, const igl_dev_col* right, int64_t n_right,
                             const igl_dev_col* left, int64_t n_left, int64_t* out_rows, void* stream);
int32_t igl_join_shuffle_output(igl_join_shuffle* js, igl_dev_col* out /* [n_left + n_right] */, int64_t* rows);
/* Device time (ms) of the phases of the last run on this rank: [0] histograms + all-gather,
 * [1] local partitioning on the call's stream (p2p: the build side only — the probe side is
 * partitioned on a second stream under [2] and [3]), [2] barrier / NCCL exchange, [3] build-side
 * pull + table build, [4] probe (fused with the probe-side transfer in p2p mode), [5] closing
 * barrier, [6] total, [7] the probe kernel alone.  stats: [0] build rows received, [1] probe rows
 * received, [2] bytes read from peer GPUs, [3] log2 of the partition count, [4] 1 = fused probe,
 * [5] output rows, [6] probe-side partitioning on the second stream (us, -1 = not overlapped),
 * [7] time the probe waited for it after the build (us), [8] probe-side bytes the copy engines
 * prefetched over NVLink while the table was built, [9] partitions prefetched, [10] hot keys:
 * when one GPU would receive more than 1.15 x its share of the probe rows, the keys behind it
 * are found from a sample, their probe rows stay on the GPU that holds them and their build rows
 * go to every GPU (0 = the hash alone decided). */
int32_t igl_join_shuffle_phase_ms(igl_join_shuffle* js, double out[8]);
int32_t igl_join_shuffle_stats(igl_join_shuffle* js, int64_t out[12]);
void igl_join_shuffle_destroy(igl_join_shuffle* js);
/* The exchange plan as pure host arithmetic (no GPU needed; igl_join_shuffle_run uses it).
 * counts[nranks][bins], bins = nranks << log2_parts, destination-major: rows of each source per
 * (destination, partition).  src_offsets[nranks][bins]: first row of every bin inside its
 * source's staging buffer (bins start on multiples of 4 rows).  For `rank`, in the order its
 * join walks them (partition by partition, source by source): seg_src / seg_row0 / seg_rows
 * [(nranks << log2_parts) / nranks * nranks entries = parts * nranks], and their total. */
int32_t igl_shuffle_plan(int32_t nranks, int32_t rank, int32_t log2_parts, const int64_t* counts,
                         int64_t* src_offsets, int32_t* seg_src, int64_t* seg_row0, int64_t* seg_rows,
                         int64_t* total_rows);
/* The same with one extra destination (broadcast = 1: bins = (nranks + 1) << log2_parts) whose
 * rows EVERY rank reads — the build rows of hot keys, whose probe rows stay on the GPU that holds
 * them.  A partition's runs are then the rank's own bins of every source followed by the
 * broadcast bins of every source: 2 * nranks entries per partition. */
int32_t igl_shuffle_plan_ex(int32_t nranks, int32_t rank, int32_t log2_parts, int32_t broadcast, const int64_t* counts,
                            int64_t* src_offsets, int32_t* seg_src, int64_t* seg_row0, int64_t* seg_rows,
                            int64_t* total_rows);

/* ---- synthetic data + multiset digest (SURVEY.md §8d generator), device side */
enum igl_gen_kind {
    IGL_GEN_I32_UNIFORM = 1, /* mulhi64(u, m)                                  */
    IGL_GEN_I64_UNIFORM = 2, /* mulhi64(u, m); m == 0 => raw u                 */
    IGL_GEN_F64_UNIT = 3,    /* (u >> 11) * 2^-53                              */
    IGL_GEN_I64_PERM = 4,    /* bijection of row index on [0, 2^bits)          */
    IGL_GEN_I64_LOOKUP = 5,  /* u(seed, col, perm^-1(key[i])): the generated
                                value of the build row holding key[i]          */
    IGL_GEN_I64_TPCH_ORDERKEY = 6, /* C4: ((row >> 3) << 5 | row & 7) + 1      */
    IGL_GEN_I64_LINE_ORDER = 7,    /* C4: lineitem row -> index of its order (m
                                      orders, 1-7 lines each, 4 on average)     */
    IGL_GEN_I64_ZIPF = 8     /* C5: Zipf(1.1) rank - 1 over [0, m) by inverse
                                CDF of u; bits > 0: mapped through the same
                                bijection as IGL_GEN_I64_PERM                  */
};
/* Kinds 1-3 and 6 with dev_keys != NULL generate the value of row dev_keys[i]
 * instead of row0 + i (expected join payloads, recomputed without a table). */
int32_t igl_gen_column(igl_ctx* ctx, int32_t kind, uint64_t seed, uint64_t col_id, int64_t row0,
                       int64_t n, uint64_t m, int32_t bits, const int64_t* dev_keys, void* dev_out,
                       void* stream);
int32_t igl_digest_rows(igl_ctx* ctx, const igl_dev_col* cols, int32_t ncols, int64_t nrows,
                        uint64_t out[3], void* stream);

/* ---- streaming operators on host Arrow batches (push / pull).
 *      `trait ExecutionPlan::execute() -> BoxStream<RecordBatch>`
 *      (physical_plan.rs:10-17): the Rust operator's stream polls igl_op_pull
 *      and feeds igl_op_push from its child stream.  Input batches are struct
 *      arrays borrowed for the duration of the call; output batches are owned
 *      by the caller through their `release` callbacks. -------------------- */
typedef struct igl_op igl_op;
int32_t igl_op_filter_project_create(igl_ctx* ctx, const struct ArrowSchema* in_schema,
                                     const uint8_t* pred, size_t pred_len,
                                     const uint8_t* const* proj, const size_t* proj_len,
                                     const char* const* proj_names, int32_t nproj, igl_op** out);
int32_t igl_op_hash_join_create(igl_ctx* ctx, int32_t join_type,
                                const struct ArrowSchema* left_schema, const struct ArrowSchema* right_schema,
                                const uint8_t* left_key, size_t left_key_len,
                                const uint8_t* right_key, size_t right_key_len, igl_op** out);
/* The same with every key pair of `join.on` (physical_planner.rs:106-120): nkeys bytecodes per side,
 * any expressions of the side's columns; Boolean and Utf8 columns may ride along (igl_join_create_ex;
 * strings by row id), and a bare Utf8 column may be a key on both sides of a pair (hash_join.rs:120
 * prints any Arrow type): the device joins on a 64-bit hash of the bytes and every joined pair is then
 * compared byte for byte; NULL == NULL as for every key. */
int32_t igl_op_hash_join_create_ex(igl_ctx* ctx, int32_t join_type,
                                   const struct ArrowSchema* left_schema, const struct ArrowSchema* right_schema,
                                   const uint8_t* const* left_keys, const size_t* left_key_lens,
                                   const uint8_t* const* right_keys, const size_t* right_key_lens, int32_t nkeys,
                                   igl_op** out);
/* ExecutionPlan::schema (physical_plan.rs:16). */
int32_t igl_op_schema(igl_op* op, struct ArrowSchema* out);
/* Join only: right (build) child batches, then finish_build (hash_join.rs:100-128). */
int32_t igl_op_push_build(igl_op* op, const struct ArrowArray* batch);
int32_t igl_op_finish_build(igl_op* op);
int32_t igl_op_push(igl_op* op, const struct ArrowArray* batch);
int32_t igl_op_finish_input(igl_op* op);
int32_t igl_op_pull(igl_op* op, struct ArrowArray* out);
void igl_op_destroy(igl_op* op);

#ifdef __cplusplus
}
#endif
#endif /* IGLOO_CUDA_H */
