//! Device-level surface of include/igloo_cuda.h: the same kernels on columns that already live in
//! HBM (a GPU-side scan, a fragment that keeps its output on the device for the next operator).
//! The streaming operators of `ops.rs` are built from exactly these calls inside the library.
//! Unbuilt like the rest of the crate (no Rust toolchain in the image this was written in).
use crate::ffi::{IglCtx, IglDevCol, IglExpr, IglJoinShuffle};
use std::ffi::c_void;

pub enum IglFp {}
pub enum IglJoin {}

extern "C" {
    pub fn igl_version() -> i32;
    pub fn igl_ctx_device(ctx: *const IglCtx) -> i32;
    pub fn igl_ctx_launch_count(ctx: *const IglCtx) -> i64;

    // device buffer pool, pinned host memory, copies
    pub fn igl_dev_alloc(ctx: *mut IglCtx, bytes: usize, out: *mut *mut c_void) -> i32;
    pub fn igl_dev_free(ctx: *mut IglCtx, ptr: *mut c_void) -> i32;
    pub fn igl_host_alloc_pinned(ctx: *mut IglCtx, bytes: usize, out: *mut *mut c_void) -> i32;
    pub fn igl_host_free_pinned(ctx: *mut IglCtx, ptr: *mut c_void) -> i32;
    pub fn igl_memcpy_h2d(ctx: *mut IglCtx, dev_dst: *mut c_void, host_src: *const c_void, bytes: usize, stream: *mut c_void) -> i32;
    pub fn igl_memcpy_d2h(ctx: *mut IglCtx, host_dst: *mut c_void, dev_src: *const c_void, bytes: usize, stream: *mut c_void) -> i32;
    pub fn igl_stream_sync(ctx: *mut IglCtx, stream: *mut c_void) -> i32;

    // expressions: Arc<dyn PhysicalExpr> as compiled bytecode (expr.rs encodes it)
    pub fn igl_expr_compile(
        ctx: *mut IglCtx, bytecode: *const u8, len: usize, col_types: *const i32, col_nullable: *const u8, ncols: i32,
        out: *mut *mut IglExpr,
    ) -> i32;
    pub fn igl_expr_type(e: *const IglExpr, type_: *mut i32, nullable: *mut i32) -> i32;
    pub fn igl_expr_destroy(e: *mut IglExpr);

    // FilterExec / ProjectionExec / Projection(Filter) fused (filter.rs:39-68, projection.rs:48-72)
    pub fn igl_fp_create(
        ctx: *mut IglCtx, in_types: *const i32, in_nullable: *const u8, n_in: i32, pred: *const IglExpr,
        proj: *const *const IglExpr, nproj: i32, out: *mut *mut IglFp,
    ) -> i32;
    pub fn igl_fp_num_outputs(fp: *const IglFp) -> i32;
    pub fn igl_fp_output_type(fp: *const IglFp, j: i32, type_: *mut i32, nullable: *mut i32) -> i32;
    pub fn igl_fp_run(fp: *mut IglFp, input: *const IglDevCol, nrows: i64, out: *mut IglDevCol, out_rows: *mut i64, stream: *mut c_void) -> i32;
    pub fn igl_fp_launch(fp: *mut IglFp, input: *const IglDevCol, nrows: i64, out: *mut IglDevCol, stream: *mut c_void) -> i32;
    pub fn igl_fp_finish(fp: *mut IglFp, out_rows: *mut i64, stream: *mut c_void) -> i32;
    /// igl_fp_run without any device allocation (fixed 88 KiB scratch inside the library): at most
    /// 4,194,304 rows per call, no boolean / nullable outputs, calls on one GPU must not overlap.
    pub fn igl_fp_run_static(fp: *mut IglFp, input: *const IglDevCol, nrows: i64, out: *mut IglDevCol, out_rows: *mut i64, stream: *mut c_void) -> i32;
    /// Compile the plan's specialised kernels now (while planning) instead of on the first large batch.
    pub fn igl_fp_prepare(fp: *mut IglFp) -> i32;
    pub fn igl_fp_is_specialised(fp: *const IglFp) -> i32;
    pub fn igl_fp_destroy(fp: *mut IglFp);

    // HashJoinExec on device-resident columns (hash_join.rs:35-76, 100-213)
    pub fn igl_join_create(
        ctx: *mut IglCtx, join_type: i32, left_types: *const i32, left_nullable: *const u8, n_left: i32,
        right_types: *const i32, right_nullable: *const u8, n_right: i32, left_key: *const IglExpr, right_key: *const IglExpr,
        out: *mut *mut IglJoin,
    ) -> i32;
    /// Every `join.on` pair (physical_planner.rs:106-120): 1 to 4 key expressions per side.
    pub fn igl_join_create_ex(
        ctx: *mut IglCtx, join_type: i32, left_types: *const i32, left_nullable: *const u8, n_left: i32,
        right_types: *const i32, right_nullable: *const u8, n_right: i32, left_keys: *const *const IglExpr,
        right_keys: *const *const IglExpr, nkeys: i32, out: *mut *mut IglJoin,
    ) -> i32;
    pub fn igl_join_build(j: *mut IglJoin, right: *const IglDevCol, nrows: i64, stream: *mut c_void) -> i32;
    pub fn igl_join_probe_count(j: *mut IglJoin, left: *const IglDevCol, nrows: i64, out_rows: *mut i64, stream: *mut c_void) -> i32;
    pub fn igl_join_probe(
        j: *mut IglJoin, left: *const IglDevCol, nrows: i64, out: *mut IglDevCol, out_capacity: i64, out_rows: *mut i64,
        stream: *mut c_void,
    ) -> i32;
    pub fn igl_join_has_duplicates(j: *const IglJoin) -> i32;
    pub fn igl_join_partition_bits(ctx: *mut IglCtx, build_rows: i64) -> i32;
    /// `bits` = the number of leading join-hash bits the caller's rows are already ordered by (0 = not ordered).
    pub fn igl_join_set_prepartitioned(j: *mut IglJoin, bits: i32) -> i32;
    pub fn igl_join_phase_ms(j: *mut IglJoin, out: *mut f64) -> i32;
    pub fn igl_join_destroy(j: *mut IglJoin);

    // measurements of the multi-GPU join (ffi.rs declares the join itself)
    pub fn igl_comm_rank(ctx: *const IglCtx, rank: *mut i32, nranks: *mut i32) -> i32;
    pub fn igl_join_shuffle_phase_ms(js: *mut IglJoinShuffle, out: *mut f64) -> i32;
    pub fn igl_join_shuffle_stats(js: *mut IglJoinShuffle, out: *mut i64) -> i32;
    /// The exchange plan as pure host arithmetic (what igl_join_shuffle_run derives on every rank).
    pub fn igl_shuffle_plan(
        nranks: i32, rank: i32, log2_parts: i32, counts: *const i64, src_offsets: *mut i64, seg_src: *mut i32,
        seg_row0: *mut i64, seg_rows: *mut i64, total_rows: *mut i64,
    ) -> i32;
}
