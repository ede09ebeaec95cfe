//! `extern "C"` surface of include/igloo_cuda.h used by the operators.
use arrow::ffi::{FFI_ArrowArray, FFI_ArrowSchema};
use datafusion::error::{DataFusionError, Result as DataFusionResult};
use std::os::raw::c_char;

#[repr(C)]
pub struct IglConfig {
    pub device: i32,
    pub flags: i32,
    pub coalesce_rows: i64,
    pub hash_load_factor: f64,
    pub pinned_ring_bytes: i64,
}
pub enum IglCtx {}
pub enum IglOp {}
pub enum IglExpr {}
pub enum IglJoinShuffle {}
/// `igl_dev_col`: a device-resident Arrow column (values, validity bitmap, Utf8 offsets).
#[repr(C)]
pub struct IglDevCol {
    pub type_: i32,
    pub reserved: i32,
    pub data: *mut std::ffi::c_void,
    pub validity: *mut u8,
    pub offsets: *mut i32,
}

pub const IGL_OK: i32 = 0;
pub const IGL_AGAIN: i32 = 1;
pub const IGL_EOS: i32 = 2;

extern "C" {
    pub fn igl_last_error() -> *const c_char;
    pub fn igl_ctx_create(cfg: *const IglConfig, out: *mut *mut IglCtx) -> i32;
    pub fn igl_ctx_destroy(ctx: *mut IglCtx);
    pub fn igl_op_filter_project_create(
        ctx: *mut IglCtx, in_schema: *const FFI_ArrowSchema, pred: *const u8, pred_len: usize,
        proj: *const *const u8, proj_len: *const usize, proj_names: *const *const c_char, nproj: i32,
        out: *mut *mut IglOp,
    ) -> i32;
    pub fn igl_op_hash_join_create(
        ctx: *mut IglCtx, join_type: i32, left: *const FFI_ArrowSchema, right: *const FFI_ArrowSchema,
        left_key: *const u8, left_key_len: usize, right_key: *const u8, right_key_len: usize,
        out: *mut *mut IglOp,
    ) -> i32;
    pub fn igl_op_hash_join_create_ex(
        ctx: *mut IglCtx, join_type: i32, left: *const FFI_ArrowSchema, right: *const FFI_ArrowSchema,
        left_keys: *const *const u8, left_key_lens: *const usize, right_keys: *const *const u8, right_key_lens: *const usize,
        nkeys: i32, out: *mut *mut IglOp,
    ) -> i32;
    // multi-GPU join of one worker per GPU: communicator and the collective join
    pub fn igl_comm_unique_id(id_out: *mut u8) -> i32;
    pub fn igl_comm_init(ctx: *mut IglCtx, nranks: i32, rank: i32, id: *const u8) -> i32;
    pub fn igl_join_shuffle_create(
        ctx: *mut IglCtx, join_type: i32, left_types: *const i32, left_nullable: *const u8, n_left: i32,
        right_types: *const i32, right_nullable: *const u8, n_right: i32, left_key: *const IglExpr, right_key: *const IglExpr,
        mode: i32, out: *mut *mut IglJoinShuffle,
    ) -> i32;
    pub fn igl_join_shuffle_run(
        js: *mut IglJoinShuffle, right: *const IglDevCol, n_right: i64, left: *const IglDevCol, n_left: i64,
        out_rows: *mut i64, stream: *mut std::ffi::c_void,
    ) -> i32;
    pub fn igl_join_shuffle_output(js: *mut IglJoinShuffle, out: *mut IglDevCol, rows: *mut i64) -> i32;
    pub fn igl_join_shuffle_destroy(js: *mut IglJoinShuffle);
    pub fn igl_op_schema(op: *mut IglOp, out: *mut FFI_ArrowSchema) -> i32;
    pub fn igl_op_push_build(op: *mut IglOp, batch: *const FFI_ArrowArray) -> i32;
    pub fn igl_op_finish_build(op: *mut IglOp) -> i32;
    pub fn igl_op_push(op: *mut IglOp, batch: *const FFI_ArrowArray) -> i32;
    pub fn igl_op_finish_input(op: *mut IglOp) -> i32;
    pub fn igl_op_pull(op: *mut IglOp, out: *mut FFI_ArrowArray) -> i32;
    pub fn igl_op_destroy(op: *mut IglOp);
}

/// Status -> DataFusionError, the mapping INTEGRATION.md §2 describes.
pub fn check(rc: i32) -> DataFusionResult<i32> {
    if rc >= 0 {
        return Ok(rc);
    }
    let msg = unsafe { std::ffi::CStr::from_ptr(igl_last_error()) }.to_string_lossy().into_owned();
    Err(match rc {
        -2 => DataFusionError::NotImplemented(msg), // IGL_ERR_UNSUPPORTED: the planner keeps the CPU operator
        -7 | -10 => DataFusionError::Internal(msg), // filter.rs:50-55 ("Filter predicate must return boolean array")
        _ => DataFusionError::Execution(msg),
    })
}

/// One GPU context per process and device; operators share it.
pub struct GpuContext(pub *mut IglCtx);
unsafe impl Send for GpuContext {}
unsafe impl Sync for GpuContext {}

impl GpuContext {
    pub fn new(device: i32) -> DataFusionResult<Self> {
        let cfg = IglConfig { device, flags: 0, coalesce_rows: 4 << 20, hash_load_factor: 0.5, pinned_ring_bytes: 0 };
        let mut ctx = std::ptr::null_mut();
        check(unsafe { igl_ctx_create(&cfg, &mut ctx) })?;
        Ok(Self(ctx))
    }
}
impl Drop for GpuContext {
    fn drop(&mut self) {
        unsafe { igl_ctx_destroy(self.0) }
    }
}

/// Owning handle of one streaming operator (single consumer: `Send`, not `Sync`).
pub struct OpHandle(pub *mut IglOp);
unsafe impl Send for OpHandle {}
impl Drop for OpHandle {
    fn drop(&mut self) {
        unsafe { igl_op_destroy(self.0) }
    }
}
