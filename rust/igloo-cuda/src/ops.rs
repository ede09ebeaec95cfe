//! GPU operators with the reference's constructors and `ExecutionPlan` trait.
//! Each `execute()` drives one `igl_op` (include/igloo_cuda.h "streaming operators"): input
//! batches are pushed as Arrow C Data Interface struct arrays, outputs are pulled the same way.
use crate::expr::to_bytecode;
use crate::ffi::*;
use arrow::array::{Array, StructArray};
use arrow::datatypes::Schema;
use arrow::ffi::{from_ffi, to_ffi, FFI_ArrowArray, FFI_ArrowSchema};
use arrow::record_batch::RecordBatch;
use datafusion::error::{DataFusionError, Result as DataFusionResult};
use datafusion::physical_expr::PhysicalExpr;
use futures::stream::{BoxStream, StreamExt};
use igloo_engine::operators::JoinType;            // hash_join.rs:16-21
use igloo_engine::physical_plan::ExecutionPlan;   // physical_plan.rs:10-17
use std::ffi::CString;
use std::sync::Arc;

type BatchStream = BoxStream<'static, DataFusionResult<RecordBatch>>;

fn push(op: &OpHandle, batch: &RecordBatch, build_side: bool) -> DataFusionResult<()> {
    let data = StructArray::from(batch.clone()).into_data();
    let (array, _schema) = to_ffi(&data)?; // borrowed by the library until the call returns
    check(unsafe { if build_side { igl_op_push_build(op.0, &array) } else { igl_op_push(op.0, &array) } })?;
    Ok(())
}

/// One output batch, `None` on IGL_AGAIN; `Err(..)`/end handled by the caller through `eos`.
fn pull(op: &OpHandle, schema: &Arc<Schema>, eos: &mut bool) -> DataFusionResult<Option<RecordBatch>> {
    let mut out = FFI_ArrowArray::empty();
    match check(unsafe { igl_op_pull(op.0, &mut out) })? {
        IGL_OK => {
            let ffi_schema = FFI_ArrowSchema::try_from(arrow::datatypes::DataType::Struct(schema.fields().clone()))?;
            // buffers are library-owned pinned host memory; the array's release callback recycles them
            let data = unsafe { from_ffi(out, &ffi_schema) }?;
            let batch = RecordBatch::from(StructArray::from(data)).with_schema(schema.clone())?;
            Ok(Some(batch))
        }
        IGL_EOS => { *eos = true; Ok(None) }
        _ => Ok(None),
    }
}

fn out_schema(op: &OpHandle) -> DataFusionResult<Arc<Schema>> {
    let mut s = FFI_ArrowSchema::empty();
    check(unsafe { igl_op_schema(op.0, &mut s) })?;
    Ok(Arc::new(Schema::try_from(&s)?))
}

fn stream_through(op: OpHandle, mut input: BatchStream, schema: Arc<Schema>) -> BatchStream {
    Box::pin(async_stream::try_stream! {
        let mut eos = false;
        while let Some(batch) = input.next().await {
            push(&op, &batch?, false)?;
            while let Some(out) = pull(&op, &schema, &mut eos)? { yield out; }
        }
        check(unsafe { igl_op_finish_input(op.0) })?;
        // after finish_input the operator drains: IGL_AGAIN now only means "a download is still in
        // flight", so hand the executor back its thread instead of spinning on the FFI call
        while !eos {
            match pull(&op, &schema, &mut eos)? {
                Some(out) => yield out,
                None if !eos => tokio::task::yield_now().await,
                None => {}
            }
        }
    })
}

// ---- Filter (filter.rs:19, 39-68), optionally with the Projection folded into it -----------
// no derive(Debug): `dyn ExecutionPlan` has no Debug bound (physical_plan.rs:11)
impl std::fmt::Debug for GpuFilterExec {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        f.debug_struct("GpuFilterExec").field("predicate", &self.predicate).field("fused_projection", &self.projection.is_some()).finish_non_exhaustive()
    }
}

#[derive(Clone)]
pub struct GpuFilterExec {
    ctx: Arc<GpuContext>,
    input: Arc<dyn ExecutionPlan>,
    predicate: Arc<dyn PhysicalExpr>,
    projection: Option<(Vec<(Arc<dyn PhysicalExpr>, String)>, Arc<Schema>)>,
}

impl GpuFilterExec {
    /// Fails with `NotImplemented` when the predicate is outside the bytecode subset: the planner
    /// then keeps `FilterExec`.
    pub fn try_new(ctx: Arc<GpuContext>, input: Arc<dyn ExecutionPlan>, predicate: Arc<dyn PhysicalExpr>) -> DataFusionResult<Self> {
        to_bytecode(&predicate)?;
        Ok(Self { ctx, input, predicate, projection: None })
    }
    /// Projection(Filter(x)) as one fused launch (physical_planner.rs:52-94 builds the two nodes).
    pub fn with_projection(mut self, exprs: Vec<(Arc<dyn PhysicalExpr>, String)>, schema: Arc<Schema>) -> DataFusionResult<Self> {
        for (e, _) in &exprs { to_bytecode(e)?; }
        self.projection = Some((exprs, schema));
        Ok(self)
    }
}

fn create_fp(ctx: &GpuContext, in_schema: &Schema, pred: Option<&Arc<dyn PhysicalExpr>>,
             proj: &[(Arc<dyn PhysicalExpr>, String)]) -> DataFusionResult<OpHandle> {
    let ffi_schema = FFI_ArrowSchema::try_from(in_schema)?;
    let pred_bc = match pred { Some(p) => to_bytecode(p)?, None => Vec::new() };
    let proj_bc: Vec<Vec<u8>> = proj.iter().map(|(e, _)| to_bytecode(e)).collect::<DataFusionResult<_>>()?;
    let names: Vec<CString> = proj.iter().map(|(_, n)| CString::new(n.as_str()).unwrap()).collect();
    let ptrs: Vec<*const u8> = proj_bc.iter().map(|v| v.as_ptr()).collect();
    let lens: Vec<usize> = proj_bc.iter().map(|v| v.len()).collect();
    let name_ptrs: Vec<*const std::os::raw::c_char> = names.iter().map(|n| n.as_ptr()).collect();
    let mut op = std::ptr::null_mut();
    check(unsafe {
        igl_op_filter_project_create(ctx.0, &ffi_schema,
            if pred.is_some() { pred_bc.as_ptr() } else { std::ptr::null() }, pred_bc.len(),
            ptrs.as_ptr(), lens.as_ptr(), name_ptrs.as_ptr(), proj.len() as i32, &mut op)
    })?;
    Ok(OpHandle(op))
}

#[async_trait::async_trait]
impl ExecutionPlan for GpuFilterExec {
    async fn execute(&self) -> DataFusionResult<BatchStream> {
        let input = self.input.execute().await?; // filter.rs:40
        let proj: &[(Arc<dyn PhysicalExpr>, String)] = self.projection.as_ref().map(|p| p.0.as_slice()).unwrap_or(&[]);
        let op = create_fp(&self.ctx, &self.input.schema(), Some(&self.predicate), proj)?;
        Ok(stream_through(op, input, self.schema())) // 0-row results are never emitted (filter.rs:59-63)
    }
    fn schema(&self) -> Arc<Schema> {
        match &self.projection { Some((_, s)) => s.clone(), None => self.input.schema() } // filter.rs:70-72
    }
}

// ---- Projection (projection.rs:19-23, 48-72) --------------------------------------------------
// no derive(Debug): `dyn ExecutionPlan` has no Debug bound (physical_plan.rs:11)
impl std::fmt::Debug for GpuProjectionExec {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        f.debug_struct("GpuProjectionExec").field("schema", &self.schema).finish_non_exhaustive()
    }
}

#[derive(Clone)]
pub struct GpuProjectionExec {
    ctx: Arc<GpuContext>,
    input: Arc<dyn ExecutionPlan>,
    expressions: Vec<(Arc<dyn PhysicalExpr>, String)>,
    schema: Arc<Schema>,
}

impl GpuProjectionExec {
    pub fn try_new(ctx: Arc<GpuContext>, input: Arc<dyn ExecutionPlan>,
                   expressions: Vec<(Arc<dyn PhysicalExpr>, String)>, schema: Arc<Schema>) -> DataFusionResult<Self> {
        for (e, _) in &expressions { to_bytecode(e)?; }
        Ok(Self { ctx, input, expressions, schema })
    }
}

#[async_trait::async_trait]
impl ExecutionPlan for GpuProjectionExec {
    async fn execute(&self) -> DataFusionResult<BatchStream> {
        let input = self.input.execute().await?;
        let op = create_fp(&self.ctx, &self.input.schema(), None, &self.expressions)?;
        Ok(stream_through(op, input, self.schema.clone())) // every batch is emitted, even 0-row (projection.rs:67)
    }
    fn schema(&self) -> Arc<Schema> { self.schema.clone() }
}

// ---- HashJoin (hash_join.rs:35-76, 100-213) ------------------------------------------------------
// no derive(Debug): `dyn ExecutionPlan` has no Debug bound (physical_plan.rs:11)
impl std::fmt::Debug for GpuHashJoinExec {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result {
        f.debug_struct("GpuHashJoinExec").field("join_type", &self.join_type).field("schema", &self.schema).finish_non_exhaustive()
    }
}

#[derive(Clone)]
pub struct GpuHashJoinExec {
    ctx: Arc<GpuContext>,
    left: Arc<dyn ExecutionPlan>,
    right: Arc<dyn ExecutionPlan>,
    left_keys: Vec<Arc<dyn PhysicalExpr>>,
    right_keys: Vec<Arc<dyn PhysicalExpr>>,
    join_type: JoinType,
    schema: Arc<Schema>,
}

impl GpuHashJoinExec {
    /// Same arguments as `HashJoinExec::new` (hash_join.rs:35-41): one (left, right) key pair per
    /// `join.on` entry, any expressions (igl_op_hash_join_create_ex).  What the library cannot
    /// take (Utf8 columns on a join side, Left joins on wide key lists) comes back as
    /// `NotImplemented` and the planner keeps the CPU join.
    pub fn try_new(ctx: Arc<GpuContext>, left: Arc<dyn ExecutionPlan>, right: Arc<dyn ExecutionPlan>,
                   left_keys: Vec<Arc<dyn PhysicalExpr>>, right_keys: Vec<Arc<dyn PhysicalExpr>>,
                   join_type: JoinType) -> DataFusionResult<Self> {
        if left_keys.is_empty() || left_keys.len() != right_keys.len() {
            return Err(DataFusionError::Plan("join keys come in pairs".into()));
        }
        let me = Self { ctx, left, right, left_keys, right_keys, join_type, schema: Arc::new(Schema::empty()) };
        // the library derives the output schema (left fields, right fields, `right_` prefix on a
        // name clash: hash_join.rs:54-63); creating the operator once here validates keys and types
        let op = me.new_op()?;
        Ok(Self { schema: out_schema(&op)?, ..me })
    }

    fn new_op(&self) -> DataFusionResult<OpHandle> {
        let l = FFI_ArrowSchema::try_from(self.left.schema().as_ref())?;
        let r = FFI_ArrowSchema::try_from(self.right.schema().as_ref())?;
        let lk: Vec<Vec<u8>> = self.left_keys.iter().map(to_bytecode).collect::<DataFusionResult<_>>()?;
        let rk: Vec<Vec<u8>> = self.right_keys.iter().map(to_bytecode).collect::<DataFusionResult<_>>()?;
        let (lp, rp): (Vec<*const u8>, Vec<*const u8>) = (lk.iter().map(|v| v.as_ptr()).collect(), rk.iter().map(|v| v.as_ptr()).collect());
        let (ll, rl): (Vec<usize>, Vec<usize>) = (lk.iter().map(|v| v.len()).collect(), rk.iter().map(|v| v.len()).collect());
        let jt = match self.join_type { JoinType::Inner => 0, JoinType::Left => 1, JoinType::Right => 2, JoinType::Full => 3 };
        let mut op = std::ptr::null_mut();
        check(unsafe {
            igl_op_hash_join_create_ex(self.ctx.0, jt, &l, &r, lp.as_ptr(), ll.as_ptr(), rp.as_ptr(), rl.as_ptr(), lk.len() as i32, &mut op)
        })?;
        Ok(OpHandle(op))
    }
}

#[async_trait::async_trait]
impl ExecutionPlan for GpuHashJoinExec {
    async fn execute(&self) -> DataFusionResult<BatchStream> {
        let op = self.new_op()?;
        // the reference drains the right child inside execute(), before returning (hash_join.rs:100-128)
        let mut right = self.right.execute().await?;
        while let Some(batch) = right.next().await {
            push(&op, &batch?, true)?;
        }
        check(unsafe { igl_op_finish_build(op.0) })?;
        let left = self.left.execute().await?;
        Ok(stream_through(op, left, self.schema.clone()))
    }
    fn schema(&self) -> Arc<Schema> { self.schema.clone() }
}
