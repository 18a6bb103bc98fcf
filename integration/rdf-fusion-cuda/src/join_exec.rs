//! GPU ExecutionPlan nodes for the operators that have no RDF-Fusion-level hook (SURVEY 8b iv): the SPARQL join and the
//! ID-level filter.  Template: KleenePlusClosureExec (lib/physical/src/paths/kleene_plus/physical.rs:50-175).  NOT COMPILED here.
use crate::data_source::{table_stream, GpuQuadPatternDataSource};
use crate::{check, ffi, Gpu, GpuTable};
use datafusion::arrow::datatypes::SchemaRef;
use datafusion::common::exec_err;
use datafusion::datasource::source::DataSourceExec;
use datafusion::error::{DataFusionError, Result as DFResult};
use datafusion::execution::{SendableRecordBatchStream, TaskContext};
use datafusion::physical_expr::{EquivalenceProperties, Partitioning};
use datafusion::physical_plan::execution_plan::{Boundedness, EmissionType};
use datafusion::physical_plan::stream::RecordBatchStreamAdapter;
use datafusion::physical_plan::{DisplayAs, DisplayFormatType, ExecutionPlan, PlanProperties};
use futures::{stream, StreamExt, TryStreamExt};
use std::any::Any;
use std::fmt::Formatter;
use std::ptr;
use std::sync::Arc;

/// A child's result as a device table: GPU children hand over their rfgpu_table (no host round trip), CPU children are
/// collected and imported through the Arrow C Data Interface.
pub(crate) async fn device_table(gpu: Arc<Gpu>, child: Arc<dyn ExecutionPlan>, ctx: Arc<TaskContext>) -> DFResult<GpuTable> {
    if let Some(j) = child.as_any().downcast_ref::<GpuSparqlJoinExec>() {
        return j.run(ctx).await;
    }
    if let Some(f) = child.as_any().downcast_ref::<GpuIdFilterExec>() {
        return f.run(ctx).await;
    }
    if let Some(ds) = child.as_any().downcast_ref::<DataSourceExec>() {
        if let Some(scan) = ds.data_source().as_any().downcast_ref::<GpuQuadPatternDataSource>() {
            return scan.device_table().await;
        }
    }
    let batches = datafusion::physical_plan::collect(child.clone(), ctx).await?;
    let batch = arrow::compute::concat_batches(&child.schema(), &batches)?;
    tokio::task::spawn_blocking(move || GpuTable::import(gpu, &batch)).await.map_err(|e| DataFusionError::External(Box::new(e)))?
}

fn properties(schema: SchemaRef) -> PlanProperties {
    PlanProperties::new(EquivalenceProperties::new(schema), Partitioning::UnknownPartitioning(1), EmissionType::Final, Boundedness::Bounded)
}

/// SparqlJoinNode on the device.  rfgpu_sparql_join makes the lowering decision of SparqlJoinLoweringRule itself
/// (join/rewrite.rs:71-221: cross / equi with NullEqualsNothing / IS_COMPATIBLE + COALESCE) from the tables' schema
/// nullability, so this node replaces HashJoinExec, NestedLoopJoinExec(IS_COMPATIBLE) and CrossJoinExec alike.
#[derive(Debug)]
pub struct GpuSparqlJoinExec {
    pub(crate) gpu: Arc<Gpu>,
    pub(crate) lhs: Arc<dyn ExecutionPlan>,
    pub(crate) rhs: Arc<dyn ExecutionPlan>,
    pub(crate) join_type: i32, // RFGPU_JOIN_INNER | RFGPU_JOIN_LEFT
    pub(crate) filter: Vec<ffi::rfgpu_filter_pred>, // SparqlJoinNode.filter as ID-level predicates over the result row, or empty
    schema: SchemaRef,
    props: PlanProperties,
}

impl GpuSparqlJoinExec {
    pub fn new(gpu: Arc<Gpu>, lhs: Arc<dyn ExecutionPlan>, rhs: Arc<dyn ExecutionPlan>, join_type: i32,
               filter: Vec<ffi::rfgpu_filter_pred>, schema: SchemaRef) -> Self {
        Self { gpu, lhs, rhs, join_type, filter, props: properties(Arc::clone(&schema)), schema }
    }

    pub(crate) async fn run(&self, ctx: Arc<TaskContext>) -> DFResult<GpuTable> {
        let (l, r) = futures::try_join!(
            device_table(Arc::clone(&self.gpu), Arc::clone(&self.lhs), Arc::clone(&ctx)),
            device_table(Arc::clone(&self.gpu), Arc::clone(&self.rhs), ctx)
        )?;
        let (gpu, jt, preds) = (Arc::clone(&self.gpu), self.join_type, self.filter.clone());
        tokio::task::spawn_blocking(move || {
            let mut out = ptr::null_mut();
            let rc = unsafe {
                ffi::rfgpu_sparql_join_filter(gpu.raw(), l.raw, r.raw, jt, preds.as_ptr(), preds.len() as i32, &mut out)
            };
            check(gpu.raw(), rc)?;
            Ok(GpuTable { gpu, raw: out })
        })
        .await
        .map_err(|e| DataFusionError::External(Box::new(e)))?
    }
}

impl DisplayAs for GpuSparqlJoinExec {
    fn fmt_as(&self, _t: DisplayFormatType, f: &mut Formatter) -> std::fmt::Result {
        write!(f, "GpuSparqlJoinExec: join_type={}", if self.join_type == ffi::RFGPU_JOIN_LEFT { "Left" } else { "Inner" })
    }
}

impl ExecutionPlan for GpuSparqlJoinExec {
    fn name(&self) -> &str { "GpuSparqlJoinExec" }
    fn as_any(&self) -> &dyn Any { self }
    fn properties(&self) -> &PlanProperties { &self.props }
    fn children(&self) -> Vec<&Arc<dyn ExecutionPlan>> { vec![&self.lhs, &self.rhs] }
    fn with_new_children(self: Arc<Self>, c: Vec<Arc<dyn ExecutionPlan>>) -> DFResult<Arc<dyn ExecutionPlan>> {
        if c.len() != 2 {
            return exec_err!("GpuSparqlJoinExec takes two children");
        }
        Ok(Arc::new(Self::new(Arc::clone(&self.gpu), Arc::clone(&c[0]), Arc::clone(&c[1]), self.join_type, self.filter.clone(),
                              Arc::clone(&self.schema))))
    }
    fn execute(&self, partition: usize, ctx: Arc<TaskContext>) -> DFResult<SendableRecordBatchStream> {
        if partition != 0 {
            return exec_err!("Only partition 0 is supported for now.");
        }
        let schema = Arc::clone(&self.schema);
        let bs = ctx.session_config().batch_size() as u64;
        let this = Self::new(Arc::clone(&self.gpu), Arc::clone(&self.lhs), Arc::clone(&self.rhs), self.join_type, self.filter.clone(),
                             Arc::clone(&self.schema));
        let batches = stream::once(async move { this.run(ctx).await })
            .map_ok(move |t| table_stream(t, Arc::clone(&schema), bs))
            .try_flatten();
        Ok(Box::pin(RecordBatchStreamAdapter::new(Arc::clone(&self.schema), batches)))
    }
}

/// DataFusion FilterExec over object-id columns (`col = c`, `col != c`, `col_i = col_j`, BOUND / NOT BOUND:
/// lib/logical/src/patterns/rewrite.rs:68-148, expr_builder.rs:58-60,931-959) on the device.
#[derive(Debug)]
pub struct GpuIdFilterExec {
    pub(crate) gpu: Arc<Gpu>,
    pub(crate) input: Arc<dyn ExecutionPlan>,
    pub(crate) preds: Vec<ffi::rfgpu_filter_pred>,
    props: PlanProperties,
}

impl GpuIdFilterExec {
    pub fn new(gpu: Arc<Gpu>, input: Arc<dyn ExecutionPlan>, preds: Vec<ffi::rfgpu_filter_pred>) -> Self {
        let props = properties(input.schema());
        Self { gpu, input, preds, props }
    }
    pub(crate) async fn run(&self, ctx: Arc<TaskContext>) -> DFResult<GpuTable> {
        let t = device_table(Arc::clone(&self.gpu), Arc::clone(&self.input), ctx).await?;
        let (gpu, preds) = (Arc::clone(&self.gpu), self.preds.clone());
        tokio::task::spawn_blocking(move || {
            let mut out = ptr::null_mut();
            check(gpu.raw(), unsafe { ffi::rfgpu_filter(gpu.raw(), t.raw, preds.as_ptr(), preds.len() as i32, &mut out) })?;
            Ok(GpuTable { gpu, raw: out })
        })
        .await
        .map_err(|e| DataFusionError::External(Box::new(e)))?
    }
}

impl DisplayAs for GpuIdFilterExec {
    fn fmt_as(&self, _t: DisplayFormatType, f: &mut Formatter) -> std::fmt::Result {
        write!(f, "GpuIdFilterExec: {} conjunct(s)", self.preds.len())
    }
}

impl ExecutionPlan for GpuIdFilterExec {
    fn name(&self) -> &str { "GpuIdFilterExec" }
    fn as_any(&self) -> &dyn Any { self }
    fn properties(&self) -> &PlanProperties { &self.props }
    fn children(&self) -> Vec<&Arc<dyn ExecutionPlan>> { vec![&self.input] }
    fn with_new_children(self: Arc<Self>, c: Vec<Arc<dyn ExecutionPlan>>) -> DFResult<Arc<dyn ExecutionPlan>> {
        Ok(Arc::new(Self::new(Arc::clone(&self.gpu), Arc::clone(&c[0]), self.preds.clone())))
    }
    fn execute(&self, partition: usize, ctx: Arc<TaskContext>) -> DFResult<SendableRecordBatchStream> {
        if partition != 0 {
            return exec_err!("Only partition 0 is supported for now.");
        }
        let schema = self.input.schema();
        let bs = ctx.session_config().batch_size() as u64;
        let this = Self::new(Arc::clone(&self.gpu), Arc::clone(&self.input), self.preds.clone());
        let out_schema = Arc::clone(&schema);
        let batches = stream::once(async move { this.run(ctx).await })
            .map_ok(move |t| table_stream(t, Arc::clone(&schema), bs))
            .try_flatten();
        Ok(Box::pin(RecordBatchStreamAdapter::new(out_schema, batches)))
    }
}
