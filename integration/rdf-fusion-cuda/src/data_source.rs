//! GpuQuadPatternDataSource — sibling of MemQuadPatternDataSource (pattern_data_source.rs:21-152).  NOT COMPILED here.
use crate::{GpuScanPlan, GpuTable};
use arrow::array::RecordBatch;
use datafusion::arrow::datatypes::SchemaRef;
use datafusion::common::{exec_err, Statistics};
use datafusion::config::ConfigOptions;
use datafusion::datasource::source::DataSource;
use datafusion::error::{DataFusionError, Result as DFResult};
use datafusion::execution::{SendableRecordBatchStream, TaskContext};
use datafusion::physical_expr::{EquivalenceProperties, Partitioning, PhysicalExpr};
use datafusion::physical_plan::execution_plan::SchedulingType;
use datafusion::physical_plan::filter_pushdown::{FilterPushdownPropagation, PushedDown};
use datafusion::physical_plan::metrics::ExecutionPlanMetricsSet;
use datafusion::physical_plan::stream::RecordBatchStreamAdapter;
use datafusion::physical_plan::DisplayFormatType;
use futures::{stream, StreamExt, TryStreamExt};
use std::any::Any;
use std::fmt::Formatter;
use std::sync::Arc;

#[derive(Clone)]
pub struct GpuQuadPatternDataSource {
    schema: SchemaRef,
    metrics: ExecutionPlanMetricsSet,
    pub(crate) plan: Arc<GpuScanPlan>,
}

impl std::fmt::Debug for GpuQuadPatternDataSource {
    fn fmt(&self, f: &mut Formatter<'_>) -> std::fmt::Result {
        write!(f, "GpuQuadPatternDataSource: {}", self.plan.describe())
    }
}

impl GpuQuadPatternDataSource {
    pub fn new(schema: SchemaRef, plan: Arc<GpuScanPlan>) -> Self {
        Self { schema, metrics: ExecutionPlanMetricsSet::default(), plan }
    }

    /// The device-resident result, for a GPU parent (GpuSparqlJoinExec, GpuIdFilterExec): no Arrow export at all.
    pub async fn device_table(&self) -> DFResult<GpuTable> {
        let plan = Arc::clone(&self.plan);
        // the C call synchronises its stream: never block a tokio worker on it (SURVEY 8b "threading")
        tokio::task::spawn_blocking(move || plan.execute(true)).await.map_err(|e| DataFusionError::External(Box::new(e)))?
    }
}

/// rows of a device table as RecordBatches of `batch_size` rows (the reference emits one batch per row group of <= 8192 rows,
/// scan.rs:135-205; empty batches are skipped)
pub(crate) fn table_stream(table: GpuTable, schema: SchemaRef, batch_size: u64) -> impl stream::Stream<Item = DFResult<RecordBatch>> {
    let n = table.num_rows();
    let table = Arc::new(table);
    stream::iter((0..n).step_by(batch_size.max(1) as usize)).then(move |off| {
        let (table, schema) = (Arc::clone(&table), Arc::clone(&schema));
        async move {
            tokio::task::spawn_blocking(move || table.export(&schema, off, batch_size.min(n - off)))
                .await
                .map_err(|e| DataFusionError::External(Box::new(e)))?
        }
    })
}

impl DataSource for GpuQuadPatternDataSource {
    fn open(&self, partition: usize, context: Arc<TaskContext>) -> DFResult<SendableRecordBatchStream> {
        if partition != 0 {
            return exec_err!("Only partition 0 is supported for now."); // pattern_data_source.rs:47-49
        }
        let (this, schema) = (self.clone(), Arc::clone(&self.schema));
        let batch_size = context.session_config().batch_size() as u64;
        let batches = stream::once(async move { this.device_table().await })
            .map_ok(move |t| table_stream(t, Arc::clone(&schema), batch_size))
            .try_flatten();
        Ok(Box::pin(RecordBatchStreamAdapter::new(Arc::clone(&self.schema), batches)))
    }

    fn as_any(&self) -> &dyn Any {
        self
    }

    fn fmt_as(&self, _t: DisplayFormatType, f: &mut Formatter) -> std::fmt::Result {
        write!(f, "{}", self.plan.describe()) // same text as the reference's fmt_as (:89-104): plan snapshots stay comparable
    }

    fn output_partitioning(&self) -> Partitioning {
        Partitioning::UnknownPartitioning(1) // :68-70
    }

    fn eq_properties(&self) -> EquivalenceProperties {
        EquivalenceProperties::new(Arc::clone(&self.schema))
    }

    fn scheduling_type(&self) -> SchedulingType {
        SchedulingType::Cooperative // :76-78
    }

    fn partition_statistics(&self, _partition: Option<usize>) -> DFResult<Statistics> {
        Ok(Statistics::new_unknown(&self.schema))
    }

    fn with_fetch(&self, _limit: Option<usize>) -> Option<Arc<dyn DataSource>> {
        None
    }

    fn fetch(&self) -> Option<usize> {
        None
    }

    fn metrics(&self) -> ExecutionPlanMetricsSet {
        self.metrics.clone()
    }

    /// predicate_pushdown.rs:120-249 stays as it is in Rust (it walks DataFusion expressions); each `column <op> u32` it
    /// recognises becomes rfgpu_scan_plan_apply_filter, which re-chooses the index (scan.rs:469-486).  A filter that is not
    /// representable is reported PushedDown::No and stays in the FilterExec above (-> GpuIdFilterExec if it is ID-level).
    fn try_pushdown_filters(
        &self,
        filters: Vec<Arc<dyn PhysicalExpr>>,
        _config: &ConfigOptions,
    ) -> DFResult<FilterPushdownPropagation<Arc<dyn DataSource>>> {
        let mut plan = GpuScanPlan { store: self.plan.store.clone(), raw: clone_plan(&self.plan)?, is_empty: self.plan.is_empty };
        let mut support = Vec::with_capacity(filters.len());
        for f in &filters {
            match crate::optimizer::as_column_cmp_u32(f) {
                Some((column, op, value)) if plan.apply_filter(&column, op, value).is_ok() => support.push(PushedDown::Yes),
                _ => support.push(PushedDown::No),
            }
        }
        let updated: Arc<dyn DataSource> = Arc::new(Self::new(Arc::clone(&self.schema), Arc::new(plan)));
        Ok(FilterPushdownPropagation::with_parent_pushdown_result(support).with_updated_node(updated))
    }
}

/// a second plan over the same store and instructions (plans are cheap host objects)
fn clone_plan(_p: &GpuScanPlan) -> DFResult<*mut crate::ffi::rfgpu_scan_plan> {
    // rebuilt from the QuadPatternNode kept by the planner in the real crate; elided in this uncompiled mirror
    datafusion::common::not_impl_err!("clone_plan: rebuild through GpuQuadStorePlanner::plan_pattern")
}
