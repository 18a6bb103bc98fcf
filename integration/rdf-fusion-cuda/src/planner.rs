//! GpuQuadStorePlanner — sibling of MemQuadStorePlanner (lib/storage/src/memory/planner.rs:31-64): plans a QuadPatternNode as
//! a DataSourceExec over the GPU-resident index.  NOT COMPILED here (README.md).
use crate::data_source::GpuQuadPatternDataSource;
use crate::{check, ffi, GpuScanPlan, GpuStore};
use async_trait::async_trait;
use datafusion::datasource::source::DataSourceExec;
use datafusion::error::Result as DFResult;
use datafusion::execution::context::SessionState;
use datafusion::logical_expr::{LogicalPlan, UserDefinedLogicalNode};
use datafusion::physical_plan::empty::EmptyExec;
use datafusion::physical_plan::ExecutionPlan;
use datafusion::physical_planner::{ExtensionPlanner, PhysicalPlanner};
use rdf_fusion_logical::quad_pattern::QuadPatternNode;
use std::ffi::CString;
use std::ptr;
use std::sync::Arc;

/// What the host-side dictionary look-ups (encode_active_graph / encode_triple_pattern, snapshot.rs:186-276) produced for a
/// pattern.  Everything after them — active graph -> level-G instruction, term patterns -> instructions, "unknown constant =>
/// empty", choose_index — happens inside librfgpu (rfgpu_scan_plan_create_pattern).
pub enum EncodedGraph { Default, All, Union(Vec<u32>), AnyNamed }
pub enum EncodedTerm { Variable(String), ObjectId(u32), Unknown }

/// The term dictionary stays on the host (object_id_mapping.rs); the planner only needs these two look-ups from it.
pub trait ObjectIdLookup: Send + Sync {
    fn encode_graph(&self, node: &QuadPatternNode) -> EncodedGraph;
    fn encode_pattern(&self, node: &QuadPatternNode) -> [EncodedTerm; 3];
}

pub struct GpuQuadStorePlanner {
    store: GpuStore, // one retained snapshot for the whole query: all scans see the same data (quad_storage.rs:24-28)
    dictionary: Arc<dyn ObjectIdLookup>,
}

impl GpuQuadStorePlanner {
    pub fn new(store: GpuStore, dictionary: Arc<dyn ObjectIdLookup>) -> Self {
        Self { store, dictionary }
    }

    fn plan_pattern(&self, node: &QuadPatternNode) -> DFResult<GpuScanPlan> {
        let graph_ids;
        let graph = match self.dictionary.encode_graph(node) {
            EncodedGraph::Default => ffi::rfgpu_active_graph { kind: ffi::RFGPU_GRAPH_DEFAULT, graphs: ptr::null(), n_graphs: 0 },
            EncodedGraph::All => ffi::rfgpu_active_graph { kind: ffi::RFGPU_GRAPH_ALL, graphs: ptr::null(), n_graphs: 0 },
            EncodedGraph::AnyNamed => ffi::rfgpu_active_graph { kind: ffi::RFGPU_GRAPH_ANY_NAMED, graphs: ptr::null(), n_graphs: 0 },
            EncodedGraph::Union(ids) => {
                graph_ids = ids;
                ffi::rfgpu_active_graph { kind: ffi::RFGPU_GRAPH_UNION, graphs: graph_ids.as_ptr(), n_graphs: graph_ids.len() as u32 }
            }
        };
        let graph_var = node.graph_variable().map(|v| CString::new(v.as_str()).unwrap());
        let terms = self.dictionary.encode_pattern(node);
        let names: Vec<Option<CString>> = terms
            .iter()
            .map(|t| if let EncodedTerm::Variable(v) = t { Some(CString::new(v.as_str()).unwrap()) } else { None })
            .collect();
        let mut spo = [ffi::rfgpu_term_pattern { kind: ffi::RFGPU_TERM_UNKNOWN, var: ptr::null(), id: 0 }; 3];
        for (k, t) in terms.iter().enumerate() {
            match t {
                EncodedTerm::Variable(_) => {
                    spo[k].kind = ffi::RFGPU_TERM_VARIABLE;
                    spo[k].var = names[k].as_ref().unwrap().as_ptr();
                }
                EncodedTerm::ObjectId(id) => {
                    spo[k].kind = ffi::RFGPU_TERM_ID;
                    spo[k].id = *id;
                }
                EncodedTerm::Unknown => {}
            }
        }
        let (mut raw, mut is_empty) = (ptr::null_mut(), 0i32);
        let ctx = self.store.gpu.raw();
        check(ctx, unsafe {
            ffi::rfgpu_scan_plan_create_pattern(
                ctx, self.store.raw, &graph, graph_var.as_ref().map_or(ptr::null(), |c| c.as_ptr()), spo.as_ptr(), &mut raw,
                &mut is_empty,
            )
        })?;
        Ok(GpuScanPlan { store: self.store.clone(), raw, is_empty: is_empty != 0 })
    }
}

#[async_trait]
impl ExtensionPlanner for GpuQuadStorePlanner {
    async fn plan_extension(
        &self,
        _planner: &dyn PhysicalPlanner,
        node: &dyn UserDefinedLogicalNode,
        _logical_inputs: &[&LogicalPlan],
        _physical_inputs: &[Arc<dyn ExecutionPlan>],
        _session_state: &SessionState,
    ) -> DFResult<Option<Arc<dyn ExecutionPlan>>> {
        let Some(node) = node.as_any().downcast_ref::<QuadPatternNode>() else {
            return Ok(None); // foreign node: another planner's business (planner.rs:61-63)
        };
        let schema = Arc::clone(node.schema().inner());
        let plan = self.plan_pattern(node)?;
        if plan.is_empty {
            // PlanPatternScanResult::Empty (snapshot.rs:101-110)
            return Ok(Some(Arc::new(EmptyExec::new(schema))));
        }
        Ok(Some(DataSourceExec::from_data_source(GpuQuadPatternDataSource::new(schema, Arc::new(plan)))))
    }
}
