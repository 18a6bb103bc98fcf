//! GpuSubtreeRule — a PhysicalOptimizerRule appended to the list built at lib/execution/src/sparql/optimizer.rs:87-91: wherever
//! a HashJoinExec / NestedLoopJoinExec(IS_COMPATIBLE) / CrossJoinExec / FilterExec sits on GPU-capable children, it is replaced
//! by the GPU node, bottom-up, so that whole scan -> join -> ... -> filter subtrees stay device-resident.  NOT COMPILED here.
use crate::data_source::GpuQuadPatternDataSource;
use crate::ffi;
use crate::join_exec::{GpuIdFilterExec, GpuSparqlJoinExec};
use crate::Gpu;
use datafusion::common::tree_node::{Transformed, TreeNode};
use datafusion::common::JoinType;
use datafusion::config::ConfigOptions;
use datafusion::datasource::source::DataSourceExec;
use datafusion::error::Result as DFResult;
use datafusion::logical_expr::Operator;
use datafusion::physical_expr::expressions::{BinaryExpr, Column, IsNotNullExpr, IsNullExpr, Literal};
use datafusion::physical_expr::PhysicalExpr;
use datafusion::physical_optimizer::PhysicalOptimizerRule;
use datafusion::physical_plan::filter::FilterExec;
use datafusion::physical_plan::joins::{CrossJoinExec, HashJoinExec, NestedLoopJoinExec};
use datafusion::physical_plan::ExecutionPlan;
use datafusion::scalar::ScalarValue;
use std::sync::Arc;

#[derive(Debug)]
pub struct GpuSubtreeRule {
    gpu: Arc<Gpu>,
}

impl GpuSubtreeRule {
    pub fn new(gpu: Arc<Gpu>) -> Self {
        Self { gpu }
    }
}

fn is_gpu(plan: &Arc<dyn ExecutionPlan>) -> bool {
    plan.as_any().is::<GpuSparqlJoinExec>()
        || plan.as_any().is::<GpuIdFilterExec>()
        || plan.as_any().downcast_ref::<DataSourceExec>().is_some_and(|d| d.data_source().as_any().is::<GpuQuadPatternDataSource>())
}

/// `column <op> u32 literal` (either side) -> (column name, rfgpu_cmp_op, value): predicate_pushdown.rs:120-157
pub(crate) fn as_column_cmp_u32(e: &Arc<dyn PhysicalExpr>) -> Option<(String, i32, u32)> {
    let b = e.as_any().downcast_ref::<BinaryExpr>()?;
    let op = |o: &Operator, flipped: bool| match (o, flipped) {
        (Operator::Eq, _) => Some(ffi::RFGPU_OP_EQ),
        (Operator::Gt, false) | (Operator::Lt, true) => Some(ffi::RFGPU_OP_GT),
        (Operator::GtEq, false) | (Operator::LtEq, true) => Some(ffi::RFGPU_OP_GTEQ),
        (Operator::Lt, false) | (Operator::Gt, true) => Some(ffi::RFGPU_OP_LT),
        (Operator::LtEq, false) | (Operator::GtEq, true) => Some(ffi::RFGPU_OP_LTEQ),
        _ => None,
    };
    let lit = |e: &Arc<dyn PhysicalExpr>| match e.as_any().downcast_ref::<Literal>()?.value() {
        ScalarValue::UInt32(Some(v)) => Some(*v),
        _ => None,
    };
    let col = |e: &Arc<dyn PhysicalExpr>| Some(e.as_any().downcast_ref::<Column>()?.name().to_owned());
    if let (Some(c), Some(v)) = (col(b.left()), lit(b.right())) {
        return Some((c, op(b.op(), false)?, v));
    }
    let (c, v) = (col(b.right())?, lit(b.left())?);
    Some((c, op(b.op(), true)?, v))
}

/// An ID-level conjunct as rfgpu_filter_pred (`col = c`, `col != c`, `col_i = col_j`, `col_i != col_j`, IS [NOT] NULL);
/// None for value-level expressions (EBV(GT(ENC_TV(..))): comparisons of ENC_TV(column) with constants or columns, joined by
/// AND / OR / NOT, compile to a typed_value::GpuTypedValues::eval program; anything else stays on DataFusion).
fn as_id_pred(e: &Arc<dyn PhysicalExpr>) -> Option<ffi::rfgpu_filter_pred> {
    let mut p = ffi::rfgpu_filter_pred { op: 0, col: 0, col2: 0, value: 0, hi: 0, ids: std::ptr::null(), n_ids: 0 };
    if let Some(n) = e.as_any().downcast_ref::<IsNullExpr>() {
        p.op = ffi::RFGPU_F_IS_NULL;
        p.col = n.arg().as_any().downcast_ref::<Column>()?.index() as i32;
        return Some(p);
    }
    if let Some(n) = e.as_any().downcast_ref::<IsNotNullExpr>() {
        p.op = ffi::RFGPU_F_IS_NOT_NULL;
        p.col = n.arg().as_any().downcast_ref::<Column>()?.index() as i32;
        return Some(p);
    }
    let b = e.as_any().downcast_ref::<BinaryExpr>()?;
    let (eq, ne) = (matches!(b.op(), Operator::Eq), matches!(b.op(), Operator::NotEq));
    if !eq && !ne {
        return None;
    }
    let l = b.left().as_any().downcast_ref::<Column>()?;
    p.col = l.index() as i32;
    if let Some(r) = b.right().as_any().downcast_ref::<Column>() {
        p.op = if eq { ffi::RFGPU_F_EQ_COL } else { ffi::RFGPU_F_NE_COL };
        p.col2 = r.index() as i32;
    } else if let ScalarValue::UInt32(Some(v)) = b.right().as_any().downcast_ref::<Literal>()?.value() {
        p.op = if eq { ffi::RFGPU_F_EQ_CONST } else { ffi::RFGPU_F_NE_CONST };
        p.value = *v;
    } else {
        return None;
    }
    Some(p)
}

impl PhysicalOptimizerRule for GpuSubtreeRule {
    fn optimize(&self, plan: Arc<dyn ExecutionPlan>, _config: &ConfigOptions) -> DFResult<Arc<dyn ExecutionPlan>> {
        let gpu = Arc::clone(&self.gpu);
        plan.transform_up(|node| {
            let any = node.as_any();
            // joins: both children already on the GPU (a CPU child of a large join is imported instead; policy by row estimate)
            let pair = if let Some(j) = any.downcast_ref::<HashJoinExec>() {
                // SparqlJoinLoweringRule only produces Inner / Left with the preserved side on the left (rewrite.rs:356-361);
                // a join that carries a value-level `filter=` (BSBM Explore Q5) keeps its CPU node
                matches!(j.join_type(), JoinType::Inner | JoinType::Left).then(|| (j.left(), j.right(), *j.join_type(), j.filter().is_none()))
            } else if let Some(j) = any.downcast_ref::<NestedLoopJoinExec>() {
                // IS_COMPATIBLE conjunction or keyless Left join: rfgpu_sparql_join re-derives the lowering from the schemas
                matches!(j.join_type(), JoinType::Inner | JoinType::Left).then(|| (j.left(), j.right(), *j.join_type(), true))
            } else if let Some(j) = any.downcast_ref::<CrossJoinExec>() {
                Some((j.left(), j.right(), JoinType::Inner, true))
            } else {
                None
            };
            if let Some((l, r, jt, no_value_filter)) = pair {
                if no_value_filter && is_gpu(l) && is_gpu(r) {
                    let jt = if jt == JoinType::Left { ffi::RFGPU_JOIN_LEFT } else { ffi::RFGPU_JOIN_INNER };
                    let exec = GpuSparqlJoinExec::new(Arc::clone(&gpu), Arc::clone(l), Arc::clone(r), jt, vec![], node.schema());
                    return Ok(Transformed::yes(Arc::new(exec) as Arc<dyn ExecutionPlan>));
                }
            }
            if let Some(f) = any.downcast_ref::<FilterExec>() {
                if is_gpu(f.input()) {
                    let conjuncts = datafusion::physical_expr::split_conjunction(f.predicate());
                    let preds: Option<Vec<_>> = conjuncts.iter().map(|c| as_id_pred(c)).collect();
                    if let Some(preds) = preds {
                        let exec = GpuIdFilterExec::new(Arc::clone(&gpu), Arc::clone(f.input()), preds);
                        return Ok(Transformed::yes(Arc::new(exec) as Arc<dyn ExecutionPlan>));
                    }
                }
            }
            Ok(Transformed::no(node))
        })
        .map(|t| t.data)
    }

    fn name(&self) -> &str {
        "GpuSubtreeRule"
    }

    fn schema_check(&self) -> bool {
        true // the GPU nodes produce exactly node.schema() (check_same_schema in join/rewrite.rs:43)
    }
}
