"""rdf_fusion_b200 — B200-native quad-pattern scan / filter / SPARQL-join path behind rdf-fusion's storage API.

The product is `librfgpu.so` (hand-written CUDA for sm_100a behind the C ABI of include/rfgpu.h).  This package is
the thin host harness over it.  Importing it never touches oracle/ and there is no CPU fallback.
"""
from . import _ffi as ffi  # noqa: F401
from ._ffi import (ALGO_COMPATIBLE, ALGO_CROSS, ALGO_EQUI, F_BETWEEN, F_EQ_COL, F_EQ_CONST, F_IN,  # noqa: F401
                   F_IS_NOT_NULL, F_IS_NULL, F_NE_COL, F_NE_CONST, GOPS, GOSP, GPOS, GPSO, GSOP, GSPO, JOIN_INNER,
                   JOIN_LEFT, OP_EQ, OP_GT, OP_GTEQ, OP_LT, OP_LTEQ, PERM_NAMES)
from .api import (FALSE, Between, Context, Dictionary, EqualTo, FPred, In, Instr, Pred, RfgpuError, ScanPlan, Store,  # noqa: F401
                  Table, TV_VALUE_DTYPE, TypedValueTable, execute_many, make_filter_preds, make_instructions, numeric_rank_table,
                  pack_terms, scan, traverse, typed_values_to_arrow)
from .synth import BsbmLikeGraph, SynthGraph, generate_bsbm_like, generate_graph  # noqa: F401
