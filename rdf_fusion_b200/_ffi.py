"""ctypes binding of include/rfgpu.h (librfgpu.so).

This is the same surface a Rust `extern "C"` block would declare.  There is no fallback:
if the CUDA library is missing the import raises, and if no GPU is present `rfgpu_ctx_create` fails with
RFGPU_E_NOGPU.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librfgpu.so")

OK, E_INVALID, E_CUDA, E_EXEC, E_UNSUPPORTED, E_NOGPU = 0, -1, -2, -3, -4, -5

G, S, P, O = 0, 1, 2, 3


def perm(a, b, c, d):
    return a | (b << 2) | (c << 4) | (d << 6)


GSPO, GPOS, GOSP = perm(G, S, P, O), perm(G, P, O, S), perm(G, O, S, P)
GPSO, GSOP, GOPS = perm(G, P, S, O), perm(G, S, O, P), perm(G, O, P, S)
PERM_NAMES = {GSPO: "GSPO", GPOS: "GPOS", GOSP: "GOSP", GPSO: "GPSO", GSOP: "GSOP", GOPS: "GOPS"}

PRED_NONE, PRED_FALSE, PRED_IN, PRED_BETWEEN, PRED_EQUAL_TO = 0, 1, 2, 3, 4
TRAVERSE, SCAN = 0, 1
OP_EQ, OP_GT, OP_GTEQ, OP_LT, OP_LTEQ = 0, 1, 2, 3, 4
F_EQ_CONST, F_NE_CONST, F_EQ_COL, F_NE_COL, F_BETWEEN, F_IN, F_IS_NULL, F_IS_NOT_NULL = range(8)
JOIN_INNER, JOIN_LEFT = 0, 1
ALGO_CROSS, ALGO_EQUI, ALGO_COMPATIBLE = 0, 1, 2
SCAN_ALLOW_VIEW = 1
PHYS_NONE, PHYS_HASH, PHYS_MERGE, PHYS_RADIX, PHYS_SORT_MERGE, PHYS_LOOKUP = 0, 1, 2, 3, 4, 5
AGG_PHYS_NONE, AGG_PHYS_HASH, AGG_PHYS_SORT = 0, 1, 2
OPT_MERGE_JOIN = 1
OPT_MERGE_THREADS = 2
OPT_RADIX_JOIN = 3
OPT_RADIX_MIN_ROWS = 4
OPT_HASH_L2_BYTES = 5
OPT_SHUFFLE_P2P = 6
OPT_SHUFFLE_ARENA_BYTES = 7
OPT_SORT_MERGE_JOIN = 8
OPT_SORT_SEMI = 9
OPT_SORT_AGG = 10
OPT_LOOKUP_JOIN = 11
OPT_LAZY_ROWS = 12
OPT_SHUFFLE_BULK = 13
SHUFFLE_LOCAL, SHUFFLE_NCCL, SHUFFLE_P2P = 0, 1, 2
UNIQUE_ID_BYTES = 128

u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)
u64p = C.POINTER(C.c_uint64)


class ScanPred(C.Structure):
    _fields_ = [("kind", C.c_int32), ("ids", u32p), ("n_ids", C.c_uint32), ("lo", C.c_uint32), ("hi", C.c_uint32),
                ("equal_to", C.c_char_p)]


class ScanInstruction(C.Structure):
    _fields_ = [("action", C.c_int32), ("var", C.c_char_p), ("pred", ScanPred)]


class ActiveGraph(C.Structure):
    _fields_ = [("kind", C.c_int32), ("graphs", u32p), ("n_graphs", C.c_uint32)]


class TermPattern(C.Structure):
    _fields_ = [("kind", C.c_int32), ("var", C.c_char_p), ("id", C.c_uint32)]


GRAPH_DEFAULT, GRAPH_ALL, GRAPH_UNION, GRAPH_ANY_NAMED = 0, 1, 2, 3
TERM_VARIABLE, TERM_ID, TERM_UNKNOWN = 0, 1, 2


class FilterPred(C.Structure):
    _fields_ = [("op", C.c_int32), ("col", C.c_int32), ("col2", C.c_int32), ("value", C.c_uint32), ("hi", C.c_uint32),
                ("ids", u32p), ("n_ids", C.c_uint32)]


class SynthDesc(C.Structure):
    _fields_ = [("n_rows", C.c_uint64), ("n_predicates", C.c_uint32), ("seed", C.c_uint64), ("pred_cdf", C.POINTER(C.c_double)),
                ("obj_cdf", C.POINTER(C.c_double)), ("named_graph_frac", C.c_double), ("keep", C.c_int32), ("rank", C.c_uint32),
                ("n_ranks", C.c_uint32), ("perm", C.c_uint32), ("lo", C.c_uint32 * 4), ("hi", C.c_uint32 * 4),
                ("has_lo", C.c_int32), ("has_hi", C.c_int32)]


PLACE_NONE, PLACE_JOINED_COMPONENT, PLACE_SUBJECT = 0, 1, 2
SYNTH_KEEP_ALL, SYNTH_KEEP_HASH_SUBJECT, SYNTH_KEEP_RANGE, SYNTH_KEEP_HASH_OBJECT = 0, 1, 2, 3


# typed values (ENC_TV): rfgpu_tv_field / rfgpu_tv_op and the structs of include/rfgpu.h
(TV_NULL, TV_NAMED_NODE, TV_BLANK_NODE, TV_STRING, TV_BOOLEAN, TV_FLOAT, TV_DOUBLE, TV_DECIMAL, TV_INT, TV_INTEGER, TV_DATE_TIME,
 TV_TIME, TV_DATE, TV_DURATION, TV_OTHER_LITERAL) = range(15)
TV_N_FIELDS = 15
TV_HAS_AUX, TV_HAS_VALUE = 1, 2
TV_EQ, TV_LT, TV_LE, TV_GT, TV_GE, TV_AND, TV_OR, TV_NOT, TV_BOUND = range(9)


class TvValue(C.Structure):
    _fields_ = [("lo", C.c_uint64), ("hi", C.c_uint64), ("aux", C.c_int64), ("field", C.c_uint8), ("flags", C.c_uint8),
                ("reserved", C.c_uint8 * 6)]


class TvInstr(C.Structure):
    _fields_ = [("op", C.c_int32), ("lhs", C.c_int32), ("rhs", C.c_int32)]


class TvConst(C.Structure):
    _fields_ = [("v", TvValue), ("str", C.c_char_p), ("str_len", C.c_uint32), ("str2", C.c_char_p), ("str2_len", C.c_uint32)]


class TvExport(C.Structure):
    _fields_ = [("type_ids", C.c_void_p), ("offsets", C.c_void_p), ("boolean_bits", C.c_void_p), ("floats", C.c_void_p),
                ("doubles", C.c_void_p), ("decimals", C.c_void_p), ("ints", C.c_void_p), ("integers", C.c_void_p),
                ("ts_value", C.c_void_p * 3), ("ts_offset", C.c_void_p * 3), ("ts_offset_valid", C.c_void_p * 3),
                ("dur_months", C.c_void_p), ("dur_months_valid", C.c_void_p), ("dur_seconds", C.c_void_p),
                ("dur_seconds_valid", C.c_void_p)]


class Metrics(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "kernel_launches", "h2d_bytes", "d2h_bytes", "scan_rows_in", "scan_rows_out", "filter_rows_in",
        "filter_rows_out", "join_build_rows", "join_probe_rows", "join_rows_out", "algorithmic_bytes")]


# name -> (restype, argtypes); every symbol include/rfgpu.h declares
SIGNATURES = {
    "rfgpu_ctx_create": (C.c_int32, [C.c_int32, C.POINTER(C.c_void_p)]),
    "rfgpu_ctx_destroy": (None, [C.c_void_p]),
    "rfgpu_last_error": (C.c_char_p, [C.c_void_p]),
    "rfgpu_ctx_stream": (C.c_void_p, [C.c_void_p]),
    "rfgpu_ctx_synchronize": (C.c_int32, [C.c_void_p]),
    "rfgpu_version": (C.c_char_p, []),
    "rfgpu_store_create": (C.c_int32, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "rfgpu_store_retain": (C.c_void_p, [C.c_void_p]),
    "rfgpu_store_release": (None, [C.c_void_p]),
    "rfgpu_store_add_index": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p), C.c_uint64, C.c_int32]),
    "rfgpu_store_len": (C.c_uint64, [C.c_void_p]),
    "rfgpu_store_num_indexes": (C.c_int32, [C.c_void_p]),
    "rfgpu_store_set_sharded": (C.c_int32, [C.c_void_p, C.c_int32]),
    "rfgpu_store_index_rows": (C.c_uint64, [C.c_void_p, C.c_int32]),
    "rfgpu_store_set_placement": (C.c_int32, [C.c_void_p, C.c_int32, C.c_uint32, C.c_uint32]),
    "rfgpu_synth_quads": (C.c_int32, [C.c_void_p, C.POINTER(SynthDesc), C.POINTER(C.c_void_p)]),
    "rfgpu_store_index_perm": (C.c_uint32, [C.c_void_p, C.c_int32]),
    "rfgpu_scan_plan_create": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(ScanInstruction), C.POINTER(C.c_void_p)]),
    "rfgpu_scan_plan_create_pattern": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(ActiveGraph), C.c_char_p,
                                                   C.POINTER(TermPattern), C.POINTER(C.c_void_p), C.POINTER(C.c_int32)]),
    "rfgpu_scan_plan_destroy": (None, [C.c_void_p]),
    "rfgpu_scan_plan_apply_filter": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_char_p, C.c_int32, C.c_uint32]),
    "rfgpu_scan_plan_apply_between": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_char_p, C.c_uint32, C.c_uint32]),
    "rfgpu_scan_plan_set_index": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint32]),
    "rfgpu_scan_plan_index": (C.c_uint32, [C.c_void_p]),
    "rfgpu_scan_plan_describe": (C.c_size_t, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "rfgpu_scan_execute": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p)]),
    "rfgpu_scan_execute_many": (C.c_int32, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int32, C.c_uint32, C.POINTER(C.c_void_p)]),
    "rfgpu_scan_last_range": (C.c_int32, [C.c_void_p, u64p, u64p]),
    "rfgpu_table_retain": (C.c_void_p, [C.c_void_p]),
    "rfgpu_table_release": (None, [C.c_void_p]),
    "rfgpu_table_num_rows": (C.c_uint64, [C.c_void_p]),
    "rfgpu_table_num_cols": (C.c_int32, [C.c_void_p]),
    "rfgpu_table_settle": (C.c_int32, [C.c_void_p, C.c_void_p]),
    "rfgpu_table_is_pending": (C.c_int32, [C.c_void_p]),
    "rfgpu_ctx_host_syncs": (C.c_uint64, [C.c_void_p]),
    "rfgpu_ctx_alloc_stats": (C.c_int32, [C.c_void_p, C.POINTER(C.c_double), C.c_int32]),
    "rfgpu_ctx_lazy_redos": (C.c_uint64, [C.c_void_p]),
    "rfgpu_table_col_name": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "rfgpu_table_col_nullable": (C.c_int32, [C.c_void_p, C.c_int32]),
    "rfgpu_table_col_device_ptr": (C.c_void_p, [C.c_void_p, C.c_int32]),
    "rfgpu_table_import": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), u8p, C.POINTER(C.c_void_p),
                                       C.POINTER(C.c_void_p), C.c_uint64, C.POINTER(C.c_void_p)]),
    "rfgpu_table_export": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_void_p),
                                       C.POINTER(C.c_void_p), u64p]),
    "rfgpu_table_col_min_max": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), u64p]),
    "rfgpu_table_export_async": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_void_p)]),
    "rfgpu_table_project": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_char_p),
                                        C.POINTER(C.c_void_p)]),
    "rfgpu_table_set_nullable": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32]),
    "rfgpu_table_import_arrow": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]),
    "rfgpu_table_export_arrow": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]),
    "rfgpu_filter": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(FilterPred), C.c_int32, C.POINTER(C.c_void_p)]),
    "rfgpu_filter_noalloc": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.c_uint64, C.POINTER(FilterPred),
                                         C.c_int32, C.POINTER(C.c_void_p), C.c_void_p]),
    "rfgpu_sparql_join": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p)]),
    "rfgpu_group_count": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_char_p,
                                      C.POINTER(C.c_void_p)]),
    "rfgpu_distinct": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]),
    "rfgpu_last_join_algo": (C.c_int32, [C.c_void_p]),
    "rfgpu_sparql_minus": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]),
    "rfgpu_sparql_exists": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p)]),
    "rfgpu_last_join_phys": (C.c_int32, [C.c_void_p]),
    "rfgpu_ctx_set_option": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int64]),
    "rfgpu_table_sorted_col": (C.c_int32, [C.c_void_p]),
    "rfgpu_table_set_sorted_col": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32]),
    "rfgpu_scan_plan_order_by": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_char_p, C.POINTER(C.c_int32)]),
    "rfgpu_comm_unique_id": (C.c_int32, [u8p]),
    "rfgpu_comm_init": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32, u8p]),
    "rfgpu_comm_rank": (C.c_int32, [C.c_void_p]),
    "rfgpu_comm_size": (C.c_int32, [C.c_void_p]),
    "rfgpu_shuffle": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_void_p)]),
    "rfgpu_shuffle_dest": (C.c_uint32, [C.c_uint32, C.c_uint32]),
    "rfgpu_last_shuffle_path": (C.c_int32, [C.c_void_p]),
    "rfgpu_partition": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_uint32, C.POINTER(C.c_void_p), u64p]),
    "rfgpu_metrics_get": (C.c_int32, [C.c_void_p, C.POINTER(Metrics)]),
    "rfgpu_metrics_reset": (C.c_int32, [C.c_void_p]),
    "rfgpu_profile_enable": (C.c_int32, [C.c_void_p, C.c_int32]),
    "rfgpu_profile_num_kernels": (C.c_int32, [C.c_void_p]),
    "rfgpu_profile_get": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), u64p, C.POINTER(C.c_double), u64p]),
    "rfgpu_profile_gap_ms": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_double)]),
    "rfgpu_profile_timeline": (C.c_int32, [C.c_void_p, C.c_int32]),
    "rfgpu_profile_timeline_len": (C.c_int32, [C.c_void_p]),
    "rfgpu_profile_timeline_get": (C.c_int32, [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                                C.POINTER(C.c_double)]),
    "rfgpu_timer_start": (C.c_int32, [C.c_void_p]),
    "rfgpu_timer_stop": (C.c_int32, [C.c_void_p, C.POINTER(C.c_double)]),
    "rfgpu_flush_l2": (C.c_int32, [C.c_void_p]),
    "rfgpu_last_agg_phys": (C.c_int32, [C.c_void_p]),
    "rfgpu_table_lookup": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_uint32, C.c_int32, C.c_char_p,
                                       C.POINTER(C.c_void_p)]),
    "rfgpu_sparql_join_filter": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(FilterPred), C.c_int32,
                                             C.POINTER(C.c_void_p)]),
    "rfgpu_dict_create": (C.c_int32, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]),
    "rfgpu_dict_retain": (C.c_void_p, [C.c_void_p]),
    "rfgpu_dict_release": (None, [C.c_void_p]),
    "rfgpu_dict_len": (C.c_uint32, [C.c_void_p]),
    "rfgpu_dict_decode": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_uint64, C.c_uint64,
                                      C.POINTER(C.c_void_p)]),
    "rfgpu_plain_terms_release": (None, [C.c_void_p]),
    "rfgpu_plain_terms_sizes": (C.c_int32, [C.c_void_p, u64p, u64p, u64p]),
    "rfgpu_plain_terms_export": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                             C.POINTER(C.c_void_p)]),
    "rfgpu_tv_table_create": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p)]),
    "rfgpu_tv_table_retain": (C.c_void_p, [C.c_void_p]),
    "rfgpu_tv_table_release": (None, [C.c_void_p]),
    "rfgpu_tv_decode": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_uint64, C.c_uint64,
                                    C.POINTER(C.c_void_p)]),
    "rfgpu_typed_values_release": (None, [C.c_void_p]),
    "rfgpu_typed_values_sizes": (C.c_int32, [C.c_void_p, u64p, u64p]),
    "rfgpu_typed_values_strings": (C.c_void_p, [C.c_void_p, C.c_int32]),
    "rfgpu_typed_values_export": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(TvExport)]),
    "rfgpu_tv_eval": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(TvInstr), C.c_int32, C.POINTER(TvConst),
                                  C.c_int32, C.c_char_p, C.POINTER(C.c_void_p)]),
    "rfgpu_table_sort": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_void_p)]),
    "rfgpu_store_build": (C.c_int32, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_uint64, C.c_int32, u32p,
                                      C.c_int32]),
}

_lib = None


def load():
    """Load librfgpu.so; raises if it has not been built (python __graft_entry__.py or make -C rdf_fusion_b200/csrc)."""
    global _lib
    if _lib is not None:
        return _lib
    global LIB_PATH
    LIB_PATH = os.environ.get("RFGPU_LIB", LIB_PATH)  # kernel-variant experiments load another build of the same library
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the CUDA library first (make -C rdf_fusion_b200/csrc). "
            "rdf_fusion_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
