/*
 * rfgpu.h — C ABI of the B200-native quad-pattern scan / filter / SPARQL-join path.
 *
 * This is the drop-in boundary a Rust `rdf-fusion-cuda` crate binds with `extern "C"`
 * (see INTEGRATION.md for the bindgen-style stub).  Every entry point cites the
 * reference interface (tobixdev/rdf-fusion, paths relative to the reference root) it
 * replaces.  No torch / C++ types cross this boundary: plain pointers and sizes only.
 *
 * Conventions
 *  - every function returning int32_t returns RFGPU_OK (0) or a negative RFGPU_E_* code;
 *    the message is available through rfgpu_last_error(ctx) (thread-local copy when ctx
 *    is NULL).  No exceptions or longjmp cross the boundary (DFResult<T> on the Rust side:
 *    lib/storage/src/memory/storage/pattern_data_source.rs:47-55).
 *  - object ids are uint32_t (lib/encoding/src/object_id/encoding.rs:131-133).  Id 0 is
 *    DEFAULT_GRAPH_ID and is surfaced as Arrow NULL (lib/storage/src/memory/object_id.rs:21-27,
 *    quad_index_data.rs:437-440).  On the device 0 *is* the NULL sentinel for every column;
 *    validity bitmaps are synthesised on export and folded in on import.
 *  - a ctx owns one CUDA device + one stream; calls on one ctx must be serialised by the
 *    caller (one ctx per tokio worker / per DataFusion partition); stores and tables are
 *    reference counted and may be shared between ctxs of the same device.
 */
#ifndef RFGPU_H
#define RFGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RFGPU_OK 0
#define RFGPU_E_INVALID (-1)     /* plan_err!: malformed argument / plan            */
#define RFGPU_E_CUDA (-2)        /* CUDA runtime failure (message has the cudaError) */
#define RFGPU_E_EXEC (-3)        /* exec_err!: runtime failure of a valid plan       */
#define RFGPU_E_UNSUPPORTED (-4) /* shape not handled by the GPU node: keep the CPU node */
#define RFGPU_E_NOGPU (-5)       /* no usable CUDA device: there is no CPU fallback  */

typedef struct rfgpu_ctx rfgpu_ctx;
typedef struct rfgpu_store rfgpu_store; /* IndexPermutations<MemQuadIndex> on the device */
typedef struct rfgpu_table rfgpu_table; /* a device-resident set of solution mappings    */
typedef struct rfgpu_scan_plan rfgpu_scan_plan; /* PlannedPatternScan                     */

/* ---- index components / permutations: lib/storage/src/index/components.rs:62-84 ---- */
enum { RFGPU_G = 0, RFGPU_S = 1, RFGPU_P = 2, RFGPU_O = 3 };
/* a permutation is 4 components packed 2 bits each, level 0 in the low bits */
#define RFGPU_PERM(a, b, c, d) ((uint32_t)((a) | ((b) << 2) | ((c) << 4) | ((d) << 6)))
#define RFGPU_GSPO RFGPU_PERM(RFGPU_G, RFGPU_S, RFGPU_P, RFGPU_O)
#define RFGPU_GPOS RFGPU_PERM(RFGPU_G, RFGPU_P, RFGPU_O, RFGPU_S)
#define RFGPU_GOSP RFGPU_PERM(RFGPU_G, RFGPU_O, RFGPU_S, RFGPU_P)
/* extra G-leading orders the reference does not keep but the store may */
#define RFGPU_GPSO RFGPU_PERM(RFGPU_G, RFGPU_P, RFGPU_S, RFGPU_O)
#define RFGPU_GSOP RFGPU_PERM(RFGPU_G, RFGPU_S, RFGPU_O, RFGPU_P)
#define RFGPU_GOPS RFGPU_PERM(RFGPU_G, RFGPU_O, RFGPU_P, RFGPU_S)

/* ---- context ---- */
/* One CUDA device + stream + stream-ordered memory pool.  Fails with RFGPU_E_NOGPU when
 * the device cannot be opened: the product path has no CPU fallback. */
int32_t rfgpu_ctx_create(int32_t device, rfgpu_ctx** out);
void rfgpu_ctx_destroy(rfgpu_ctx* ctx);
const char* rfgpu_last_error(const rfgpu_ctx* ctx);
/* the ctx's cudaStream_t (as void*), so a host can order its own work / events on it */
void* rfgpu_ctx_stream(rfgpu_ctx* ctx);
int32_t rfgpu_ctx_synchronize(rfgpu_ctx* ctx);
const char* rfgpu_version(void);

/* ---- scan instructions: lib/storage/src/memory/storage/scan_instructions.rs ---- */
typedef enum {
  RFGPU_PRED_NONE = 0,
  RFGPU_PRED_FALSE = 1,    /* MemIndexScanPredicate::False   :159 */
  RFGPU_PRED_IN = 2,       /* MemIndexScanPredicate::In      :161 */
  RFGPU_PRED_BETWEEN = 3,  /* MemIndexScanPredicate::Between :163 (inclusive) */
  RFGPU_PRED_EQUAL_TO = 4  /* MemIndexScanPredicate::EqualTo :165 (repeated variable) */
} rfgpu_pred_kind;

typedef struct {
  int32_t kind;           /* rfgpu_pred_kind */
  const uint32_t* ids;    /* IN: the id set (any order, duplicates tolerated) */
  uint32_t n_ids;
  uint32_t lo, hi;        /* BETWEEN: inclusive bounds */
  const char* equal_to;   /* EQUAL_TO: variable name scanned at another level */
} rfgpu_scan_pred;

typedef enum { RFGPU_TRAVERSE = 0, RFGPU_SCAN = 1 } rfgpu_scan_action;

typedef struct {
  int32_t action;         /* Traverse(pred) | Scan(var, pred) — scan_instructions.rs:247-252 */
  const char* var;        /* SCAN: variable (= output column) name */
  rfgpu_scan_pred pred;
} rfgpu_scan_instruction;

/* comparison operators of pushed-down filters: predicate_pushdown.rs:24-58 */
typedef enum { RFGPU_OP_EQ = 0, RFGPU_OP_GT = 1, RFGPU_OP_GTEQ = 2, RFGPU_OP_LT = 3, RFGPU_OP_LTEQ = 4 } rfgpu_cmp_op;

/* ---- store: IndexPermutations (lib/storage/src/index/permutations.rs:54-128) ---- */
int32_t rfgpu_store_create(rfgpu_ctx* ctx, rfgpu_store** out);
rfgpu_store* rfgpu_store_retain(rfgpu_store* s); /* one read snapshot per plan: snapshot.rs:35-37 */
void rfgpu_store_release(rfgpu_store* s);
/* Adds one permutation from columns already sorted (and de-duplicated) in `perm` order;
 * cols[i] is the column of level i.  `on_device` != 0: the pointers are device pointers on
 * the ctx's device and are copied device-to-device.  Sortedness/uniqueness is verified by a
 * kernel; violation → RFGPU_E_INVALID.  Replaces MemQuadIndex::insert (quad_index.rs:132-143)
 * + MemIndexData::insert (quad_index_data.rs:287-332) for bulk load. */
int32_t rfgpu_store_add_index(rfgpu_ctx* ctx, rfgpu_store* s, uint32_t perm,
                              const uint32_t* const cols[4], uint64_t n_rows, int32_t on_device);
/* Sharded store (multi-GPU; north_star / SURVEY 8e): every permutation of this store is ONE RANK'S SHARD of the global
 * permutation, and the permutations are sharded independently — a contiguous row slice of the globally sorted permutation
 * ("the index is range-partitioned by leading key"), or the quads whose subject (GSPO, GPSO, GSOP) / object (GPOS, GOSP,
 * GOPS) hashes to the rank, so that a scan comes out already partitioned on the variable it is joined on.  The permutations
 * therefore hold different quads and different row counts.  A scan is the same binary search over the local shard; the union
 * of the ranks' results is the global scan.  Must be set before the second rfgpu_store_add_index / rfgpu_store_build. */
int32_t rfgpu_store_set_sharded(rfgpu_store* s, int32_t on);
/* Declares how the quads of a sharded store were placed, so that scans can tag their output: under rfgpu_shuffle's
 * destination (id mod n_ranks) the placement column of a permutation holds only ids = rank (mod n_ranks), and a join's
 * direct-address table over such a column needs one slot per n_ranks ids.  Verified on the device by every later
 * rfgpu_store_add_index / rfgpu_store_build (RFGPU_E_INVALID if a quad does not belong to this rank).  Call before adding
 * permutations. */
typedef enum {
  RFGPU_PLACE_NONE = 0,
  RFGPU_PLACE_JOINED_COMPONENT = 1, /* subject for GSPO / GPSO / GSOP, object for GPOS / GOSP / GOPS */
  RFGPU_PLACE_SUBJECT = 2           /* every permutation holds the quads whose subject belongs to the rank */
} rfgpu_placement_mode;
int32_t rfgpu_store_set_placement(rfgpu_store* s, int32_t mode, uint32_t n_ranks, uint32_t rank);
uint64_t rfgpu_store_len(const rfgpu_store* s);
/* rows of permutation i (they differ in a range-sharded store) */
uint64_t rfgpu_store_index_rows(const rfgpu_store* s, int32_t i);
int32_t rfgpu_store_num_indexes(const rfgpu_store* s);
uint32_t rfgpu_store_index_perm(const rfgpu_store* s, int32_t i);

/* ---- scan planning: PlannedPatternScan (scan.rs:414-513) ---- */
/* `gspo` = the four instructions in G,S,P,O order (MemIndexScanInstructions::new_gspo,
 * scan_instructions.rs:19-51: a variable scanned twice becomes Traverse(EqualTo(var))).
 * The permutation is picked by IndexPermutations::choose_index (permutations.rs:81-96). */
int32_t rfgpu_scan_plan_create(rfgpu_ctx* ctx, rfgpu_store* s,
                               const rfgpu_scan_instruction gspo[4], rfgpu_scan_plan** out);
/* Pattern -> plan, the whole of MemQuadStorageSnapshot::plan_pattern_evaluation (snapshot.rs:84-131) after the dictionary
 * look-ups, so that a binder does not re-derive it: level G from the active graph (MemIndexScanInstruction::from_active_graph,
 * scan_instructions.rs:323-355: Default -> In{0}, AllGraphs -> no predicate, Union -> In{ids}, AnyNamed -> Between(1, MAX);
 * Scan(graph_variable) when the graph name is bound to a variable, else Traverse), levels S, P, O from the encoded term
 * patterns (From<EncodedTermPattern>, :358-369).  A constant the dictionary does not know (RFGPU_TERM_UNKNOWN), or a union in
 * which no graph could be encoded (n_graphs = 0), makes the pattern provably empty: *is_empty = 1 — the reference returns
 * PlanPatternScanResult::Empty -> EmptyExec (snapshot.rs:101-110) — and the returned plan yields no rows with the right schema. */
typedef enum { RFGPU_GRAPH_DEFAULT = 0, RFGPU_GRAPH_ALL = 1, RFGPU_GRAPH_UNION = 2, RFGPU_GRAPH_ANY_NAMED = 3 } rfgpu_graph_kind;
typedef struct {
  int32_t kind;            /* rfgpu_graph_kind */
  const uint32_t* graphs;  /* UNION: the encoded graph ids (0 = default graph) */
  uint32_t n_graphs;
} rfgpu_active_graph;
typedef enum { RFGPU_TERM_VARIABLE = 0, RFGPU_TERM_ID = 1, RFGPU_TERM_UNKNOWN = 2 } rfgpu_term_kind;
typedef struct {
  int32_t kind;     /* rfgpu_term_kind */
  const char* var;  /* VARIABLE (blank nodes of a query are variables: BlankNodeMatchingMode::Variable) */
  uint32_t id;      /* ID */
} rfgpu_term_pattern;
int32_t rfgpu_scan_plan_create_pattern(rfgpu_ctx* ctx, rfgpu_store* s, const rfgpu_active_graph* graph,
                                       const char* graph_variable, const rfgpu_term_pattern spo[3], rfgpu_scan_plan** out,
                                       int32_t* is_empty);
void rfgpu_scan_plan_destroy(rfgpu_scan_plan* p);
/* PlannedPatternScan::apply_filter (scan.rs:469-480) for `column <op> value`; re-chooses the
 * index afterwards (try_find_better_index, scan.rs:483-486). */
int32_t rfgpu_scan_plan_apply_filter(rfgpu_ctx* ctx, rfgpu_scan_plan* p, const char* column,
                                     int32_t op /*rfgpu_cmp_op*/, uint32_t value);
/* same for an already-combined Between (dynamic join filters: predicate_pushdown.rs:253-268) */
int32_t rfgpu_scan_plan_apply_between(rfgpu_ctx* ctx, rfgpu_scan_plan* p, const char* column,
                                      uint32_t lo, uint32_t hi);
/* force a permutation (tests / benchmarks); RFGPU_E_INVALID if the store does not hold it */
int32_t rfgpu_scan_plan_set_index(rfgpu_ctx* ctx, rfgpu_scan_plan* p, uint32_t perm);
uint32_t rfgpu_scan_plan_index(const rfgpu_scan_plan* p);
/* "[GPOS] subject=?subject, ..., additional_filters=[object in (2..9)]"-style text of the
 * instructions in GSPO order (pattern_data_source.rs:89-104 fmt_as); returns bytes needed. */
size_t rfgpu_scan_plan_describe(const rfgpu_scan_plan* p, char* buf, size_t cap);

#define RFGPU_SCAN_ALLOW_VIEW 1u /* unfiltered ranges come back as zero-copy views (scan.rs:146-170) */
/* MemQuadIndexScanIterator (scan.rs:101-213) + prune_relevant_row_groups
 * (quad_index_data.rs:155-284) + compute_selection_vector (scan.rs:264-340), whole range in
 * one pass.  Output columns: scanned variables in first-occurrence G,S,P,O order
 * (lib/logical/src/patterns/mod.rs:68-107); graph column nullable, others not. */
int32_t rfgpu_scan_execute(rfgpu_ctx* ctx, const rfgpu_scan_plan* p, uint32_t flags, rfgpu_table** out);
/* Batched form: the range look-ups of all `n` plans run in ONE launch (2 CTAs per plan) and one host hand-off, then every
 * plan's gather / filter is issued back to back.  Same results as n calls of rfgpu_scan_execute; out[i] is plan i's table.
 * For callers that hold several patterns at once (the scans of a BGP: logical_plan_builder_context.rs:233-251). */
int32_t rfgpu_scan_execute_many(rfgpu_ctx* ctx, const rfgpu_scan_plan* const* plans, int32_t n, uint32_t flags,
                                rfgpu_table** out);
/* the contiguous index range pruning narrowed to, and rows kept by the residual predicates */
int32_t rfgpu_scan_last_range(const rfgpu_ctx* ctx, uint64_t* from, uint64_t* to);
/* Interesting orders: among the permutations that tie for the best choose_index score, prefer the one whose scan
 * output comes out sorted by `column` (rows of a scan are in index order, scan.rs:135-205), so that a following
 * join on that variable can run as a streaming merge join instead of a hash join.  *achieved = 1 if such a
 * permutation exists (plan switched to it), 0 otherwise (plan unchanged).  Results are permutation-independent. */
int32_t rfgpu_scan_plan_order_by(rfgpu_ctx* ctx, rfgpu_scan_plan* p, const char* column, int32_t* achieved);

/* ---- tables (solution mappings on the device) ---- */
rfgpu_table* rfgpu_table_retain(rfgpu_table* t);
void rfgpu_table_release(rfgpu_table* t);
/* Row counts are LAZY (RFGPU_OPT_LAZY_ROWS, default on): a join or filter whose output size is only known on the device
 * returns without waiting for it, so that a chain of operators is queued back to back the way the reference streams record
 * batches from operator to operator (lib/storage/src/memory/storage/scan.rs:101-213 — no barrier between operators there).
 * The count is read the first time the host needs it: rfgpu_table_num_rows, any export, or an operator that has to know
 * its input size on the host.  That first read synchronises the ctx stream once (not at all if it was synchronised since)
 * and, if the operator's optimistic output capacity turned out too small, runs it a second time at the exact size — the
 * caller never sees the difference.  rfgpu_table_num_rows has no error channel (0 on failure, message in
 * rfgpu_last_error); rfgpu_table_settle is the same read with a status code.  rfgpu_table_is_pending: 1 while the count
 * has not been read.  rfgpu_ctx_host_syncs / rfgpu_ctx_lazy_redos: how often this ctx waited for its stream / re-ran an
 * operator because its capacity guess was too small. */
uint64_t rfgpu_table_num_rows(const rfgpu_table* t);
int32_t rfgpu_table_settle(rfgpu_ctx* ctx, const rfgpu_table* t);
int32_t rfgpu_table_is_pending(const rfgpu_table* t);
uint64_t rfgpu_ctx_host_syncs(const rfgpu_ctx* ctx);
/* device allocations of this ctx since the last reset: out = {calls, host microseconds inside cudaMallocAsync, the longest call,
 * calls that took more than 20 us (the stream-ordered pool could not reuse a block and went to the driver)} */
int32_t rfgpu_ctx_alloc_stats(rfgpu_ctx* ctx, double out[4], int32_t reset);
uint64_t rfgpu_ctx_lazy_redos(const rfgpu_ctx* ctx);
int32_t rfgpu_table_num_cols(const rfgpu_table* t);
const char* rfgpu_table_col_name(const rfgpu_table* t, int32_t col);
int32_t rfgpu_table_col_nullable(const rfgpu_table* t, int32_t col);
/* device pointer of a column (uint32_t[num_rows], 0 = NULL), valid while the table lives */
const uint32_t* rfgpu_table_col_device_ptr(const rfgpu_table* t, int32_t col);
/* Import host columns (an Arrow RecordBatch's UInt32 value buffers + optional validity
 * bitmaps, LSB-first as in Arrow).  `pinned` != 0 promises page-locked buffers. */
int32_t rfgpu_table_import(rfgpu_ctx* ctx, int32_t n_cols, const char* const* names,
                           const uint8_t* nullable, const uint32_t* const* values,
                           const uint8_t* const* validity, uint64_t n_rows, rfgpu_table** out);
/* Export rows [row_off, row_off+n_rows) into caller buffers; validity[i] may be NULL (then
 * nothing is written for it), else receives an Arrow bitmap (bit = value != 0).  Returns the
 * null count per column in null_counts when non-NULL. */
int32_t rfgpu_table_export(rfgpu_ctx* ctx, const rfgpu_table* t, uint64_t row_off, uint64_t n_rows,
                           uint32_t* const* values, uint8_t* const* validity, uint64_t* null_counts);
/* Asynchronous export of the value buffers (no validity: NULL = 0 in the values): the device→host copies run on the ctx's
 * copy stream once the work queued so far has produced `t`, and overlap later kernels and host→device imports.  The caller's
 * buffers (page-locked for a real overlap) are valid after the next rfgpu_ctx_synchronize; `t` is retained until then. */
int32_t rfgpu_table_export_async(rfgpu_ctx* ctx, rfgpu_table* t, uint64_t row_off, uint64_t n_rows,
                                 uint32_t* const* values);
/* Bounds of the non-NULL values of a column: what HashJoinExec publishes as its dynamic filter (build-side min/max →
 * Between on the probe-side scan: pattern_data_source.rs:107-151, predicate_pushdown.rs:253-268).  The host feeds the result
 * to rfgpu_scan_plan_apply_between before the probe-side scan runs.  No non-NULL value: min = UINT32_MAX, max = 0. */
int32_t rfgpu_table_col_min_max(rfgpu_ctx* ctx, const rfgpu_table* t, int32_t col, uint32_t* min_out, uint32_t* max_out,
                                uint64_t* non_null);
/* zero-copy projection / rename (DataFusion ProjectionExec over plain columns) */
int32_t rfgpu_table_project(rfgpu_ctx* ctx, const rfgpu_table* t, int32_t n_cols,
                            const int32_t* cols, const char* const* new_names, rfgpu_table** out);
/* override schema nullability (a field's is_nullable is a schema property: join/rewrite.rs:131-139) */
int32_t rfgpu_table_set_nullable(rfgpu_table* t, int32_t col, int32_t nullable);
/* column the rows are known to be sorted by (ascending, NULL first), or -1.  Set by scans, kept by filters and joins. */
int32_t rfgpu_table_sorted_col(const rfgpu_table* t);
/* Declare an imported table sorted by `col` (the child plan's output_ordering); verified on the device
 * (RFGPU_E_INVALID if the column is not ascending).  col = -1 clears the property. */
int32_t rfgpu_table_set_sorted_col(rfgpu_ctx* ctx, rfgpu_table* t, int32_t col);

/* ---- Arrow C Data Interface (what arrow-rs `arrow::ffi` speaks) ---- */
struct ArrowSchema;
struct ArrowArray;
/* struct-typed array (a RecordBatch) of UInt32 children → table; does not take ownership
 * of the input (caller still calls release). */
int32_t rfgpu_table_import_arrow(rfgpu_ctx* ctx, const struct ArrowArray* array,
                                 const struct ArrowSchema* schema, rfgpu_table** out);
/* rows [row_off,row_off+n_rows) → newly allocated host buffers owned by the returned
 * ArrowArray (release callback frees them); one struct array with one UInt32 child per column. */
int32_t rfgpu_table_export_arrow(rfgpu_ctx* ctx, const rfgpu_table* t, uint64_t row_off,
                                 uint64_t n_rows, struct ArrowArray* out_array,
                                 struct ArrowSchema* out_schema);

/* ---- ID-level filters: DataFusion FilterExec over object-id columns ---- */
typedef enum {
  RFGPU_F_EQ_CONST = 0,   /* col = c      (lib/logical/src/patterns/rewrite.rs:68-110)   */
  RFGPU_F_NE_CONST = 1,   /* col != c     (BSBM explore Q5 `product@0 != <oid>`)          */
  RFGPU_F_EQ_COL = 2,     /* col = col2   (sameTerm / filter_same_variable :112-148)      */
  RFGPU_F_NE_COL = 3,     /* col != col2                                                  */
  RFGPU_F_BETWEEN = 4,    /* lo <= col <= hi                                              */
  RFGPU_F_IN = 5,         /* col IN ids                                                   */
  RFGPU_F_IS_NULL = 6,    /* NOT BOUND(col)  (lib/logical/src/expr_builder.rs:58-60)      */
  RFGPU_F_IS_NOT_NULL = 7 /* BOUND(col)                                                   */
} rfgpu_filter_op;

typedef struct {
  int32_t op;           /* rfgpu_filter_op */
  int32_t col, col2;    /* column indexes in the input table */
  uint32_t value, hi;   /* EQ/NE const: value; BETWEEN: [value, hi] */
  const uint32_t* ids;  /* IN */
  uint32_t n_ids;
} rfgpu_filter_pred;

/* Conjunction of `n` predicates under SQL three-valued logic (a NULL comparison drops the
 * row), order-preserving compaction of all columns. */
int32_t rfgpu_filter(rfgpu_ctx* ctx, const rfgpu_table* in, const rfgpu_filter_pred* preds,
                     int32_t n, rfgpu_table** out);

/* The same filter for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  It calls no cudaMalloc /
 * cudaFree / allocating routine: the caller passes device pointers for the n_cols input columns (uint32_t[n_rows]), for the
 * n_cols output columns (capacity n_rows each; rows kept, in input order) and for the row count (one uint64_t in device
 * memory, written by the kernel; read it after synchronising the ctx stream).  preds[i].col / col2 index `cols`.  The
 * kernel's look-back scratch is a static device array of 128 KB that every GPU loading the library holds until exit:
 * n_rows > RFGPU_FILTER_NOALLOC_MAX_ROWS -> RFGPU_E_UNSUPPORTED, as are IN lists of more than 8 ids (they would need a
 * device copy).  All calls on one device share that array: calls that overlap on the GPU (two ctxs / streams of one device)
 * race; calls on one ctx are stream-ordered and safe back to back.  Asynchronous: the work is queued on the ctx stream. */
#define RFGPU_FILTER_NOALLOC_MAX_ROWS (1ull << 26)
int32_t rfgpu_filter_noalloc(rfgpu_ctx* ctx, int32_t n_cols, const uint32_t* const* cols, uint64_t n_rows,
                             const rfgpu_filter_pred* preds, int32_t n_preds, uint32_t* const* out_cols,
                             uint64_t* out_rows_device);

/* ---- SPARQL join: SparqlJoinNode + SparqlJoinLoweringRule ---- */
typedef enum { RFGPU_JOIN_INNER = 0, RFGPU_JOIN_LEFT = 1 } rfgpu_join_type;
typedef enum {
  RFGPU_JOIN_ALGO_CROSS = 0,     /* disjoint schemas: join/rewrite.rs:76-96                 */
  RFGPU_JOIN_ALGO_EQUI = 1,      /* try_build_regular_join :126-168 (NullEqualsNothing)     */
  RFGPU_JOIN_ALGO_COMPATIBLE = 2 /* build_join_with_is_compatible :183-221 (+COALESCE)      */
} rfgpu_join_algo;

/* Joins on the variables both schemas share (get_join_keys, join/rewrite.rs:363-379).
 * Lowering follows rewrite_sparql_join (:71-111): no shared variable → cross / keyless left;
 * no shared column nullable by schema → equi hash join; otherwise compatible-mapping join
 * with COALESCE(lhs.k, rhs.k).  Output schema: all lhs columns, then rhs-only columns
 * (join/logical.rs:251-279; rhs-only columns nullable for Left).  The multiset of rows is
 * what DataFusion's HashJoinExec / NestedLoopJoinExec / CrossJoinExec would produce. */
int32_t rfgpu_sparql_join(rfgpu_ctx* ctx, const rfgpu_table* lhs, const rfgpu_table* rhs,
                          int32_t join_type, rfgpu_table** out);
/* The same join with SparqlJoinNode's optional filter (lib/logical/src/join/logical.rs:50; join/rewrite.rs:149-161 hands it to
 * DataFusion as the join filter, :206-218 ANDs it to the IS_COMPATIBLE conjunction): a conjunction of ID-level predicates
 * over the RESULT row — preds[i].col / col2 index the output schema (lhs columns, then rhs-only columns), a shared variable of
 * a compatible join is its COALESCE (value_from_joined, :321-352).  Inner: the join followed by the filter.  Left
 * (OPTIONAL { P FILTER(e) }): an lhs row is padded with NULLs iff no compatible rhs row passes the filter (SPARQL 1.1 §18.5
 * LeftJoin).  Disjoint schemas with a filter are a filtered product (rewrite.rs:76 takes the cross-join shortcut only
 * without a filter).  n_preds = 0 is rfgpu_sparql_join.  Row order unspecified. */
int32_t rfgpu_sparql_join_filter(rfgpu_ctx* ctx, const rfgpu_table* lhs, const rfgpu_table* rhs, int32_t join_type,
                                 const rfgpu_filter_pred* preds, int32_t n_preds, rfgpu_table** out);
/* SPARQL MINUS: MinusNode + MinusLoweringRule (lib/logical/src/minus/rewrite.rs:58-141) — a LeftAnti
 * join that removes an lhs row iff some rhs row is compatible with it on every shared variable (equal, or NULL on either
 * side) AND at least one shared variable is bound on both sides.  No shared variable → the lhs unchanged (:61-64).
 * Output: the lhs schema, surviving rows in lhs order.  rfgpu_last_join_algo reports EQUI (no shared column nullable),
 * COMPATIBLE, or CROSS (no shared variable).  Up to 4 shared variables, at most 2 bound on both sides of a class pair. */
int32_t rfgpu_sparql_minus(rfgpu_ctx* ctx, const rfgpu_table* lhs, const rfgpu_table* rhs, rfgpu_table** out);
/* FILTER EXISTS { rhs } (negated = 0) / FILTER NOT EXISTS { rhs } (negated = 1) over the solutions `lhs`: the correlated
 * sub-query of RdfFusionExprBuilderContext::exists / not_exists (lib/logical/src/expr_builder_context.rs:203-338), i.e. the
 * LeftSemi / LeftAnti join DataFusion decorrelates it into: an lhs row has a partner iff some rhs row is compatible with it
 * on every shared variable (equal, or NULL on either side) — no "bound on both sides" requirement, unlike MINUS.
 * Disjoint schemas: every lhs row has a partner iff rhs has a row (:246-262).  Output: lhs schema, lhs order. */
int32_t rfgpu_sparql_exists(rfgpu_ctx* ctx, const rfgpu_table* lhs, const rfgpu_table* rhs, int32_t negated,
                            rfgpu_table** out);
/* which lowering the last rfgpu_sparql_join / rfgpu_sparql_minus on this ctx took (rfgpu_join_algo) */
int32_t rfgpu_last_join_algo(const rfgpu_ctx* ctx);
/* physical operator the last equi (sub-)join used */
typedef enum {
  RFGPU_JOIN_PHYS_NONE = 0,
  RFGPU_JOIN_PHYS_HASH = 1,  /* one open-addressing table in HBM */
  RFGPU_JOIN_PHYS_MERGE = 2, /* both inputs sorted on the key */
  RFGPU_JOIN_PHYS_RADIX = 3, /* radix-partitioned, per-partition tables in shared memory */
  RFGPU_JOIN_PHYS_SORT_MERGE = 4, /* unsorted side(s) radix-sorted on the key, then the merge join (RFGPU_OPT_SORT_MERGE_JOIN);
                                    for MINUS / EXISTS: sorted rhs key list + binary search (RFGPU_OPT_SORT_SEMI) */
  RFGPU_JOIN_PHYS_LOOKUP = 5      /* dense build keys: direct-address (identity hash) table, one u32 per possible key */
} rfgpu_join_phys;
int32_t rfgpu_last_join_phys(const rfgpu_ctx* ctx);

/* ---- aggregation on id columns: DataFusion AggregateExec(mode=Single, gby=[..], aggr=[COUNT]) ---- */
/* GROUP BY key_cols with one COUNT: counted_col = -1 → COUNT(*), else COUNT(col) (non-NULL values).  NULL keys form a
 * group of their own (SQL GROUP BY).  Output: the key columns, then a non-nullable column `count_name` holding the count
 * as a uint32 NUMBER (not an object id; INT64_AS_TERM turns it into a term on the host:
 * lib/logical/src/logical_plan_builder.rs:235-264).  Row order unspecified. */
int32_t rfgpu_group_count(rfgpu_ctx* ctx, const rfgpu_table* in, const int32_t* key_cols, int32_t n_keys,
                          int32_t counted_col, const char* count_name, rfgpu_table** out);
/* SELECT DISTINCT over all columns (ReplaceDistinctWithAggregate: a GROUP BY on every column). */
int32_t rfgpu_distinct(rfgpu_ctx* ctx, const rfgpu_table* in, rfgpu_table** out);

/* ---- object id -> plain term decode: a device dictionary gather ---- */
/* Replaces MemObjectIdMapping::decode_array (lib/storage/src/memory/object_id_mapping.rs:331-378) behind ENC_PT
 * (lib/functions/src/builtin/encoding/with_plain_term_encoding.rs:60-100).  The dictionary holds, for the ids 1..n_ids
 * (arrays are indexed by id, entry 0 is unused; id 0 = NULL decodes to a null struct):
 *   term_type[id]   PlainTermType: 0 named node, 1 blank node, 2 literal (lib/encoding/src/plain_term/encoding.rs:108-116)
 *   value_off       n_ids + 2 ascending byte offsets: the lexical value of id is value_bytes[value_off[id], value_off[id+1])
 *   datatype[id]    literals: index (1..n_aux) of the data type IRI in the string table (rdf:langString for language-tagged
 *                   literals, as LiteralRef::datatype() reports); 0 for named / blank nodes
 *   language[id]    language-tagged literals: index of the tag in the string table; 0 otherwise
 *   aux_off         n_aux + 2 ascending offsets into aux_bytes (entry 0 unused)
 * All pointers are host memory; the dictionary is copied to the device once.  The encode direction stays on the host. */
typedef struct rfgpu_dict rfgpu_dict;
typedef struct rfgpu_plain_terms rfgpu_plain_terms;
int32_t rfgpu_dict_create(rfgpu_ctx* ctx, uint32_t n_ids, const uint8_t* term_type, const uint64_t* value_off,
                          const uint8_t* value_bytes, const uint32_t* datatype, const uint32_t* language, uint32_t n_aux,
                          const uint32_t* aux_off, const uint8_t* aux_bytes, rfgpu_dict** out);
rfgpu_dict* rfgpu_dict_retain(rfgpu_dict* d);
void rfgpu_dict_release(rfgpu_dict* d);
uint32_t rfgpu_dict_len(const rfgpu_dict* d);
/* Decodes rows [row_off, row_off + n_rows) of column `col` into the four children of the PlainTerm struct array
 * (lib/encoding/src/plain_term/element_builder.rs:29-134), device-resident.  An id that is neither 0 nor in 1..n_ids is
 * RFGPU_E_INVALID (the reference panics: "Missing EncodedObjectId").  More than 2^31-2 bytes in one child: RFGPU_E_UNSUPPORTED
 * (Utf8 offsets are i32) — decode a smaller row range. */
int32_t rfgpu_dict_decode(rfgpu_ctx* ctx, const rfgpu_dict* d, const rfgpu_table* t, int32_t col, uint64_t row_off,
                          uint64_t n_rows, rfgpu_plain_terms** out);
void rfgpu_plain_terms_release(rfgpu_plain_terms* p);
/* rows; byte sizes of the value / data_type / language_tag children; null counts of the struct (= term_type = value
 * children), the data_type and the language_tag children */
int32_t rfgpu_plain_terms_sizes(const rfgpu_plain_terms* p, uint64_t* rows, uint64_t bytes[3], uint64_t nulls[3]);
/* Copies the Arrow buffers to host memory: term_type[rows]; per child k (value, data_type, language_tag) the validity
 * bitmap ((rows + 7) / 8 bytes; validity[0] is also the struct's and term_type's), rows + 1 i32 offsets and the bytes. */
int32_t rfgpu_plain_terms_export(rfgpu_ctx* ctx, const rfgpu_plain_terms* p, uint8_t* term_type, uint8_t* const validity[3],
                                 int32_t* const offsets[3], uint8_t* const data[3]);

/* Value-level filters without decoding (the other half of §8f-2: every value FILTER of the reference starts with ENC_TV(col),
 * lib/functions/src/builtin/encoding/with_typed_value_encoding.rs): appends the column `name` = table[t.col[i]] to `t` (ids
 * outside [0, n_table), NULL ids and zero entries read as NULL).  With a table that maps an object id to the 1-based RANK of
 * its numeric value among the dictionary's numeric literals (0 for every other term; built by the host next to the
 * dictionary), `?x > c` is RFGPU_F_BETWEEN(rank(c) + 1, MAX) on the new column — an ID-level predicate for rfgpu_filter or
 * rfgpu_sparql_join_filter — and a non-numeric ?x fails the comparison like SPARQL's type error does.  `table` is host
 * memory, or device memory when on_device != 0. */
int32_t rfgpu_table_lookup(rfgpu_ctx* ctx, const rfgpu_table* t, int32_t col, const uint32_t* table, uint32_t n_table,
                           int32_t on_device, const char* name, rfgpu_table** out);

/* ---- object id -> typed value: ENC_TV (the rest of SURVEY §8f-2) ---- */
/* Replaces MemObjectIdMapping::decode_array_to_typed_value (lib/storage/src/memory/object_id_mapping.rs:376-402) behind
 * ENC_TV (lib/functions/src/builtin/encoding/with_typed_value_encoding.rs:63-84), and the comparison operators evaluated
 * on its result (lib/functions/src/scalar/comparison/{equal,less_than,less_or_equal,greater_than,greater_or_equal}.rs over
 * TypedValueRef::partial_cmp, lib/model/src/typed_value.rs:162-262).
 *
 * The reference keeps an EncodedTypedValue per object id (lib/storage/src/memory/encoding/typed_value.rs:9-25), parsed once
 * from the term's lexical form when the term is interned (TypedValueRef::try_from, lib/model/src/typed_value.rs:337-410).
 * The device twin is a table of one 32-byte record per id — one sector per gathered id — holding the fixed-width part of
 * that value; the strings of the string-valued variants stay in the rfgpu_dict the table is created over.  Parsing stays
 * with the reference's own parsers on the host: the Rust binding fills rfgpu_tv_value from EncodedTypedValue. */
typedef struct rfgpu_tv_table rfgpu_tv_table;
typedef struct rfgpu_typed_values rfgpu_typed_values;

/* TypedValueEncodingField = the type ids of the dense union (lib/encoding/src/typed_value/encoding.rs:241-262, :343-362) */
typedef enum {
  RFGPU_TV_NULL = 0,        /* unbound, or a literal whose lexical form is not in its datatype's lexical space (Invalid) */
  RFGPU_TV_NAMED_NODE = 1,  /* Utf8: the dictionary's value */
  RFGPU_TV_BLANK_NODE = 2,  /* Utf8 */
  RFGPU_TV_STRING = 3,      /* Struct{value: Utf8, language: Utf8 nullable}: simple literals / xsd:string and language-tagged strings */
  RFGPU_TV_BOOLEAN = 4,     /* Boolean: lo = 0 / 1 */
  RFGPU_TV_FLOAT = 5,       /* Float32: lo = the f32 bits */
  RFGPU_TV_DOUBLE = 6,      /* Float64: lo = the f64 bits */
  RFGPU_TV_DECIMAL = 7,     /* Decimal128: (hi, lo) = the i128 of lib/model/src/xsd/decimal.rs (value * 10^18);
                               aux = the f64 bits of Double::from(decimal) (decimal.rs:445-462), used when it meets a float / double */
  RFGPU_TV_INT = 8,         /* Int32: lo = the value, sign-extended to 64 bits */
  RFGPU_TV_INTEGER = 9,     /* Int64: lo */
  RFGPU_TV_DATE_TIME = 10,  /* Struct{value: Decimal128, offset: Int16 nullable}: (hi, lo) = Timestamp::value(), */
  RFGPU_TV_TIME = 11,       /*   aux = the timezone offset in minutes when RFGPU_TV_HAS_AUX is set */
  RFGPU_TV_DATE = 12,       /*   (lib/encoding/src/typed_value/element_builder.rs:385-398) */
  RFGPU_TV_DURATION = 13,   /* Struct{months: Int64 nullable, seconds: Decimal128 nullable}: aux = months when RFGPU_TV_HAS_AUX,
                               (hi, lo) = seconds when RFGPU_TV_HAS_VALUE (xsd:duration both, yearMonthDuration / dayTimeDuration one) */
  RFGPU_TV_OTHER_LITERAL = 14 /* Struct{value: Utf8, datatype: Utf8} */
} rfgpu_tv_field;
#define RFGPU_TV_N_FIELDS 15
#define RFGPU_TV_HAS_AUX 1
#define RFGPU_TV_HAS_VALUE 2

/* One typed value in its fixed-width form (32 bytes, no implicit padding). */
typedef struct {
  uint64_t lo, hi;
  int64_t aux;
  uint8_t field;  /* rfgpu_tv_field */
  uint8_t flags;  /* RFGPU_TV_HAS_AUX | RFGPU_TV_HAS_VALUE */
  uint8_t reserved[6];
} rfgpu_tv_value;

/* values[0 .. n_ids]: indexed by object id, entry 0 unused; n_ids must equal rfgpu_dict_len(d).  `d` is retained.  The field
 * of a string-valued entry must agree with the dictionary's term type (named node / blank node / literal): RFGPU_E_INVALID. */
int32_t rfgpu_tv_table_create(rfgpu_ctx* ctx, rfgpu_dict* d, const rfgpu_tv_value* values, uint32_t n_ids,
                              rfgpu_tv_table** out);
rfgpu_tv_table* rfgpu_tv_table_retain(rfgpu_tv_table* tv);
void rfgpu_tv_table_release(rfgpu_tv_table* tv);

/* ENC_TV(col): rows [row_off, row_off + n_rows) of column `col` as the dense union array of the TypedValue encoding, device
 * resident: type_ids, offsets, and every child compacted in row order (TypedValueArrayElementBuilder,
 * lib/encoding/src/typed_value/element_builder.rs:83-398).  NULL ids and Invalid values land in the null child.  An id beyond
 * the dictionary is RFGPU_E_INVALID (the reference panics: "Missing EncodedObjectId"). */
int32_t rfgpu_tv_decode(rfgpu_ctx* ctx, const rfgpu_tv_table* tv, const rfgpu_table* t, int32_t col, uint64_t row_off,
                        uint64_t n_rows, rfgpu_typed_values** out);
void rfgpu_typed_values_release(rfgpu_typed_values* v);
/* rows and the length of each of the 15 children */
int32_t rfgpu_typed_values_sizes(const rfgpu_typed_values* v, uint64_t* rows, uint64_t counts[RFGPU_TV_N_FIELDS]);
/* The strings of a string-valued child (field 1, 2, 3 or 14) in PlainTerm form, one row per child element (borrowed: valid
 * until `v` is released; NULL for an empty child or another field): named / blank node = its `value` child; string = value +
 * language_tag; other literal = value + data_type.  Read them with rfgpu_plain_terms_sizes / _export. */
const rfgpu_plain_terms* rfgpu_typed_values_strings(const rfgpu_typed_values* v, int32_t field);
/* Host destinations of the fixed-width buffers; a NULL member is skipped.  Decimal128 values are 16 bytes each, little endian
 * (Arrow's layout); validity bitmaps are LSB-first, (count + 7) / 8 bytes. */
typedef struct {
  int8_t* type_ids;            /* [rows] */
  int32_t* offsets;            /* [rows] */
  uint8_t* boolean_bits;       /* (counts[4] + 7) / 8 bytes */
  float* floats;               /* [counts[5]] */
  double* doubles;             /* [counts[6]] */
  uint8_t* decimals;           /* [counts[7]] x 16 B */
  int32_t* ints;               /* [counts[8]] */
  int64_t* integers;           /* [counts[9]] */
  uint8_t* ts_value[3];        /* date_time, time, date: [counts[10 + k]] x 16 B */
  int16_t* ts_offset[3];       /* [counts[10 + k]] */
  uint8_t* ts_offset_valid[3];
  int64_t* dur_months;         /* [counts[13]] */
  uint8_t* dur_months_valid;
  uint8_t* dur_seconds;        /* [counts[13]] x 16 B */
  uint8_t* dur_seconds_valid;
} rfgpu_tv_export;
int32_t rfgpu_typed_values_export(rfgpu_ctx* ctx, const rfgpu_typed_values* v, const rfgpu_tv_export* dst);

/* A value FILTER evaluated where the ids are: a postfix program over typed values whose result is the effective boolean
 * value of the expression per row.  Comparisons push one truth value (SPARQL's three-valued logic: true / false / error),
 * AND / OR / NOT combine the top of the stack as SPARQL 1.1 §17.2 defines (error || true = true, error && false = false).
 * The output is the input table plus the nullable column `name`: 1 = true, 2 = false, NULL = error — so FILTER(e) is
 * RFGPU_F_EQ_CONST(name, 1) for rfgpu_filter / rfgpu_sparql_join_filter and FILTER(!e) is RFGPU_F_EQ_CONST(name, 2).
 * A comparison whose operand is NULL or Invalid is an error (dispatch/binary.rs: error_op), as is every pair partial_cmp
 * leaves unordered.  Duration-against-duration comparisons (calendar arithmetic on four reference dates,
 * lib/model/src/xsd/duration.rs:271-308) are not evaluated on the device: RFGPU_E_UNSUPPORTED if a row needs one. */
typedef enum {
  RFGPU_TV_EQ = 0,  /* partial_cmp == Equal   (equal.rs:57) */
  RFGPU_TV_LT = 1,  /* == Less                (less_than.rs:55) */
  RFGPU_TV_LE = 2,  /* Equal or Less          (less_or_equal.rs:55) */
  RFGPU_TV_GT = 3,  /* == Greater             (greater_than.rs:55) */
  RFGPU_TV_GE = 4,  /* Equal or Greater       (greater_or_equal.rs:55) */
  RFGPU_TV_AND = 5, /* pops two, pushes one */
  RFGPU_TV_OR = 6,
  RFGPU_TV_NOT = 7, /* pops one, pushes one */
  RFGPU_TV_BOUND = 8 /* BOUND(lhs column): true / false, never an error */
} rfgpu_tv_op;
typedef struct {
  int32_t op;        /* rfgpu_tv_op */
  int32_t lhs, rhs;  /* comparison operands: >= 0 a column of the table, < 0 the constant -(k + 1) */
} rfgpu_tv_instr;
typedef struct {
  rfgpu_tv_value v;
  const uint8_t* str;   /* string-valued fields: the value bytes (host memory) */
  uint32_t str_len;
  const uint8_t* str2;  /* STRING with RFGPU_TV_HAS_AUX: the language tag; OTHER_LITERAL: the datatype IRI */
  uint32_t str2_len;
} rfgpu_tv_const;
#define RFGPU_TV_MAX_INSTRS 16
#define RFGPU_TV_MAX_CONSTS 8
#define RFGPU_TV_MAX_CONST_BYTES 1024 /* all constant strings of one program together */
int32_t rfgpu_tv_eval(rfgpu_ctx* ctx, const rfgpu_tv_table* tv, const rfgpu_table* t, const rfgpu_tv_instr* prog,
                      int32_t n_instr, const rfgpu_tv_const* consts, int32_t n_consts, const char* name, rfgpu_table** out);

/* physical operator of the last rfgpu_group_count / rfgpu_distinct */
typedef enum {
  RFGPU_AGG_PHYS_NONE = 0,
  RFGPU_AGG_PHYS_HASH = 1, /* representative-row table in HBM */
  RFGPU_AGG_PHYS_SORT = 2  /* radix sort by the keys + run-length encoding (RFGPU_OPT_SORT_AGG) */
} rfgpu_agg_phys;
int32_t rfgpu_last_agg_phys(const rfgpu_ctx* ctx);

/* ---- sorting on ids (sort.cu): the primitive under the index build and the sort-based operators ---- */
/* Rows of `t` ordered ascending by key_cols (major first; NULL = 0 first), ties in input order (stable LSD radix sort).
 * The result has sorted_col = key_cols[0], so a following single-key rfgpu_sparql_join on that column streams through
 * the merge join.  Ordering is by object id, not by term value: this is an execution primitive, not SPARQL ORDER BY. */
int32_t rfgpu_table_sort(rfgpu_ctx* ctx, const rfgpu_table* t, const int32_t* key_cols, int32_t n_keys, rfgpu_table** out);
/* Bulk load: builds the permutations `perms` of the quads (quads[0..3] = graph, subject, predicate, object
 * id columns, any order, duplicates allowed; host pointers, or device pointers when on_device != 0) on the device: per
 * permutation a stable radix sort from the least significant level up, then de-duplication (the store is a set).
 * Replaces MemQuadIndex::insert / MemIndexData::insert for every permutation (lib/storage/src/memory/storage/
 * quad_index.rs:132-156, quad_index_data.rs:287-376, lib/storage/src/index/permutations.rs:102-128). */
int32_t rfgpu_store_build(rfgpu_ctx* ctx, rfgpu_store* s, const uint32_t* const quads[4], uint64_t n_rows,
                          int32_t on_device, const uint32_t* perms, int32_t n_perms);

/* ---- synthetic graphs in id space (benchmark / test set-up) ---- */
/* G(N, P, s, seed): the deterministic graph every benchmark and large parity test runs on, generated on the device so that
 * each rank of a multi-GPU run produces exactly the rows it owns without communication.  The reference generates its own
 * benchmark data on the host (bench/src/benchmarks/windfarm/generate/generate_time_series.rs:23; BSBM through the Java
 * bsbm-tools); this is the counterpart for the synthetic configs of BASELINE.json.  Bit-identical with the numpy generator
 * rdf_fusion_b200/synth.py::generate_graph (the specification, and what the oracle consumes): row i draws
 * splitmix64(i ^ splitmix64(seed * 256 + stream)); subjects uniform over S = N / 10, predicates Zipf by rank through
 * `pred_cdf`, objects half entities (uniform, or Zipf through `obj_cdf`) half literals (uniform over L = N / 4).
 * Duplicates are NOT removed (rfgpu_store_build does that). */
typedef enum {
  RFGPU_SYNTH_KEEP_ALL = 0,
  RFGPU_SYNTH_KEEP_HASH_SUBJECT = 1, /* rows with rfgpu_shuffle_dest(subject, n_ranks) == rank */
  RFGPU_SYNTH_KEEP_RANGE = 2,        /* rows whose quad, in the level order of `perm`, is in [lo, hi): one rank's contiguous
                                        slice of the globally sorted permutation (range partitioning by leading key) */
  RFGPU_SYNTH_KEEP_HASH_OBJECT = 3   /* rows with rfgpu_shuffle_dest(object, n_ranks) == rank */
} rfgpu_synth_keep_mode;
typedef struct {
  uint64_t n_rows;          /* N: rows of the GLOBAL graph */
  uint32_t n_predicates;    /* P */
  uint64_t seed;
  const double* pred_cdf;   /* host, [P]: cumulative Zipf weights, last = 1 */
  const double* obj_cdf;    /* host, [max(N / 10, 1)], or NULL for uniform entity objects */
  double named_graph_frac;  /* > 0: that fraction of the quads is placed in the named graphs 1..16 */
  int32_t keep;             /* rfgpu_synth_keep_mode */
  uint32_t rank, n_ranks;   /* HASH_SUBJECT / HASH_OBJECT */
  uint32_t perm;            /* RANGE: RFGPU_PERM(..) */
  uint32_t lo[4], hi[4];    /* RANGE: bounds in level order */
  int32_t has_lo, has_hi;   /* RANGE: 0 = unbounded on that side (first / last rank) */
} rfgpu_synth_desc;
/* -> a table (graph, subject, predicate, object) of the kept rows, device-resident, in no particular order */
int32_t rfgpu_synth_quads(rfgpu_ctx* ctx, const rfgpu_synth_desc* desc, rfgpu_table** out);

/* ---- options ---- */
typedef enum {
  RFGPU_OPT_MERGE_JOIN = 1, /* 1 (default): one-key equi joins whose inputs are both sorted on the key run as a
                              streaming merge join; 0: always build/probe the hash table */
  RFGPU_OPT_RADIX_JOIN = 3,     /* 1: equi joins with at least RADIX_MIN_ROWS build rows are radix-partitioned (default 0:
                                   experimental) */
  RFGPU_OPT_RADIX_MIN_ROWS = 4, /* default 65536 */
  RFGPU_OPT_HASH_L2_BYTES = 5,  /* hash join: table bytes per pass (default: unlimited = one pass); larger build sides are
                                   joined in up to 8 passes over disjoint hash ranges so each pass's table stays in L2 */
  RFGPU_OPT_MERGE_THREADS = 2, /* CTA size of the merge-join kernel: 64, 128 (default) or 256 threads, 8 probe rows each */
  RFGPU_OPT_SHUFFLE_P2P = 6,   /* 1 (default): rfgpu_shuffle partitions straight into the destination ranks' receive arenas
                                  over NVLink peer memory (one fused kernel); 0: partition locally, then ncclSend/ncclRecv */
  RFGPU_OPT_SORT_MERGE_JOIN = 8, /* 1: a single-key equi join whose inputs are not both sorted on the key sorts the
                                    unsorted side(s) and runs the merge join instead of the hash join (default 0) */
  RFGPU_OPT_SORT_SEMI = 9,       /* 1: single-key MINUS / [NOT] EXISTS without nullable shared columns searches a sorted
                                    list of the rhs keys instead of probing a hash table (default 0) */
  RFGPU_OPT_SORT_AGG = 10,       /* 1 (default): GROUP BY / COUNT / DISTINCT over at most 4 columns by sort + run-length
                                    encoding; 0: representative-row hash table */
  RFGPU_OPT_LOOKUP_JOIN = 11,    /* 1 (default): single-key equi joins, MINUS / EXISTS and GROUP BY whose (build) keys span a
                                    range of at most 8 slots per input row use a direct-address table (slot = id - min id)
                                    instead of the open-addressing hash table; 0: always hash */
  RFGPU_OPT_SHUFFLE_BULK = 13,   /* 1 (default): the fused shuffle kernel writes each destination's run with bulk copies (copy engine,
                                    shared -> peer memory); 0: one 4-byte store per row and thread */
  RFGPU_OPT_LAZY_ROWS = 12,      /* 1 (default): joins and filters return before their output row count has reached the host
                                    (see rfgpu_table_num_rows); 0: every operator waits for its count */
  RFGPU_OPT_SHUFFLE_ARENA_BYTES = 7 /* size of this rank's receive arena (two halves, used alternately); read by
                                       rfgpu_comm_init, default 1 GiB.  A shuffle whose rows do not fit one half on some
                                       rank falls back to the NCCL exchange on every rank */
} rfgpu_option;
int32_t rfgpu_ctx_set_option(rfgpu_ctx* ctx, int32_t option, int64_t value);

/* ---- multi-GPU: one process (ctx) per GPU, NCCL over NVLink ---- */
#define RFGPU_UNIQUE_ID_BYTES 128
int32_t rfgpu_comm_unique_id(uint8_t id[RFGPU_UNIQUE_ID_BYTES]);
int32_t rfgpu_comm_init(rfgpu_ctx* ctx, int32_t rank, int32_t n_ranks,
                        const uint8_t id[RFGPU_UNIQUE_ID_BYTES]);
int32_t rfgpu_comm_rank(const rfgpu_ctx* ctx);
int32_t rfgpu_comm_size(const rfgpu_ctx* ctx);
/* Hash-partition `in` on column `key_col` (dest = id mod n_ranks — ids are dense dictionary ids; NULL is an ordinary key value and goes to rank 0) and
 * exchange the rows between the ranks: the exchange DataFusion's RepartitionExec(Hash) does between tokio tasks.
 * Collective: every rank of the communicator calls it, in the same order.  Default path: ONE kernel that computes the
 * destinations, reserves space in each destination's receive arena with an atomic over NVLink and writes the rows there
 * through peer memory (cudaIpc mappings set up by rfgpu_comm_init), followed by one 8-byte-per-rank all-gather that is
 * both the completion barrier and the row-count exchange.  Fallback (option off, IPC unavailable, arena too small):
 * partition kernels + count-matrix all-gather + grouped ncclSend/ncclRecv.  Row order of the result is unspecified. */
int32_t rfgpu_shuffle(rfgpu_ctx* ctx, const rfgpu_table* in, int32_t key_col, rfgpu_table** out);
/* which exchange the last rfgpu_shuffle of this ctx used */
#define RFGPU_SHUFFLE_LOCAL 0 /* one rank: partition only */
#define RFGPU_SHUFFLE_NCCL 1  /* partition kernels + grouped ncclSend/ncclRecv */
#define RFGPU_SHUFFLE_P2P 2   /* fused partition + peer-memory scatter kernel */
int32_t rfgpu_last_shuffle_path(const rfgpu_ctx* ctx);
/* the destination rank of an id under the shuffle's hash (host mirror, for tests) */
uint32_t rfgpu_shuffle_dest(uint32_t id, uint32_t n_ranks);
/* partition only (no communication): rows grouped by destination, counts[n_parts] filled */
int32_t rfgpu_partition(rfgpu_ctx* ctx, const rfgpu_table* in, int32_t key_col, uint32_t n_parts,
                        rfgpu_table** out, uint64_t* counts);

/* ---- metrics: the BaselineMetrics analogue (stream.rs:43-61) + kernel evidence ---- */
typedef struct {
  uint64_t kernel_launches;   /* launches of rfgpu's own kernels since the last reset */
  uint64_t h2d_bytes, d2h_bytes;
  uint64_t scan_rows_in, scan_rows_out;
  uint64_t filter_rows_in, filter_rows_out;
  uint64_t join_build_rows, join_probe_rows, join_rows_out;
  uint64_t algorithmic_bytes; /* SURVEY §8(d) formulas, summed over operators */
} rfgpu_metrics;
int32_t rfgpu_metrics_get(const rfgpu_ctx* ctx, rfgpu_metrics* out);
int32_t rfgpu_metrics_reset(rfgpu_ctx* ctx);
/* per-kernel device time (CUDA events on the ctx stream) — off by default */
int32_t rfgpu_profile_enable(rfgpu_ctx* ctx, int32_t on);
int32_t rfgpu_profile_num_kernels(rfgpu_ctx* ctx);
int32_t rfgpu_profile_get(rfgpu_ctx* ctx, int32_t i, const char** name, uint64_t* launches,
                          double* total_ms, uint64_t* algorithmic_bytes);
/* stream time that passed between the end of the previously profiled kernel and the start of kernel i's launches, summed
 * over its launches: host hand-offs, memsets and copies in between — where a step's time goes besides the kernels */
int32_t rfgpu_profile_gap_ms(rfgpu_ctx* ctx, int32_t i, double* gap_ms);
/* launch timeline (with profiling enabled): rfgpu_profile_timeline(ctx, 1) sets the origin of the time axis on the ctx stream
 * and starts recording the start / end of every kernel launch in microseconds; _len / _get read the records back in launch
 * order (kernels of the range look-up stream overlap those of the ctx stream); rfgpu_profile_timeline(ctx, 0) stops */
int32_t rfgpu_profile_timeline(rfgpu_ctx* ctx, int32_t on);
int32_t rfgpu_profile_timeline_len(rfgpu_ctx* ctx);
int32_t rfgpu_profile_timeline_get(rfgpu_ctx* ctx, int32_t i, const char** name, double* start_us, double* end_us,
                                   double* queued_us /* optional: host clock when the launch returned, same origin */);
/* stream-ordered stopwatch on the ctx stream (bench.py times steps with it) */
int32_t rfgpu_timer_start(rfgpu_ctx* ctx);
int32_t rfgpu_timer_stop(rfgpu_ctx* ctx, double* elapsed_ms);
/* write a buffer larger than L2 (flush between timed iterations) */
int32_t rfgpu_flush_l2(rfgpu_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* RFGPU_H */
