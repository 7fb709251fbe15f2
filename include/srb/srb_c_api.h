/*
 * srb_c_api.h — C ABI of libsrb.so: the B200-native scan -> filter -> hash-aggregate / hash-join
 * operators that sit directly under the StarRocks BE pipeline Operator surface.
 *
 * The reference has no FFI for this path: operators are C++ classes linked into starrocks_be
 * (be/src/exec/pipeline/operator.h:43-344). This ABI is therefore the level one below that surface;
 * each entry point names the reference interface it stands in for. Host C++ operators that subclass
 * the mirrored Operator (include/srb/compat/operator.h) and the ctypes test harness are the callers.
 *
 * Conventions
 *  - every function returns an srb_status (values of TStatusCode, gensrc/thrift/StatusCode.thrift:53-83);
 *    SRB_END_OF_FILE from pull is "done", as Status::EndOfFile is in pipeline_driver.cpp:329.
 *  - no C++ types, no exceptions, no torch types cross this boundary; plain pointers and sizes only.
 *  - threading is the reference's: one thread at a time per srb_op (pipeline_driver.cpp:246), different
 *    ops concurrently. Each op owns its CUDA stream.
 *  - there is no CPU fallback anywhere behind this ABI: without a CUDA device srb_ctx_create fails.
 */
#ifndef SRB_C_API_H
#define SRB_C_API_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes: TStatusCode (gensrc/thrift/StatusCode.thrift:53-83) --------------------------- */
typedef int32_t srb_status;
#define SRB_OK 0
#define SRB_CANCELLED 1
#define SRB_NOT_IMPLEMENTED 3
#define SRB_RUNTIME_ERROR 4
#define SRB_MEM_LIMIT_EXCEEDED 5
#define SRB_INTERNAL_ERROR 6
#define SRB_END_OF_FILE 30
#define SRB_CORRUPTION 32
#define SRB_INVALID_ARGUMENT 33

/* ---- logical types: LogicalType (be/src/types/logical_type.h:27-77), same numeric values -------- */
#define SRB_TYPE_TINYINT 1
#define SRB_TYPE_SMALLINT 3
#define SRB_TYPE_INT 5
#define SRB_TYPE_BIGINT 7
#define SRB_TYPE_LARGEINT 9
#define SRB_TYPE_DOUBLE 11
#define SRB_TYPE_VARCHAR 17
#define SRB_TYPE_BOOLEAN 24
#define SRB_TYPE_VARBINARY 46
#define SRB_TYPE_DECIMAL32 47
#define SRB_TYPE_DECIMAL64 48
#define SRB_TYPE_DECIMAL128 49
#define SRB_TYPE_DATE 50 /* int32 Julian day, types/date_value.h:113 */

/* where a column's buffers live */
#define SRB_MEM_HOST 0   /* pageable host memory (std::vector-backed Buffer<T>, column/vectorized_fwd.h:47) */
#define SRB_MEM_PINNED 1 /* page-locked host memory */
#define SRB_MEM_DEVICE 2 /* device memory of the ctx's GPU. The library launches on streams of its own: whatever wrote the buffer on another
                            stream must have completed (stream / device synchronised by the caller) before the chunk is pushed. */

/*
 * One column of a chunk. Layouts are the reference's:
 *  - FixedLengthColumnBase<T>::get_data() flat T[n]        (column/fixed_length_column_base.h:213)
 *  - DecimalV3Column<T>: same + precision/scale            (column/decimalv3_column.h:27-59)
 *  - NullableColumn: data column + uint8 null[n], 1 = null (column/nullable_column.h:23-31)
 *  - BinaryColumn: bytes + uint32 offsets[n+1]             (column/binary_column.h:34-37)
 */
typedef struct srb_column {
    int32_t type; /* SRB_TYPE_* */
    int32_t precision;       /* decimals; CHAR / VARCHAR: the declared length (TypeDescriptor::len) when known, 0 otherwise — a GROUP BY
                                over such a column sizes its key area for it (serialized keys of up to 64 bytes; undeclared: 16) */
    int32_t scale;
    int32_t mem_kind;        /* SRB_MEM_* */
    const void* data;        /* T[n_rows]; for VARCHAR/VARBINARY the byte buffer */
    const uint8_t* nulls;    /* NULL when the column has no null map */
    const uint32_t* offsets; /* VARCHAR/VARBINARY only: n_rows + 1 entries */
    int64_t n_bytes;         /* VARCHAR/VARBINARY only: size of the byte buffer */
} srb_column;

/* Chunk (column/chunk.h:50-318): columns addressed by slot id. */
typedef struct srb_chunk {
    int64_t n_rows;
    int32_t n_cols;
    const int32_t* slot_ids;
    const srb_column* cols;
} srb_chunk;

/* ---- predicates: storage/column_predicate.h:66-84 (PredicateType values) -------------------------- */
#define SRB_PRED_EQ 1
#define SRB_PRED_NE 2
#define SRB_PRED_GT 3
#define SRB_PRED_GE 4
#define SRB_PRED_LT 5
#define SRB_PRED_LE 6
#define SRB_PRED_IS_NULL 9
#define SRB_PRED_NOT_NULL 10

/*
 * One `column <op> constant` predicate (ColumnPredicateCmpBase, storage/column_predicate_cmp.cpp:217-239).
 * A NULL row never satisfies a comparison (:233-237). Predicates that share or_group != 0 are OR-ed
 * (DisjunctivePredicates / PredicateTree OR node); the groups and all or_group == 0 predicates are
 * AND-ed in order (ConjunctivePredicates::evaluate, storage/conjunctive_predicates.cpp:35-52).
 */
typedef struct srb_predicate {
    int32_t slot_id;
    int32_t op;       /* SRB_PRED_* */
    int32_t type;     /* SRB_TYPE_* of the column; the literal is already in the column's representation */
    int32_t or_group; /* 0 = plain conjunct */
    int64_t lit_lo;   /* literal, low 64 bits (sign-extended into lit_hi for <=64-bit types) */
    int64_t lit_hi;
    /* VARCHAR / VARBINARY columns (BinaryColumnPredicateCmpBase, storage/column_predicate_cmp.cpp:559-700): the literal's bytes, compared
     * like Slice::compare (memcmp over the common length, then the shorter one is smaller). At most 16 bytes; NULL / 0 for other types. */
    const void* lit_bytes;
    int64_t lit_len;
} srb_predicate;

/* ---- scalar expressions: flattened Expr tree (exprs/expr.h, exprs/arithmetic_expr.cpp:66-128) ----- */
#define SRB_EXPR_SLOT_REF 1
#define SRB_EXPR_LITERAL 2
#define SRB_EXPR_CAST 3 /* decimal<->decimal / int->decimal casts that keep or raise the scale */
#define SRB_EXPR_ADD 4
#define SRB_EXPR_SUB 5
#define SRB_EXPR_MUL 6

/*
 * Node i may only reference nodes < i. Arithmetic on DECIMAL types follows
 * VectorizedArithmeticExpr::evaluate_checked (exprs/arithmetic_expr.cpp:103-128): a MUL whose children are
 * both same-scale casts of DECIMAL64 (or DECIMAL32) into DECIMAL128 takes the i64 x i64 -> i128 fast path
 * (:70-101, never overflows); every other decimal ADD/SUB/MUL runs in the result type with
 * OverflowMode::OUTPUT_NULL — an overflowing row becomes NULL (exprs/decimal_binary_function.h:94-109).
 * Integer (INT/BIGINT) arithmetic wraps (exprs/arithmetic_operation.h:127-134).
 */
typedef struct srb_expr_node {
    int32_t op;   /* SRB_EXPR_* */
    int32_t type; /* result SRB_TYPE_* */
    int32_t precision;
    int32_t scale;
    int32_t lhs;     /* node index; CAST uses lhs only */
    int32_t rhs;     /* node index or -1 */
    int32_t slot_id; /* SLOT_REF */
    int32_t _pad;
    int64_t lit_lo; /* LITERAL value in the node's own type/scale */
    int64_t lit_hi;
} srb_expr_node;

/* ---- aggregation: exec/aggregator.h:256-618, exprs/agg/{sum,avg,count,maxmin}.h -------------------- */
#define SRB_AGG_SUM 1
#define SRB_AGG_AVG 2
#define SRB_AGG_COUNT 3 /* count(expr): non-null rows; arg == -1 is count(*) (exprs/agg/count.h:36-120) */
#define SRB_AGG_MIN 4
#define SRB_AGG_MAX 5

typedef struct srb_agg_fn {
    int32_t fn;  /* SRB_AGG_* */
    int32_t arg; /* expr node index, -1 for count(*) */
} srb_agg_fn;

/* TAggregationOp phases as wired by AggregateBlockingNode (exec/aggregate/aggregate_blocking_node.cpp:178-254) */
#define SRB_AGG_UPDATE_FINALIZE 0  /* single phase: raw rows in, final values out */
#define SRB_AGG_UPDATE_SERIALIZE 1 /* phase 1: raw rows in, intermediate states out (aggregator.cpp:858) */
#define SRB_AGG_MERGE_FINALIZE 2   /* phase 2: intermediate states in (aggregator.cpp:871-875), final out */
#define SRB_AGG_MERGE_SERIALIZE 3

/*
 * A pipeline fragment scan-filter -> project -> aggregate, fused into one operator. Any stage may be empty:
 *  n_preds == 0          : no filtering
 *  n_aggs == 0, n_keys==0: a pure filter operator; pull returns the compacted input columns (Chunk::filter,
 *                          column/chunk.cpp:302) in `out_slots` order
 *  n_keys == 0, n_aggs>0 : no-group-by aggregation (Aggregator::compute_single_agg_state, aggregator.cpp:833)
 */
typedef struct srb_scan_agg_plan {
    int32_t n_preds;
    const srb_predicate* preds;
    int32_t n_nodes;
    const srb_expr_node* nodes;
    int32_t n_keys;
    const int32_t* key_nodes; /* group-by exprs: node indexes (SLOT_REFs) */
    int32_t n_aggs;
    const srb_agg_fn* aggs;
    int32_t phase;       /* SRB_AGG_* phase */
    int32_t n_out_slots; /* filter-only mode: columns to emit */
    const int32_t* out_slots;
    int64_t group_hint; /* expected number of groups (FE cardinality estimate); 0 = unknown */
} srb_scan_agg_plan;

/* ---- hash join: exec/hash_joiner.h:169, exec/join_hash_map.h:101-214 -------------------------------- */
#define SRB_JOIN_INNER 0     /* TJoinOp::INNER_JOIN */
#define SRB_JOIN_LEFT_SEMI 2 /* TJoinOp::LEFT_SEMI_JOIN (gensrc/thrift/PlanNodes.thrift:696-715: values count from 0) */
#define SRB_JOIN_LEFT_ANTI 8 /* TJoinOp::LEFT_ANTI_JOIN: probe rows without a match, NULL probe keys included (join_hash_map.tpp:1349-1420);
                                probe columns only */
/* Outer joins and the right-side existence joins (join_hash_map.tpp:1187-1986; dispatch :851-870). "left" = the probe side, "right" = the
 * build side, as in the reference. Columns of the side that may be missing come back nullable. RIGHT OUTER / FULL OUTER / RIGHT SEMI /
 * RIGHT ANTI remember which build rows matched across all probe chunks and emit the build-side remainder after
 * srb_join_probe_set_finishing (the probe side then has output once more; push an empty probe chunk first if no probe chunk ever
 * arrived, so that the probe columns' types are known). NULL keys never match: such probe rows are "unmatched" for LEFT / FULL OUTER,
 * such build rows for RIGHT / FULL OUTER and RIGHT ANTI. emit_indexes is ignored for these types. */
#define SRB_JOIN_LEFT_OUTER 1  /* every probe row; build columns NULL where nothing matched */
#define SRB_JOIN_RIGHT_OUTER 3 /* matches, then every build row that never matched with NULL probe columns */
#define SRB_JOIN_FULL_OUTER 4  /* both */
#define SRB_JOIN_RIGHT_SEMI 7  /* build rows with at least one match, build columns only */
#define SRB_JOIN_RIGHT_ANTI 9  /* build rows without a match, build columns only */
#define SRB_JOIN_NULL_AWARE_LEFT_ANTI 10 /* NOT IN (join_hash_map.tpp:1356-1383, hash_joiner.h:386-396): like LEFT ANTI, but a NULL probe key
                                            is not kept, and a NULL key anywhere on a non-empty build side means no row qualifies; against an
                                            empty build side every probe row is kept. Without other-conjuncts. */

/* One other-join-conjunct: a comparison between two columns of the joined row, e.g. `l.shipdate > o.orderdate` beside the equi-join keys
 * (HashJoiner::_process_other_conjunct, exec/hash_joiner.cpp:488-507: for an INNER join the conjuncts filter the joined chunk). Both slots
 * must be key or output columns of the plan (either side), of the same fixed-width integer / DATE / DECIMAL type; a NULL on either side
 * fails the comparison. op: SRB_PRED_EQ .. SRB_PRED_LE. INNER joins only (the outer / semi / anti variants re-derive matched flags from the
 * conjuncts, hash_joiner.cpp:446-486: not built). */
typedef struct srb_join_conjunct {
    int32_t left_slot;
    int32_t op;
    int32_t right_slot;
    int32_t _pad;
} srb_join_conjunct;

typedef struct srb_join_plan {
    int32_t join_type;
    int32_t n_keys; /* equi-join key pairs; all fixed-width, packed size <= 16 bytes */
    const int32_t* build_key_slots;
    const int32_t* probe_key_slots;
    int32_t n_build_out; /* build-side payload columns gathered for matched rows (late materialisation) */
    const int32_t* build_out_slots;
    int32_t n_probe_out;
    const int32_t* probe_out_slots;
    int32_t emit_indexes; /* 1: also emit probe_index/build_index UInt32 columns under slots -4 / -3
                             (HASH_JOIN_PROBE_INDEX_SLOT_ID / HASH_JOIN_BUILD_INDEX_SLOT_ID, column/chunk.h:52-58) */
    int32_t n_conjuncts;  /* other-join-conjuncts (at most 8), 0 = none */
    const srb_join_conjunct* conjuncts;
} srb_join_plan;

/* ---- TopN: exec/chunks_sorter_topn.cpp, exec/sorting/sorting.h:86-101 ------------------- */
#define SRB_TOPN_MAX_ROWS 1024 /* offset + limit the device TopN keeps (beyond that it is a full sort) */
typedef struct srb_topn_plan {
    int32_t n_keys; /* ORDER BY columns, most significant first; fixed-width types, composite key <= 32 bytes */
    int32_t _pad;
    const int32_t* key_slots;
    const uint8_t* is_asc;      /* per key: 1 ascending, 0 descending */
    const uint8_t* nulls_first; /* per key: 1 = NULLs first in the output whatever the direction (SortDesc::is_null_first) */
    int64_t offset;
    int64_t limit; /* rows [offset, offset + limit) of the sorted order are produced */
} srb_topn_plan;

/* ---- handles ------------------------------------------------------------------------------------- */
typedef struct srb_ctx srb_ctx; /* device, streams, scratch pools, optional NCCL communicator */
typedef struct srb_op srb_op;   /* one operator instance (one driver's operator) */

typedef struct srb_config {
    int32_t device;         /* CUDA device ordinal */
    int32_t chunk_size;     /* rows per pulled chunk; RuntimeState::chunk_size(), config.h:776 (4096) */
    int64_t staging_rows;   /* host-chunk staging batch before a launch (0 = default 4 Mi rows) */
    int32_t rank;           /* multi-GPU: this process's rank, 0 when single */
    int32_t world_size;     /* 1 when single */
    int64_t mem_limit;      /* bytes of device memory this ctx may hold (0 = unlimited) */
} srb_config;

srb_status srb_ctx_create(const srb_config* cfg, srb_ctx** out);
void srb_ctx_destroy(srb_ctx* ctx);
/* message of the last non-OK status returned on this thread ("" if none); never NULL */
const char* srb_last_error(void);
/* library/device facts for harnesses: fills n_sms, HBM bytes, compute capability (major*10+minor) */
srb_status srb_ctx_device_info(srb_ctx* ctx, int32_t* n_sms, int64_t* hbm_bytes, int32_t* sm_arch);
/* returns the ctx's cached (idle) device buffers to the driver (MemTracker release between queries; operators closed on this ctx
 * park their buffers in a per-ctx pool of at most 48 GB) */
srb_status srb_ctx_trim(srb_ctx* ctx);
/* harness diagnostics: how often this ctx reached cudaMalloc / cudaFree (a steady loop of operators should live off the pool) and for how
 * long, and the bytes parked in the pool now */
srb_status srb_ctx_alloc_stats(srb_ctx* ctx, int64_t* n_malloc, int64_t* n_free, int64_t* malloc_us, int64_t* free_us, int64_t* pooled_bytes);
/* the CUDA stream an op launches on (cudaStream_t as void*), so harnesses can time with events on it */
void* srb_op_stream(srb_op* op);

/*
 * Run-time compiled scan / aggregate kernels. A plan that none of the library's built-in kernel shapes matches is described to the same
 * kernel template at run time (NVRTC, in the background; the general kernel serves the operator until the compiled one is ready; results are
 * the same bits either way). The reference picks one of its ahead-of-time template instantiations per plan instead
 * (exec/aggregate/agg_hash_variant.h, exprs/agg/factory/): there is no reference entry point these replace.
 *   SRB_JIT = off | async (default) | sync     SRB_JIT_MIN_ROWS (default 4 Mi rows per batch)     SRB_JIT_CACHE_DIR (default ~/.cache/srb_jit)
 */
/* blocks until every compilation started so far has finished (benchmarks and tests: the next operator of a compiled shape uses it at once) */
srb_status srb_jit_wait(void);
/* process-wide counts: shapes compiled here, loaded from the disk cache, failed (the text of one failure in last_error), still compiling */
srb_status srb_jit_stats(int64_t* compiled, int64_t* from_disk_cache, int64_t* failed, int64_t* pending, char* last_error, int64_t last_error_len);
/* lowers `plan` against the column types of `first_chunk` (no data is read), generates its kernel shape and compiles it on the calling thread;
 * needs no GPU. SRB_OK: `description` says which shape; SRB_NOT_IMPLEMENTED: the plan is outside what the static kernel covers (the general
 * kernel runs it); SRB_INTERNAL_ERROR: the compiler's log is the error message. */
srb_status srb_jit_check_plan(const srb_scan_agg_plan* plan, const srb_chunk* first_chunk, char* description, int64_t description_len);

/* ---- operators ----------------------------------------------------------------------------------- */
/* replaces OlapScan predicate evaluation + Project + AggregateBlockingSink/Source (or the streaming pair)
 * for one driver: be/src/exec/pipeline/scan/olap_chunk_source.cpp:586-624,
 * pipeline/aggregate/aggregate_blocking_sink_operator.cpp:97-134, aggregate_blocking_source_operator.cpp:47-72 */
srb_status srb_scan_agg_create(srb_ctx* ctx, const srb_scan_agg_plan* plan, srb_op** out);

/* replaces HashJoinBuildOperator + HashJoinProbeOperator sharing one HashJoiner:
 * pipeline/hashjoin/hash_join_build_operator.cpp:40-213, hash_join_probe_operator.cpp:55-104.
 * Build side: srb_op_push_chunk(op,...) then srb_op_set_finishing(op). Probe side: srb_join_probe_* below. */
srb_status srb_join_create(srb_ctx* ctx, const srb_join_plan* plan, srb_op** out);
/* ChunksSorterTopn behind the operator contract: push_chunk takes host or device chunks (every column of the first chunk's schema
 * is carried; fixed-width columns only), set_finishing ends the input, one pull_chunk / pull_all returns rows
 * [offset, offset + limit) of the input sorted by the plan's keys, ties in arrival order. */
srb_status srb_topn_create(srb_ctx* ctx, const srb_topn_plan* plan, srb_op** out);

/* Operator::push_chunk (operator.h:121). Columns may be HOST/PINNED (staged + copied) or DEVICE (used in place,
 * must stay valid until the next srb_op_sync / set_finishing). n_rows is not limited to chunk_size. */
srb_status srb_op_push_chunk(srb_op* op, const srb_chunk* chunk);
/* ---- encoded segment pages straight into a scan operator -----------------------------------------------------------
 * What SegmentIterator / ScalarColumnIterator hand the scan on the CPU after decoding (segment_iterator.cpp:1411-1760,
 * FrameOfReferencePageDecoder::next_batch frame_of_reference_page.h:107-223, PlainPageDecoder::next_batch plain_page.h:133-276) is decoded
 * here on the device: the ENCODED page bytes are what crosses PCIe. One call = one row range of the table: for every column its data pages
 * covering exactly those n_rows rows, in order. encoding = EncodingTypePB (gensrc/proto/segment.proto:49-58): SRB_ENC_PLAIN ([u32 count]
 * [values]) or SRB_ENC_FOR (frame-of-reference, frame_of_reference_coding.h:78-101); fixed-width NOT NULL columns (type widths 1..16).
 * Page bytes may be HOST / PINNED (copied, overlapped with the previous call's kernels) or DEVICE. A corrupt page (footer / frame sizes
 * that do not fit) makes srb_op_set_finishing fail with SRB_CORRUPTION, as Status::Corruption does in the decoder's init(); its frames decode
 * to nothing and are never read out of bounds. DEVICE-resident page bytes need 16 readable bytes after the last page. */
#define SRB_ENC_PLAIN 2
#define SRB_ENC_FOR 7
typedef struct srb_page {
    int64_t offset; /* first byte of the page inside srb_page_column.data */
    int64_t bytes;
    int64_t n_rows; /* values in the page (the footer / header says the same) */
} srb_page;
typedef struct srb_page_column {
    int32_t type;      /* srb_logical_type */
    int32_t precision; /* decimals */
    int32_t scale;
    int32_t encoding;  /* SRB_ENC_* */
    const void* data;  /* the pages' bytes; pages need not be contiguous or ordered by offset */
    int64_t bytes;     /* length of data (+ at least 0 bytes of slack: the library pads its own copy) */
    int32_t mem_kind;  /* srb_mem_kind of data */
    int32_t n_pages;
    const srb_page* pages; /* in row order; their n_rows add up to the call's n_rows */
} srb_page_column;
srb_status srb_op_push_pages(srb_op* op, int32_t n_cols, const int32_t* slot_ids, const srb_page_column* cols, int64_t n_rows);
/* Operator::set_finishing (operator.h:66), idempotent */
srb_status srb_op_set_finishing(srb_op* op);
/* Operator::has_output / need_input / is_finished (operator.h:95-105): 1 / 0 */
int32_t srb_op_has_output(srb_op* op);
int32_t srb_op_need_input(srb_op* op);
int32_t srb_op_is_finished(srb_op* op);
/* Operator::pull_chunk (operator.h:118): at most chunk_size rows; buffers are library-owned pinned host memory
 * valid until the next pull/close on this op. Returns SRB_END_OF_FILE after the last chunk. */
srb_status srb_op_pull_chunk(srb_op* op, srb_chunk* out);
/* whole-result variant for harnesses: all remaining rows in one chunk (no chunk_size slicing) */
srb_status srb_op_pull_all(srb_op* op, srb_chunk* out);
/* device-resident variant for operator chaining on the GPU (scan/filter -> join probe -> aggregate without a host
 * round trip): all remaining rows, columns are SRB_MEM_DEVICE buffers owned by the op and valid until its next
 * push/pull/close. Fixed-width columns only. */
srb_status srb_op_pull_device(srb_op* op, srb_chunk* out);
/* Operator::set_cancelled (operator.h:84) / close (operator.h:92) */
srb_status srb_op_cancel(srb_op* op);
void srb_op_close(srb_op* op);
/* wait for everything this op has enqueued */
srb_status srb_op_sync(srb_op* op);
/* rows pushed / rows that passed the predicates / kernels launched so far (RuntimeProfile counters, operator.h:310-327) */
srb_status srb_op_counters(srb_op* op, int64_t* rows_in, int64_t* rows_selected, int64_t* kernel_launches);

/* device time (CUDA events on the op's stream) spent in the main scan/aggregate kernel(s), the number of timed batches,
 * how many of them ran a statically specialised (TMA-staged) kernel, and which shape matched (-1 = generic interpreter) */
srb_status srb_op_profile(srb_op* op, double* kernel_ms, int64_t* batches, int64_t* static_launches, int32_t* shape);

/* probe side of a join op: HashJoinProbeOperator::push_chunk / pull_chunk (hash_join_probe_operator.cpp:79-88).
 * The probe state is separate from the build operator state so that has_output/need_input keep the reference's
 * meaning on each side. */
srb_status srb_join_probe_push_chunk(srb_op* op, const srb_chunk* chunk);
srb_status srb_join_probe_set_finishing(srb_op* op);
int32_t srb_join_probe_has_output(srb_op* op);
int32_t srb_join_probe_need_input(srb_op* op);
int32_t srb_join_probe_is_finished(srb_op* op);
srb_status srb_join_probe_pull_chunk(srb_op* op, srb_chunk* out);
srb_status srb_join_probe_pull_all(srb_op* op, srb_chunk* out);
srb_status srb_join_probe_pull_device(srb_op* op, srb_chunk* out);

/* ---- streaming pre-aggregation ---------------------------------------------------------------------------------- */
/* Streaming pre-aggregation: the first phase of a two-phase GROUP BY in front of an exchange
 * (AggregateStreamingSinkOperator, be/src/exec/pipeline/aggregate/aggregate_streaming_sink_operator.cpp:73-307). `mode` is
 * TStreamingPreaggregationMode (gensrc/thrift/PlanNodes.thrift:689-694): SRB_STREAMING_AUTO runs the AggrAutoState machine
 * (aggregator.h:170-193) per pushed range, SRB_STREAMING_FORCE_STREAMING passes every row through in intermediate format
 * (Aggregator::output_chunk_by_streaming), SRB_STREAMING_FORCE_PREAGGREGATION aggregates every row (= the blocking sink).
 * Valid before the first push_chunk on an aggregate operator with GROUP BY keys in SRB_AGG_UPDATE_SERIALIZE; the output is the
 * pass-through rows followed by the table's groups, all in the plan's intermediate schema, available after set_finishing.
 * The result of phase 2 over that output does not depend on the mode. */
#define SRB_STREAMING_AUTO 0
#define SRB_STREAMING_FORCE_STREAMING 1
#define SRB_STREAMING_FORCE_PREAGGREGATION 2
#define SRB_STREAMING_LIMITED_MEM 3 /* _push_chunk_by_limited_memory (aggregate_streaming_sink_operator.cpp:311-319): AUTO while the group table
                                       is smaller than the limit below, FORCE_STREAMING once it has reached it (LimitedMemAggState::has_limited,
                                       exec/aggregator.h:209-212, 664-666); a limit of 0 never limits */
srb_status srb_op_set_streaming_mode(srb_op* op, int32_t mode);
/* the limit of SRB_STREAMING_LIMITED_MEM in bytes of group table (slots x slot size: what Aggregator::memory_usage counts for the hash map) */
srb_status srb_op_set_streaming_memory_limit(srb_op* op, int64_t bytes);
/* rows passed through so far, groups held by the table, current AggrAutoState (0 INIT_PREAGG, 1 ADJUST, 2 PASS_THROUGH,
 * 3 FORCE_PREAGG, 4 PREAGG, 5 SELECTIVE_PREAGG); any pointer may be NULL */
srb_status srb_op_streaming_stats(srb_op* op, int64_t* rows_passed_through, int64_t* groups_in_table, int32_t* auto_state);

/* ---- runtime filters: the join's build side narrows the probe side's scan ---------------------------------------
 * HashJoinBuildOperator::set_finishing creates partial runtime filters from the build keys (hash_join_build_operator.cpp:97-208,
 * hash_joiner.cpp:516-611): a RuntimeBloomFilter = SimdBlockFilter (exprs/runtime_filter.h:79-222, runtime_filter.cpp:26-35,114-123)
 * + min / max + has_null (runtime_filter.h:1192-1209). The probe-side scan evaluates them on its key column and drops rows that
 * cannot match (Operator::eval_runtime_bloom_filters, pipeline/operator.h:185; RuntimeBloomFilter::_t_evaluate, runtime_filter.h:
 * 1369-1422: selection[i] &= test(key[i]); NULL keys pass iff has_null). The directory holds exactly the bits the reference's
 * filter of the same size would hold for the same keys. Key types: integers up to 8 bytes, DATE, DECIMAL32/64. */
typedef struct srb_runtime_filter srb_runtime_filter;
/* bits of srb_join_build_runtime_filter's `hash_mix` argument */
#define SRB_RF_HASH_MIX 1     /* hash_mix = 1 below */
#define SRB_RF_MAYBE_BITSET 2 /* the reference's maybe_use_bitset_filter (exec/pipeline/runtime_filter_types.cpp:149-152): this builder holds the
                                 whole build side (a broadcast join), so RuntimeFilterHelper::create_join_runtime_filter's rule
                                 (exprs/runtime_filter_bank.cpp:160-205) may pick a RuntimeBitsetFilter — one bit per value of [min, max],
                                 exact membership (exprs/runtime_filter.h:1583-1700) — instead of the bloom filter: when the bitset is no
                                 larger than one byte per build row, or at most 4x that and fits the cache (here: the GPU's L2). Like the
                                 reference's, a bitset filter cannot be merged across ranks. Ignored with expected_rows > 0. */
#define SRB_RF_KIND_BLOOM 1   /* RuntimeFilterSerializeType (exprs/runtime_filter.h:46-51) */
#define SRB_RF_KIND_BITSET 3
/* `join` must have finished its build (srb_op_set_finishing). key_index: which build key column. expected_rows sizes the filter as
 * SimdBlockFilter::init(expected_rows) does (0: the rows this operator built from); ranks that will merge their filters pass the
 * same value. hash_mix = 0: the key's hash is RuntimeBloomFilter::compute_hash as the reference tree builds it — phmap_mix<8>
 * of std::hash, and phmap_mix is the identity unless ENABLE_PHMAP_MIX is defined (util/phmap/phmap_utils.h:72-124; nothing in the tree
 * defines it), so the directory is bit-identical to a BE-built filter and can be merged with one. hash_mix = 1: the multiplicative
 * phmap_mix<8> of the same header (:85-95). Which rows a filter lets through by accident differs between the two; no query result
 * does. Sparse key domains (TPC-H order keys use 8 of every 32 values) fill a quarter of the buckets under the identity hash.
 * The filter is independent of the operator afterwards; free it with srb_runtime_filter_destroy. */
srb_status srb_join_build_runtime_filter(srb_op* join, int32_t key_index, int64_t expected_rows, int32_t hash_mix, srb_runtime_filter** out);
/* attach to a scan-filter operator (a srb_scan_agg plan without aggregates) BEFORE its first srb_op_push_chunk: rows whose column
 * `slot_id` fails the filter are dropped together with those failing the plan's predicates. The filter must outlive the operator. */
srb_status srb_op_add_runtime_filter(srb_op* scan, const srb_runtime_filter* rf, int32_t slot_id);
srb_status srb_runtime_filter_info(const srb_runtime_filter* rf, int64_t* n_buckets, int64_t* min_value, int64_t* max_value, int32_t* has_null,
                                   int64_t* inserted);
/* copies the directory (n_buckets x 8 uint32) to host memory: what the tests compare with the oracle's SimdBlockFilter */
srb_status srb_runtime_filter_directory(const srb_runtime_filter* rf, void* dst, int64_t bytes);
/* which filter the probe side tests (SRB_RF_KIND_*), and for a bitset its size in bytes (whole 32-bit words over [min, max]) */
srb_status srb_runtime_filter_kind(const srb_runtime_filter* rf, int32_t* kind, int64_t* bitset_bytes);
/* copies the bitset to host memory: bit (x - min) of a little-endian bit string, as Bitset::insert sets it (exprs/runtime_filter.cpp:287-292) */
srb_status srb_runtime_filter_bitset(const srb_runtime_filter* rf, void* dst, int64_t bytes);
void srb_runtime_filter_destroy(srb_runtime_filter* rf);

/* ---- exchange: what replaces ExchangeSinkOperator / ExchangeSourceOperator + brpc between GPUs ---------------------- */
typedef struct srb_partitioner srb_partitioner;
typedef struct srb_comm srb_comm; /* one NCCL communicator over the ctx's rank / world_size */

/* HASH_PARTITIONED row routing (exchange_sink_operator.cpp:569-633): per-row FNV-1 chained over the partition columns
 * (fixed_length_column_base.cpp:223-227; NULL rows nullable_column.cpp:346-369), channel = (hash * n) >> 32
 * (shuffler.h:70-85). `in` is a device-resident chunk of fixed-width columns; `out` has the same columns with every
 * destination's rows contiguous, destination 0 first (library-owned device buffers, valid until the next run/destroy);
 * counts[d] = rows for destination d. */
srb_status srb_partition_create(srb_ctx* ctx, int32_t n_partitions, int32_t reserved, srb_partitioner** out);
srb_status srb_partition_run(srb_partitioner* p, const srb_chunk* in, int32_t n_key_slots, const int32_t* key_slots, srb_chunk* out,
                             int64_t* counts);
void srb_partition_destroy(srb_partitioner* p);

/* The same routing for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). NO COUNTERPART IN THE REFERENCE. The call allocates nothing (no
 * cudaMalloc / cudaFree / allocating library routine), never waits for the device and copies nothing to the host: the pass-1
 * scratch (one destination byte per row + 1.5 KB of totals / cursors / first rows) is a file-scope `static __device__` array of
 * 2 KB + SRB_PARTITION_NOALLOC_MAX_ROWS bytes that every GPU which loads the library holds until exit; larger chunks are rejected
 * with SRB_INVALID_ARGUMENT (split them, or use srb_partition_run). Everything is enqueued on `cuda_stream` (a cudaStream_t; NULL =
 * the legacy default stream) of device `device`, which becomes the calling thread's current device.
 *   in          device-resident chunk of fixed-width columns (<= 16), n_partitions <= 64, <= 4 partition columns
 *   out         the CALLER's device buffers: same n_cols / slot_ids / types as `in`, room for in->n_rows rows (out->n_rows >= in->n_rows),
 *               a null buffer for every column of `in` that has one; filled like srb_partition_run's `out` (destination 0 first)
 *   counts_dev  the caller's DEVICE array of n_partitions int64: rows per destination
 * Results are valid once the work enqueued on the stream has completed. All calls on one device share the one scratch array (the
 * call clears what must start at zero): calls whose kernels may overlap on the GPU — different streams without ordering, several
 * threads — race and must be serialised by the caller. */
#define SRB_PARTITION_NOALLOC_MAX_ROWS (1 << 20)
srb_status srb_partition_run_noalloc(int32_t device, void* cuda_stream, int32_t n_partitions, const srb_chunk* in, int32_t n_key_slots,
                                     const int32_t* key_slots, const srb_chunk* out, int64_t* counts_dev);

/* Cancellation that reaches in-flight collectives (PipelineDriver::cancel -> operators' set_cancelled, pipeline_driver.cpp:316; the brpc
 * exchange fails its pending RPCs): ncclCommAbort on this rank's communicator, callable from another thread while a collective is queued or
 * running. Afterwards every call on the communicator returns SRB_CANCELLED; destroy it. srb_comm_check reports the communicator's asynchronous
 * NCCL error state (a dead peer) without blocking. */
srb_status srb_comm_abort(srb_comm* c);
srb_status srb_comm_check(srb_comm* c);

/* set_finishing of a low-cardinality aggregate operator on EVERY rank of the communicator at once: the ranks' group tables (their raw
 * aggregate states) are all-gathered on the device and merged into every rank's table, then the operator finalises as usual — every rank
 * pulls the GLOBAL result. One call stands for AggregateBlockingSink (update, serialize) -> ExchangeSink / ExchangeSource (gather,
 * exchange_sink_operator.cpp:508-547) -> AggregateBlockingSink (merge, aggregator.cpp:871-875) -> finalize, without the intermediate rows
 * ever visiting the host: states are what serialize carries (sum.h:113-117, avg.h:161-181, count.h:92-96, maxmin.h), and integer states
 * merge exactly. Collective: every rank must call it, with the same plan. Tables of at most 1024 slots (a few hundred groups); a rank whose
 * table outgrew that makes the call fail with SRB_NOT_IMPLEMENTED on every rank (use the serialize -> srb_comm_all_gather -> merge operators
 * then). The plan's phase must be SRB_AGG_UPDATE_FINALIZE. */
srb_status srb_op_set_finishing_merged(srb_op* op, srb_comm* comm);

/* communicator bootstrap: rank 0 calls srb_comm_unique_id, the 128 bytes travel out of band (the harness uses
 * torch.distributed's store), every rank calls srb_comm_create with them. */
srb_status srb_comm_unique_id(void* out128);
srb_status srb_comm_create(srb_ctx* ctx, const void* id128, srb_comm** out);
void srb_comm_destroy(srb_comm* c);
void* srb_comm_stream(srb_comm* c);
/* the shuffle itself (replaces transmit_chunk over brpc, data_stream_sender/recvr): `send` rows are grouped by destination
 * rank with send_counts[world]; `recv` (library-owned device buffers, valid until the next call on this comm) holds the
 * rows from rank 0, 1, ... in that order with recv_counts[world]. Counts travel first (ncclAllGather), then one grouped
 * ncclSend/ncclRecv per column and peer. */
srb_status srb_comm_all_to_all(srb_comm* c, const srb_chunk* send, const int64_t* send_counts, srb_chunk* recv, int64_t* recv_counts);
/* Partitioning FUSED with the exchange (ExchangeSinkOperator::push_chunk HASH_PARTITIONED + transmit + ExchangeSourceOperator::pull_chunk,
 * exchange_sink_operator.cpp:569-633): rows of the device chunk `in` are routed by the same shuffle hash as srb_partition_run and the
 * scatter kernel stores each one straight into the destination rank's receive arena over NVLink peer memory (cudaIpc-mapped), so there
 * is no grouped staging copy and no send/recv; the W x W count matrix travels in one ncclAllGather, and a second small collective behind
 * the kernel publishes completion. `recv` / `recv_counts` as srb_comm_all_to_all (rows of source rank 0, 1, ... in that order; order inside
 * a source's range unspecified). Needs peer access between the GPUs; fails with INTERNAL_ERROR otherwise. */
srb_status srb_comm_shuffle(srb_comm* c, const srb_chunk* in, int32_t n_key_slots, const int32_t* key_slots, srb_chunk* recv,
                            int64_t* recv_counts);
/* global runtime filter: OR of every rank's partial filter (SimdBlockFilter::merge), min / max / has_null combined; every rank ends
 * up with the merged filter in place. All ranks must have built theirs with the same expected_rows. */
srb_status srb_comm_merge_runtime_filter(srb_comm* c, srb_runtime_filter* rf);
/* kernels launched by this communicator's fused shuffles so far */
srb_status srb_comm_counters(srb_comm* c, int64_t* kernel_launches);
/* UNPARTITIONED / broadcast exchange (exchange_sink_operator.cpp:508-547): every rank's rows replicated on every rank */
srb_status srb_comm_all_gather(srb_comm* c, const srb_chunk* mine, srb_chunk* out);

/* ---- device memory helpers for harnesses (plain cudaMalloc/cudaMemcpy wrappers) ------------------- */
srb_status srb_device_alloc(srb_ctx* ctx, int64_t bytes, void** out);
void srb_device_free(srb_ctx* ctx, void* p);
srb_status srb_pinned_alloc(srb_ctx* ctx, int64_t bytes, void** out);
void srb_pinned_free(srb_ctx* ctx, void* p);
srb_status srb_memcpy_h2d(srb_ctx* ctx, void* dst, const void* src, int64_t bytes);
srb_status srb_memcpy_d2h(srb_ctx* ctx, void* dst, const void* src, int64_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* SRB_C_API_H */
