// include/srb/gpu_operators.h — the C++ host side above the C ABI: pipeline operators with the reference's Operator
// interface (be/src/exec/pipeline/operator.h:43-344) that forward to libsrb.so. These are the classes INTEGRATION.md shows
// a BE maintainer adding next to AggregateBlockingSink/SourceOperator (exec/pipeline/aggregate/) and HashJoinBuild/
// ProbeOperator (exec/pipeline/hashjoin/); here they compile against the minimal mirror in srb/compat/ and are driven by
// tests/cpp/test_operators.cpp. No computation happens on this side: it converts Chunk <-> srb_chunk and statuses.
#pragma once
#include <cstring>
#include <memory>
#include <vector>

#include "compat/operator.h"
#include "srb_c_api.h"

namespace starrocks::pipeline {

inline Status srb_to_status(srb_status st) { return st == SRB_OK ? Status::OK() : Status((int)st, srb_last_error()); }

// ---- Chunk -> srb_chunk (no copy: the descriptors point into the columns' own buffers) ---------------------------------------
struct SrbChunkView {
    std::vector<srb_column> cols;
    std::vector<int32_t> slots;
    srb_chunk chunk{};
};
inline void to_srb(const ColumnPtr& col, srb_column* c) {
    std::memset(c, 0, sizeof(*c));
    const Column* data = col.get();
    if (col->is_nullable()) {
        auto* n = static_cast<const NullableColumn*>(col.get());
        if (n->has_null()) c->nulls = n->null_column()->get_data().data(); // a null map without NULLs is not passed (has_null())
        data = n->data_column().get();
    }
    c->type = data->logical_type();
    c->precision = data->precision();
    c->scale = data->scale();
    c->mem_kind = SRB_MEM_HOST;
    c->data = data->raw_data();
    if (data->is_binary()) {
        auto* b = static_cast<const BinaryColumn*>(data);
        c->offsets = b->get_offset().data();
        c->n_bytes = (int64_t)b->get_bytes().size();
    }
}
inline void to_srb(const Chunk& chunk, SrbChunkView* v) {
    v->cols.resize(chunk.num_columns());
    v->slots = chunk.slots_in_order();
    for (size_t i = 0; i < chunk.num_columns(); i++) to_srb(chunk.get_column_by_index(i), &v->cols[i]);
    v->chunk.n_rows = (int64_t)chunk.num_rows();
    v->chunk.n_cols = (int32_t)v->cols.size();
    v->chunk.slot_ids = v->slots.data();
    v->chunk.cols = v->cols.data();
}

// ---- srb_chunk -> Chunk (copies: the library's buffers are only valid until the operator's next pull / close) ----------------
template <typename T, typename ColT>
inline ColumnPtr srb_copy_fixed(std::shared_ptr<ColT> col, const srb_column& c, int64_t n) {
    col->get_data().resize((size_t)n);
    if (n) std::memcpy(col->get_data().data(), c.data, (size_t)n * sizeof(T));
    return col;
}
inline StatusOr<ColumnPtr> from_srb(const srb_column& c, int64_t n) {
    ColumnPtr data;
    switch (c.type) {
    case SRB_TYPE_BOOLEAN:
    case SRB_TYPE_TINYINT: data = srb_copy_fixed<int8_t>(std::make_shared<FixedLengthColumn<int8_t>>(c.type), c, n); break;
    case SRB_TYPE_SMALLINT: data = srb_copy_fixed<int16_t>(std::make_shared<FixedLengthColumn<int16_t>>(c.type), c, n); break;
    case SRB_TYPE_INT:
    case SRB_TYPE_DATE: data = srb_copy_fixed<int32_t>(std::make_shared<Int32Column>(c.type), c, n); break;
    case SRB_TYPE_BIGINT: data = srb_copy_fixed<int64_t>(std::make_shared<Int64Column>(c.type), c, n); break;
    case SRB_TYPE_DOUBLE: data = srb_copy_fixed<double>(std::make_shared<FixedLengthColumn<double>>(c.type), c, n); break;
    case SRB_TYPE_LARGEINT: data = srb_copy_fixed<int128_t>(std::make_shared<FixedLengthColumn<int128_t>>(c.type), c, n); break;
    case SRB_TYPE_DECIMAL32: data = srb_copy_fixed<int32_t>(std::make_shared<DecimalV3Column<int32_t>>(c.precision, c.scale), c, n); break;
    case SRB_TYPE_DECIMAL64: data = srb_copy_fixed<int64_t>(std::make_shared<DecimalV3Column<int64_t>>(c.precision, c.scale), c, n); break;
    case SRB_TYPE_DECIMAL128: data = srb_copy_fixed<int128_t>(std::make_shared<DecimalV3Column<int128_t>>(c.precision, c.scale), c, n); break;
    case SRB_TYPE_VARCHAR:
    case SRB_TYPE_VARBINARY: {
        auto b = std::make_shared<BinaryColumn>(c.type);
        b->get_offset().assign(c.offsets, c.offsets + n + 1);
        const uint8_t* p = (const uint8_t*)c.data;
        b->get_bytes().assign(p, p + c.offsets[n]);
        data = b;
        break;
    }
    default: return Status::NotSupported("column type " + std::to_string(c.type));
    }
    if (!c.nulls) return data;
    auto nulls = std::make_shared<NullColumn>(SRB_TYPE_BOOLEAN);
    nulls->get_data().assign(c.nulls, c.nulls + n);
    return ColumnPtr(std::make_shared<NullableColumn>(data, nulls));
}
inline StatusOr<ChunkPtr> from_srb(const srb_chunk& ch) {
    auto out = std::make_shared<Chunk>();
    for (int i = 0; i < ch.n_cols; i++) {
        auto col = from_srb(ch.cols[i], ch.n_rows);
        if (!col.ok()) return col.status();
        out->append_column(col.value(), ch.slot_ids[i]);
    }
    return out;
}

// A plan whose arrays the operator owns (the factory lowers TPlanNode into this once; aggregate_blocking_node.cpp:178-254).
struct ScanAggPlan {
    std::vector<srb_predicate> preds;
    std::vector<srb_expr_node> nodes;
    std::vector<int32_t> key_nodes;
    std::vector<srb_agg_fn> aggs;
    std::vector<int32_t> out_slots;
    int32_t phase = SRB_AGG_UPDATE_FINALIZE;
    int64_t group_hint = 0;
    srb_scan_agg_plan c_plan() const {
        srb_scan_agg_plan p{};
        p.n_preds = (int32_t)preds.size();
        p.preds = preds.data();
        p.n_nodes = (int32_t)nodes.size();
        p.nodes = nodes.data();
        p.n_keys = (int32_t)key_nodes.size();
        p.key_nodes = key_nodes.data();
        p.n_aggs = (int32_t)aggs.size();
        p.aggs = aggs.data();
        p.phase = phase;
        p.n_out_slots = (int32_t)out_slots.size();
        p.out_slots = out_slots.data();
        p.group_hint = group_hint;
        return p;
    }
};

// scan-side conjuncts + projection + (blocking) aggregation fused into one operator: takes the place of
// AggregateBlockingSinkOperator + AggregateBlockingSourceOperator of one pipeline pair, or of a filter (Chunk::filter) when
// the plan has no aggregates.
class GpuScanAggregateOperator final : public Operator {
public:
    GpuScanAggregateOperator(srb_ctx* ctx, ScanAggPlan plan) : Operator("gpu_scan_aggregate"), _ctx(ctx), _plan(std::move(plan)) {}
    ~GpuScanAggregateOperator() override { close(nullptr); }

    Status prepare(RuntimeState* state) override {
        if (_op) return Status::OK();
        const srb_scan_agg_plan p = _plan.c_plan();
        SRB_RETURN_IF_ERROR(srb_to_status(srb_scan_agg_create(_ctx, &p, &_op)));
        if (_streaming_mode >= 0) {
            SRB_RETURN_IF_ERROR(srb_to_status(srb_op_set_streaming_mode(_op, _streaming_mode)));
            if (_streaming_mode == SRB_STREAMING_LIMITED_MEM) return srb_to_status(srb_op_set_streaming_memory_limit(_op, _streaming_memory_limit));
        }
        return Status::OK();
    }
    // AggregateStreamingSinkOperator in place of the blocking sink (the first phase of a two-phase aggregation in front of an
    // exchange): TAggregationNode::streaming_preaggregation_mode (gensrc/thrift/PlanNodes.thrift:851) picks how rows meet the table.
    // Call before prepare(); the plan must be in SRB_AGG_UPDATE_SERIALIZE with GROUP BY keys.
    void set_streaming_preaggregation_mode(int32_t mode) { _streaming_mode = mode; }
    // LimitedMemAggState::limited_memory_size (exec/aggregator.h:209-212) for SRB_STREAMING_LIMITED_MEM: bytes of group table
    void set_streaming_memory_limit(int64_t bytes) { _streaming_memory_limit = bytes; }
    // the scan waits for the joins whose runtime filters it applies (Operator::set_precondition_ready, operator.h:185): no input is
    // taken before every such build side has finished; the filters are attached right before the first chunk
    void wait_for_runtime_filter(std::shared_ptr<class GpuHashJoiner> joiner, int32_t slot_id);
    bool need_input() const override { return _op && runtime_filters_ready() && srb_op_need_input(_op); }
    bool has_output() const override { return _op && srb_op_has_output(_op); }
    bool is_finished() const override { return !_op || srb_op_is_finished(_op); }
    Status push_chunk(RuntimeState* state, const ChunkPtr& chunk) override {
        if (state && state->is_cancelled()) return Status::Cancelled("query cancelled"); // aggregate_blocking_sink_operator.cpp:63
        SRB_RETURN_IF_ERROR(attach_runtime_filters());
        SrbChunkView v;
        to_srb(*chunk, &v);
        return srb_to_status(srb_op_push_chunk(_op, &v.chunk)); // host columns are staged into HBM before this returns control of them
    }
    Status set_finishing(RuntimeState*) override { return srb_to_status(srb_op_set_finishing(_op)); }
    Status set_cancelled(RuntimeState*) override { return srb_to_status(srb_op_cancel(_op)); }
    StatusOr<ChunkPtr> pull_chunk(RuntimeState*) override {
        srb_chunk out{};
        const srb_status st = srb_op_pull_chunk(_op, &out);
        if (st == SRB_END_OF_FILE) return Status::EndOfFile("gpu aggregate drained");
        if (st != SRB_OK) return srb_to_status(st);
        return from_srb(out);
    }
    void close(RuntimeState*) override {
        if (_op) {
            _rows_selected = rows_selected(); // the profile counter outlives the device state (RuntimeProfile is read after close)
            srb_op_close(_op);
        }
        _op = nullptr;
    }
    int64_t rows_selected() const {
        if (!_op) return _rows_selected;
        int64_t in = 0, sel = 0, launches = 0;
        srb_op_counters(_op, &in, &sel, &launches);
        return sel;
    }

private:
    bool runtime_filters_ready() const;
    Status attach_runtime_filters();
    srb_ctx* _ctx;
    ScanAggPlan _plan;
    srb_op* _op = nullptr;
    std::vector<std::pair<std::shared_ptr<class GpuHashJoiner>, int32_t>> _rf_sources;
    bool _rf_attached = false;
    int64_t _rows_selected = 0;
    int32_t _streaming_mode = -1; // -1: blocking sink
    int64_t _streaming_memory_limit = 0;
};

// ---- TopN: PartitionSortSinkOperator + LocalMergeSortSourceOperator over a ChunksSorterTopn, as one operator ---------------------
// (exec/pipeline/sort/partition_sort_sink_operator.cpp, exec/chunks_sorter_topn.cpp; TSortInfo::is_asc_order / nulls_first,
// TSortNode::offset / limit)
struct TopNPlan {
    std::vector<int32_t> key_slots;
    std::vector<uint8_t> is_asc, nulls_first;
    int64_t offset = 0, limit = 0;
    srb_topn_plan c_plan() const {
        srb_topn_plan p{};
        p.n_keys = (int32_t)key_slots.size();
        p.key_slots = key_slots.data();
        p.is_asc = is_asc.data();
        p.nulls_first = nulls_first.data();
        p.offset = offset;
        p.limit = limit;
        return p;
    }
};

class GpuTopNOperator final : public Operator {
public:
    GpuTopNOperator(srb_ctx* ctx, TopNPlan plan) : Operator("gpu_topn"), _ctx(ctx), _plan(std::move(plan)) {}
    ~GpuTopNOperator() override { close(nullptr); }
    Status prepare(RuntimeState*) override {
        if (_op) return Status::OK();
        const srb_topn_plan p = _plan.c_plan();
        return srb_to_status(srb_topn_create(_ctx, &p, &_op));
    }
    bool need_input() const override { return _op && srb_op_need_input(_op); }
    bool has_output() const override { return _op && srb_op_has_output(_op); }
    bool is_finished() const override { return !_op || srb_op_is_finished(_op); }
    Status push_chunk(RuntimeState* state, const ChunkPtr& chunk) override {
        if (state && state->is_cancelled()) return Status::Cancelled("query cancelled");
        SrbChunkView v;
        to_srb(*chunk, &v);
        return srb_to_status(srb_op_push_chunk(_op, &v.chunk)); // the survivors' rows are copied into operator state before this returns
    }
    Status set_finishing(RuntimeState*) override { return srb_to_status(srb_op_set_finishing(_op)); }
    StatusOr<ChunkPtr> pull_chunk(RuntimeState*) override {
        srb_chunk out{};
        const srb_status st = srb_op_pull_chunk(_op, &out);
        if (st == SRB_END_OF_FILE) return Status::EndOfFile("gpu topn drained");
        if (st != SRB_OK) return srb_to_status(st);
        return from_srb(out);
    }
    void close(RuntimeState*) override {
        if (_op) srb_op_close(_op);
        _op = nullptr;
    }

private:
    srb_ctx* _ctx;
    TopNPlan _plan;
    srb_op* _op = nullptr;
};

// ---- hash join: the build and probe operators share one joiner, like HashJoiner (exec/hash_joiner.h:169) -------------------------
struct JoinPlan {
    int32_t join_type = SRB_JOIN_INNER;
    std::vector<int32_t> build_key_slots, probe_key_slots, build_out_slots, probe_out_slots;
    bool emit_indexes = false;
    std::vector<srb_join_conjunct> conjuncts; // HashJoiner::_other_join_conjunct_ctxs of an INNER join as column-to-column comparisons
    srb_join_plan c_plan() const {
        srb_join_plan p{};
        p.join_type = join_type;
        p.n_keys = (int32_t)build_key_slots.size();
        p.build_key_slots = build_key_slots.data();
        p.probe_key_slots = probe_key_slots.data();
        p.n_build_out = (int32_t)build_out_slots.size();
        p.build_out_slots = build_out_slots.data();
        p.n_probe_out = (int32_t)probe_out_slots.size();
        p.probe_out_slots = probe_out_slots.data();
        p.emit_indexes = emit_indexes ? 1 : 0;
        p.n_conjuncts = (int32_t)conjuncts.size();
        p.conjuncts = conjuncts.empty() ? nullptr : conjuncts.data();
        return p;
    }
};
class GpuHashJoiner {
public:
    GpuHashJoiner(srb_ctx* ctx, JoinPlan plan) : _ctx(ctx), _plan(std::move(plan)) {}
    ~GpuHashJoiner() {
        if (_op) srb_op_close(_op);
        if (_rf) srb_runtime_filter_destroy(_rf);
    }
    Status prepare() {
        if (_op) return Status::OK();
        const srb_join_plan p = _plan.c_plan();
        return srb_to_status(srb_join_create(_ctx, &p, &_op));
    }
    srb_op* op() const { return _op; }
    bool build_done = false; // HashJoinPhase::BUILD -> PROBE (hash_joiner.h:60-65)

    // HashJoiner::create_runtime_filters (hash_joiner.cpp:516-611): called by the build operator's set_finishing when the plan
    // asks for a runtime filter on build key `runtime_filter_key` (-1: none). The probe-side scan picks it up with runtime_filter().
    int runtime_filter_key = -1;
    // PartialRuntimeFilterMerger's maybe_use_bitset_filter (exec/pipeline/runtime_filter_types.cpp:149-152): set for broadcast joins, where
    // this joiner holds the whole build side and create_join_runtime_filter's rule may pick a RuntimeBitsetFilter
    bool maybe_use_bitset_filter = false;
    Status create_runtime_filter() {
        if (runtime_filter_key < 0 || _rf) return Status::OK();
        return srb_to_status(srb_join_build_runtime_filter(_op, runtime_filter_key, 0, maybe_use_bitset_filter ? SRB_RF_MAYBE_BITSET : 0, &_rf));
    }
    const srb_runtime_filter* runtime_filter() const { return _rf; }

private:
    srb_ctx* _ctx;
    JoinPlan _plan;
    srb_op* _op = nullptr;
    srb_runtime_filter* _rf = nullptr;
};
using GpuHashJoinerPtr = std::shared_ptr<GpuHashJoiner>;

inline void GpuScanAggregateOperator::wait_for_runtime_filter(GpuHashJoinerPtr joiner, int32_t slot_id) { _rf_sources.emplace_back(std::move(joiner), slot_id); }
inline bool GpuScanAggregateOperator::runtime_filters_ready() const {
    for (auto& s : _rf_sources)
        if (!s.first->build_done) return false;
    return true;
}
inline Status GpuScanAggregateOperator::attach_runtime_filters() {
    if (_rf_attached) return Status::OK();
    _rf_attached = true;
    for (auto& s : _rf_sources)
        if (s.first->runtime_filter()) SRB_RETURN_IF_ERROR(srb_to_status(srb_op_add_runtime_filter(_op, s.first->runtime_filter(), s.second)));
    return Status::OK();
}

class GpuHashJoinBuildOperator final : public Operator { // hashjoin/hash_join_build_operator.cpp:40-213 (a sink)
public:
    explicit GpuHashJoinBuildOperator(GpuHashJoinerPtr j) : Operator("gpu_hash_join_build"), _j(std::move(j)) {}
    Status prepare(RuntimeState*) override { return _j->prepare(); }
    bool need_input() const override { return !_j->build_done; }
    bool has_output() const override { return false; }
    bool is_finished() const override { return _j->build_done; }
    Status push_chunk(RuntimeState*, const ChunkPtr& chunk) override {
        SrbChunkView v;
        to_srb(*chunk, &v);
        return srb_to_status(srb_op_push_chunk(_j->op(), &v.chunk));
    }
    Status set_finishing(RuntimeState*) override {
        if (_j->build_done) return Status::OK();
        SRB_RETURN_IF_ERROR(srb_to_status(srb_op_set_finishing(_j->op()))); // JoinHashTable::build (join_hash_map.cpp:529)
        SRB_RETURN_IF_ERROR(_j->create_runtime_filter());                     // hash_join_build_operator.cpp:97-208
        _j->build_done = true;
        return Status::OK();
    }
    StatusOr<ChunkPtr> pull_chunk(RuntimeState*) override { return Status::InternalError("build operator is a sink"); }

private:
    GpuHashJoinerPtr _j;
};

class GpuHashJoinProbeOperator final : public Operator { // hashjoin/hash_join_probe_operator.cpp:55-104
public:
    explicit GpuHashJoinProbeOperator(GpuHashJoinerPtr j) : Operator("gpu_hash_join_probe"), _j(std::move(j)) {}
    Status prepare(RuntimeState*) override { return _j->prepare(); }
    bool need_input() const override { return _j->build_done && srb_join_probe_need_input(_j->op()); } // blocked until the build side is done
    bool has_output() const override { return _j->build_done && srb_join_probe_has_output(_j->op()); }
    bool is_finished() const override { return _j->build_done && srb_join_probe_is_finished(_j->op()); }
    Status push_chunk(RuntimeState*, const ChunkPtr& chunk) override {
        SrbChunkView v;
        to_srb(*chunk, &v);
        return srb_to_status(srb_join_probe_push_chunk(_j->op(), &v.chunk));
    }
    Status set_finishing(RuntimeState*) override { return srb_to_status(srb_join_probe_set_finishing(_j->op())); }
    StatusOr<ChunkPtr> pull_chunk(RuntimeState*) override {
        srb_chunk out{};
        const srb_status st = srb_join_probe_pull_chunk(_j->op(), &out);
        if (st == SRB_END_OF_FILE) return Status::EndOfFile("gpu probe drained");
        if (st != SRB_OK) return srb_to_status(st);
        return from_srb(out);
    }

private:
    GpuHashJoinerPtr _j;
};

} // namespace starrocks::pipeline
