// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_arith.h header).
// Driver of the chunk-at-a-time scan -> filter -> project -> aggregate pipeline and its result emission:
//   PipelineDriver::process pull/push loop, rows <= chunk_size     be/src/exec/pipeline/pipeline_driver.cpp:286-373
//   Aggregator::convert_to_chunk_no_groupby (empty input => NULL)   be/src/exec/aggregator.cpp:924-959
//   Aggregator::convert_hash_map_to_chunk                           be/src/exec/aggregator.cpp:1421-1514
//   AvgAggregateFunction::finalize_to_column / serialize_to_column  be/src/exprs/agg/avg.h:168-245
//   DOP threads with private tables + merge                         be/src/exec/aggregate/aggregate_blocking_node.cpp:256-300
#include "orc_agg.h"

#include <thread>

namespace orc {

static void merge_intermediate_rows(Aggregator& ag, const srb_chunk& ch, int64_t base, size_t rows) {
    // phase 2: every row is (keys..., intermediate states...); no predicates (aggregator.cpp:871-875)
    ag.rows_in += (int64_t)rows;
    ag.rows_selected += (int64_t)rows;
    std::vector<uint32_t> idx(rows);
    for (size_t i = 0; i < rows; i++) idx[i] = (uint32_t)i;
    const srb_scan_agg_plan* plan = ag.plan;
    ExprEval ev(plan->nodes, plan->n_nodes, &ch, base, &idx);
    std::vector<const Col*> keys(plan->n_keys), args(ag.n_aggs, nullptr);
    for (int k = 0; k < plan->n_keys; k++) keys[k] = &ev.eval(plan->key_nodes[k]);
    for (int a = 0; a < ag.n_aggs; a++) args[a] = &ev.eval(plan->aggs[a].arg);
    if (!ag.arg_meta_known) {
        for (int a = 0; a < ag.n_aggs; a++) {
            const srb_expr_node& nd = plan->nodes[plan->aggs[a].arg];
            if (plan->aggs[a].fn == SRB_AGG_AVG) { // blob column: node.precision carries the original arg type
                ag.arg_type[a] = nd.precision;
                ag.arg_scale[a] = nd.scale;
            } else {
                ag.arg_type[a] = args[a]->type;
                ag.arg_scale[a] = args[a]->scale;
            }
        }
        ag.arg_meta_known = true;
    }
    if (!ag.key_meta_known) {
        for (int k = 0; k < plan->n_keys; k++)
            ag.key_meta.push_back({keys[k]->type, keys[k]->precision, keys[k]->scale, keys[k]->nullable});
        ag.key_meta_known = true;
    }
    std::string kb;
    for (size_t i = 0; i < rows; i++) {
        size_t g = 0;
        if (plan->n_keys > 0) {
            kb.clear();
            for (int k = 0; k < plan->n_keys; k++) {
                const Col& c = *keys[k];
                bool isnull = c.is_null(i);
                if (ag.key_meta[k].nullable) kb.push_back((char)isnull);
                if (is_binary(c.type)) {
                    uint32_t len = isnull ? 0 : (uint32_t)c.s[i].size();
                    kb.append((const char*)&len, 4);
                    if (!isnull) kb.append(c.s[i]);
                } else if (c.type == SRB_TYPE_DOUBLE) {
                    double x = isnull ? 0 : c.d[i];
                    kb.append((const char*)&x, 8);
                } else {
                    int128_t x = isnull ? 0 : c.v[i];
                    kb.append((const char*)&x, type_width(c.type));
                }
            }
            bool created;
            g = ag.table.emplace(kb, &created);
            if (created) ag.states.resize(ag.states.size() + ag.n_aggs);
        }
        for (int a = 0; a < ag.n_aggs; a++) {
            AggState& st = ag.states[g * ag.n_aggs + a];
            const Col& c = *args[a];
            int fn = plan->aggs[a].fn;
            if (c.is_null(i)) continue;
            switch (fn) {
            case SRB_AGG_COUNT: st.n += (int64_t)c.v[i]; break; // count.h:92-96
            case SRB_AGG_SUM:                                   // sum.h:113-117
                if (c.type == SRB_TYPE_DOUBLE)
                    st.dsum += c.d[i];
                else
                    st.isum = (int128_t)((uint128_t)st.isum + (uint128_t)c.v[i]);
                st.n++;
                break;
            case SRB_AGG_AVG: { // avg.h:161-166: {ImmediateType sum; int64 count}
                const std::string& b = c.s[i];
                int64_t cnt;
                if (b.size() == 24) {
                    int128_t s;
                    memcpy(&s, b.data(), 16);
                    memcpy(&cnt, b.data() + 16, 8);
                    st.isum = (int128_t)((uint128_t)st.isum + (uint128_t)s);
                } else {
                    double s;
                    memcpy(&s, b.data(), 8);
                    memcpy(&cnt, b.data() + 8, 8);
                    st.dsum += s;
                }
                st.n += cnt;
                break;
            }
            default: Aggregator::update_one(fn, st, &c, i, false);
            }
        }
    }
}

void Aggregator::emit(Result* r) const {
    const bool finalize = plan->phase == SRB_AGG_UPDATE_FINALIZE || plan->phase == SRB_AGG_MERGE_FINALIZE;
    size_t ng = plan->n_keys ? table.keys.size() : 1;
    r->rows_in = rows_in;
    r->rows_selected = rows_selected;
    r->n_rows = (int64_t)ng;
    int nk = plan->n_keys;
    r->cols.resize(nk + n_aggs);
    for (int k = 0; k < nk; k++) {
        OutCol& c = r->cols[k];
        r->slot_ids.push_back(k);
        if (key_meta_known) {
            c.type = key_meta[k].type;
            c.precision = key_meta[k].precision;
            c.scale = key_meta[k].scale;
        } else {
            c.type = plan->nodes[plan->key_nodes[k]].type;
        }
    }
    // keys: walk the groups in insertion order (the reference walks its state arena, aggregator.cpp:1443-1449)
    for (size_t g = 0; g < (nk ? ng : 0); g++) {
        const std::string& kb = table.keys[g];
        size_t off = 0;
        for (int k = 0; k < nk; k++) {
            OutCol& c = r->cols[k];
            bool isnull = false;
            if (key_meta[k].nullable) isnull = kb[off++] != 0;
            if (key_meta[k].nullable) {
                c.nulls.push_back(isnull);
                c.has_nulls = true;
            }
            if (is_binary(c.type)) {
                uint32_t len;
                memcpy(&len, kb.data() + off, 4);
                off += 4;
                c.push_bytes(kb.data() + off, len);
                off += len;
            } else {
                int w = type_width(c.type);
                size_t o = c.data.size();
                c.data.resize(o + w);
                memcpy(c.data.data() + o, kb.data() + off, w);
                off += w;
            }
        }
    }
    for (int a = 0; a < n_aggs; a++) {
        OutCol& c = r->cols[nk + a];
        r->slot_ids.push_back(nk + a);
        int fn = plan->aggs[a].fn;
        int at = arg_type[a], as = arg_scale[a];
        if (!arg_meta_known && plan->aggs[a].arg >= 0) { // no row ever arrived: types from the plan
            const srb_expr_node& nd = plan->nodes[plan->aggs[a].arg];
            at = nd.type;
            as = nd.scale;
            if (merge_input && fn == SRB_AGG_AVG) at = nd.precision;
        }
        int ft, fp, fs;
        agg_result_type(fn, at, as, &ft, &fp, &fs);
        bool avg_blob = !finalize && fn == SRB_AGG_AVG;
        if (avg_blob) {
            c.type = SRB_TYPE_VARBINARY;
            c.precision = at; // carries the arg type / scale to the merge phase
            c.scale = as;
        } else if (!finalize && fn == SRB_AGG_SUM) {
            c.type = ft;
            c.precision = fp;
            c.scale = fs;
        } else {
            c.type = ft;
            c.precision = fp;
            c.scale = fs;
        }
        bool nullable_result = fn != SRB_AGG_COUNT;
        for (size_t g = 0; g < ng; g++) {
            const AggState& st = states[g * n_aggs + a];
            bool isnull = nullable_result && st.n == 0;
            if (nullable_result) {
                c.nulls.push_back(isnull);
                c.has_nulls = true;
            }
            bool dbl_arg = at == SRB_TYPE_DOUBLE;
            switch (fn) {
            case SRB_AGG_COUNT: c.push_int(st.n); break;
            case SRB_AGG_SUM:
                if (dbl_arg)
                    c.push_double(isnull ? 0 : st.dsum);
                else
                    c.push_int(isnull ? 0 : st.isum);
                break;
            case SRB_AGG_AVG:
                if (avg_blob) {
                    if (Aggregator::avg_double(at)) {
                        char b[16];
                        memcpy(b, &st.dsum, 8);
                        memcpy(b + 8, &st.n, 8);
                        c.push_bytes(b, isnull ? 0 : 16);
                    } else {
                        char b[24];
                        memcpy(b, &st.isum, 16);
                        memcpy(b + 16, &st.n, 8);
                        c.push_bytes(b, isnull ? 0 : 24);
                    }
                } else if (isnull) {
                    if (ft == SRB_TYPE_DOUBLE)
                        c.push_double(0);
                    else
                        c.push_int(0);
                } else if (is_decimal(at)) {
                    c.push_int(decimal_div_integer(st.isum, (int128_t)st.n, as)); // avg.h:236-243
                } else if (at == SRB_TYPE_DATE) {
                    c.push_int(st.isum / st.n); // avg.h:231
                } else if (at == SRB_TYPE_LARGEINT) {
                    c.push_double((double)st.isum / (double)st.n);
                } else {
                    c.push_double(st.dsum / (double)st.n); // avg.h:233
                }
                break;
            default:
                if (dbl_arg)
                    c.push_double(isnull ? 0 : st.dmm);
                else
                    c.push_int(isnull ? 0 : st.imm);
            }
        }
    }
    r->finish();
}

// filter-only plan: compacted input columns (Chunk::filter, column/chunk.cpp:302)
static void run_filter_only(const srb_scan_agg_plan* plan, const srb_chunk& in, int chunk_size, Result* r) {
    int no = plan->n_out_slots;
    r->cols.resize(no);
    std::vector<const srb_column*> src(no);
    for (int i = 0; i < no; i++) {
        src[i] = find_slot(in, plan->out_slots[i]);
        r->slot_ids.push_back(plan->out_slots[i]);
        r->cols[i].type = src[i]->type;
        r->cols[i].precision = src[i]->precision;
        r->cols[i].scale = src[i]->scale;
        r->cols[i].has_nulls = src[i]->nulls != nullptr;
    }
    std::vector<uint8_t> sel;
    for (int64_t base = 0; base < in.n_rows; base += chunk_size) {
        size_t rows = (size_t)std::min<int64_t>(chunk_size, in.n_rows - base);
        eval_predicates(plan->preds, plan->n_preds, in, base, rows, sel);
        r->rows_in += (int64_t)rows;
        for (size_t i = 0; i < rows; i++) {
            if (!sel[i]) continue;
            r->rows_selected++;
            int64_t row = base + (int64_t)i;
            for (int c = 0; c < no; c++) {
                OutCol& o = r->cols[c];
                const srb_column& s = *src[c];
                if (s.nulls) o.nulls.push_back(s.nulls[row]);
                if (is_binary(s.type))
                    o.push_bytes((const char*)s.data + s.offsets[row], s.offsets[row + 1] - s.offsets[row]);
                else {
                    int w = type_width(s.type);
                    size_t off = o.data.size();
                    o.data.resize(off + w);
                    memcpy(o.data.data() + off, (const uint8_t*)s.data + (size_t)w * row, w);
                }
            }
        }
    }
    r->n_rows = r->rows_selected;
    r->finish();
}

Result* run_scan_agg(const srb_scan_agg_plan* plan, const srb_chunk* in, int n_threads, int chunk_size) {
    auto* r = new Result();
    if (plan->n_aggs == 0 && plan->n_keys == 0) {
        run_filter_only(plan, *in, chunk_size, r);
        return r;
    }
    if (n_threads < 1) n_threads = 1;
    int64_t n_chunks = (in->n_rows + chunk_size - 1) / chunk_size;
    if (n_chunks < n_threads) n_threads = (int)std::max<int64_t>(1, n_chunks);
    std::vector<std::unique_ptr<Aggregator>> ags;
    for (int t = 0; t < n_threads; t++) ags.emplace_back(new Aggregator(plan));
    std::vector<std::string> errors(n_threads);
    auto work = [&](int t) {
        try {
            int64_t c0 = n_chunks * t / n_threads, c1 = n_chunks * (t + 1) / n_threads;
            for (int64_t c = c0; c < c1; c++) {
                int64_t base = c * chunk_size;
                size_t rows = (size_t)std::min<int64_t>(chunk_size, in->n_rows - base);
                if (ags[t]->merge_input)
                    merge_intermediate_rows(*ags[t], *in, base, rows);
                else
                    ags[t]->push_rows(*in, base, rows);
            }
        } catch (const std::exception& e) {
            errors[t] = e.what();
        }
    };
    if (n_threads == 1)
        work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; t++) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    for (auto& e : errors)
        if (!e.empty()) {
            delete r;
            throw std::runtime_error(e);
        }
    for (int t = 1; t < n_threads; t++) ags[0]->merge_from(*ags[t]);
    ags[0]->emit(r);
    return r;
}

// ---- streaming pre-aggregation sink ---------------------------------------------------------------------------
// AggregateStreamingSinkOperator::push_chunk (pipeline/aggregate/aggregate_streaming_sink_operator.cpp:73-98) dispatches on
// TStreamingPreaggregationMode (gensrc/thrift/PlanNodes.thrift:689-694): FORCE_STREAMING passes every row through in
// intermediate format (:93-103), FORCE_PREAGGREGATION aggregates every row (:105-122), AUTO runs the AggrAutoState machine
// (:147-307) over AggrAutoContext's constants (exec/aggregator.h:172-193). The output is the pass-through rows in arrival order
// with the table's groups interleaved at their creation positions; phase 2 merges both kinds alike, so the final result does not
// depend on the mode or on which rows were passed through.
namespace {
struct AutoContext { // exec/aggregator.h:172-193, aggregator.cpp:104-118
    static constexpr size_t ContinuousUpperLimit = 10000;
    static constexpr int ForcePreaggLimit = 3, PreaggLimit = 100, StableLimit = 5;
    static constexpr double LowReduction = 0.2, HighReduction = 0.9;
    static constexpr size_t MaxHtSize = 64 * 1024 * 1024;
    size_t init_preagg_count = 0, adjust_count = 0, pass_through_count = 0, preagg_count = 0, selective_preagg_count = 0;
    size_t continuous_limit = 100;
    void update_continuous_limit() { continuous_limit = continuous_limit * 2 > ContinuousUpperLimit ? ContinuousUpperLimit : continuous_limit * 2; }
    static bool is_high_reduction(size_t agg_count, size_t chunk_size) { return agg_count >= HighReduction * chunk_size; }
    static bool is_low_reduction(size_t agg_count, size_t chunk_size) { return agg_count <= LowReduction * chunk_size; }
};
enum AutoState { INIT_PREAGG = 0, ADJUST, PASS_THROUGH, FORCE_PREAGG, PREAGG, SELECTIVE_PREAGG };

// Aggregator::should_expand_preagg_hash_tables (aggregator.cpp:770-802) with STREAMING_HT_MIN_REDUCTION (aggregator.h:200-204)
bool should_expand_preagg_hash_tables(size_t prev_row_returned, size_t input_chunk_size, int64_t ht_mem, int64_t ht_rows, int64_t num_rows_returned) {
    static const struct {
        int64_t min_ht_mem;
        double min_reduction;
    } table[] = {{0, 0.0}, {256 * 1024, 1.1}, {2 * 1024 * 1024, 2.0}};
    if (ht_rows == 0) return true;
    int level = 0;
    while (level + 1 < 3 && ht_mem >= table[level + 1].min_ht_mem) level++;
    const int64_t input_rows = (int64_t)prev_row_returned - (int64_t)input_chunk_size;
    const int64_t aggregated_input_rows = input_rows - num_rows_returned;
    const double current_reduction = (double)aggregated_input_rows / (double)ht_rows;
    if (aggregated_input_rows <= 0) return true;
    return current_reduction > table[level].min_reduction;
}
} // namespace

Result* run_scan_agg_streaming(const srb_scan_agg_plan* plan, const srb_chunk* in, int chunk_size, int mode, int64_t* stats) {
    if (plan->n_keys == 0 || plan->phase != SRB_AGG_UPDATE_SERIALIZE)
        throw std::runtime_error("streaming pre-aggregation needs GROUP BY keys and the update+serialize phase");
    auto* r = new Result();
    Aggregator ag(plan);
    AutoContext ac;
    AutoState st = INIT_PREAGG;
    int64_t n_by_state[6] = {0, 0, 0, 0, 0, 0};
    // bytes of the hash table as hash_map_variant().allocated_memory_usage() would see them: slots x (key record + states)
    auto allocated_bytes = [&]() { return (size_t)ag.table.slots.size() * (size_t)(16 + 24 * plan->n_aggs); };
    const int64_t n_chunks = (in->n_rows + chunk_size - 1) / chunk_size;
    for (int64_t c = 0; c < n_chunks; c++) {
        const int64_t base = c * chunk_size;
        const size_t rows = (size_t)std::min<int64_t>(chunk_size, in->n_rows - base);
        if (mode == 1) { // FORCE_STREAMING
            ag.push_rows(*in, base, rows, ROWS_STREAM);
            continue;
        }
        if (mode == 2) { // FORCE_PREAGGREGATION
            ag.push_rows(*in, base, rows, ROWS_AGGREGATE);
            continue;
        }
        // AUTO (and LIMITED_MEM below its limit): _push_chunk_by_auto
        const size_t mem = allocated_bytes();
        const size_t continuous_limit = ac.continuous_limit;
        n_by_state[st]++;
        switch (st) {
        case INIT_PREAGG: {
            const bool needs_expansion = ag.table.need_expand(rows);
            ac.init_preagg_count++;
            if (!needs_expansion || should_expand_preagg_hash_tables((size_t)ag.rows_in + rows, rows, (int64_t)mem, (int64_t)ag.table.n_indexed, ag.rows_streamed)) {
                ag.push_rows(*in, base, rows, ROWS_AGGREGATE);
                break;
            }
            st = ADJUST;
            ac.adjust_count = 0;
        }
            [[fallthrough]];
        case ADJUST: {
            ac.adjust_count++;
            // the reference first looks every key up (build_hash_map_with_selection), then picks how the whole chunk is handled by
            // the hit count; the lookup here is a dry run of the same table
            size_t hit_count = ag.count_hits(*in, base, rows);
            if (ac.adjust_count < continuous_limit && AutoContext::is_low_reduction(hit_count, rows)) {
                ag.push_rows(*in, base, rows, ROWS_STREAM);
                ac.pass_through_count++;
                ac.preagg_count = 0;
                ac.selective_preagg_count = 0;
                if (ac.pass_through_count == (size_t)AutoContext::StableLimit) st = PASS_THROUGH;
            } else if (ac.adjust_count < continuous_limit && AutoContext::is_high_reduction(hit_count, rows) && mem < AutoContext::MaxHtSize) {
                ag.push_rows(*in, base, rows, ROWS_AGGREGATE);
                ac.preagg_count++;
                ac.pass_through_count = 0;
                ac.selective_preagg_count = 0;
                if (ac.preagg_count == (size_t)AutoContext::StableLimit) {
                    st = PREAGG;
                    ac.preagg_count = 0;
                }
            } else {
                ag.push_rows(*in, base, rows, ROWS_SELECTIVE);
                ac.selective_preagg_count++;
                ac.pass_through_count = 0;
                ac.preagg_count = 0;
                if (ac.selective_preagg_count == (size_t)AutoContext::StableLimit) st = SELECTIVE_PREAGG;
            }
            break;
        }
        case PASS_THROUGH: {
            ag.push_rows(*in, base, rows, ROWS_STREAM);
            ac.pass_through_count++;
            if (ac.pass_through_count > continuous_limit) {
                st = mem < AutoContext::MaxHtSize ? FORCE_PREAGG : ADJUST;
                ac.pass_through_count = 0;
                ac.preagg_count = 0;
                ac.adjust_count = 0;
                ac.update_continuous_limit();
            }
            break;
        }
        case FORCE_PREAGG:
        case PREAGG: {
            ag.push_rows(*in, base, rows, ROWS_AGGREGATE);
            ac.preagg_count++;
            const size_t limit = st == FORCE_PREAGG ? AutoContext::ForcePreaggLimit : AutoContext::PreaggLimit;
            if (ac.preagg_count > limit) {
                st = ADJUST;
                ac.preagg_count = 0;
                ac.adjust_count = 0;
            }
            break;
        }
        case SELECTIVE_PREAGG: {
            ag.push_rows(*in, base, rows, ROWS_SELECTIVE);
            ac.selective_preagg_count++;
            if (ac.selective_preagg_count > continuous_limit) {
                st = ADJUST;
                ac.selective_preagg_count = 0;
                ac.adjust_count = 0;
                ac.update_continuous_limit();
            }
            break;
        }
        }
    }
    ag.emit(r);
    if (stats) {
        stats[0] = ag.rows_streamed;
        stats[1] = (int64_t)ag.table.n_indexed;
        for (int i = 0; i < 6; i++) stats[2 + i] = n_by_state[i];
        stats[8] = (int64_t)st;
    }
    return r;
}

} // namespace orc
