// srb_stream_policy.h — the AUTO mode of the streaming pre-aggregation sink as host-only code (no CUDA): the reference's
// AggrAutoState machine (pipeline/aggregate/aggregate_streaming_sink_operator.cpp:147-307) over AggrAutoContext's constants
// (exec/aggregator.h:172-193) and Aggregator::should_expand_preagg_hash_tables (exec/aggregator.cpp:770-802), restated for a device
// table. One *unit* of rows stands for one chunk. next() says how the unit's rows meet the table; observe() feeds the unit's hit
// count back. Where the reference looks every key of the chunk up first (ADJUST: build_hash_map_with_selection) and then treats the
// whole chunk one of three ways, the device runs the selective pass directly — found rows update their group, the others pass
// through — and the unit is classified by its hit count afterwards: the decision costs no extra pass over the rows.
// Kept free of CUDA so that tests/cpp/test_stream_policy.cpp can replay it on the CPU.
#pragma once
#include <cstddef>
#include <cstdint>

namespace srb {

enum StreamRowMode { STREAM_ROWS_AGGREGATE = 0, STREAM_ROWS_PASS = 1, STREAM_ROWS_SELECTIVE = 2 };
enum StreamAutoState { STREAM_INIT_PREAGG = 0, STREAM_ADJUST, STREAM_PASS_THROUGH, STREAM_FORCE_PREAGG, STREAM_PREAGG, STREAM_SELECTIVE_PREAGG };

struct StreamPolicy {
    // exec/aggregator.h:172-193
    static constexpr size_t ContinuousUpperLimit = 10000;
    static constexpr int ForcePreaggLimit = 3, PreaggLimit = 100, StableLimit = 5;
    static constexpr double LowReduction = 0.2, HighReduction = 0.9;
    // The reference's thresholds stand for CPU cache sizes (STREAMING_HT_MIN_REDUCTION steps at 256 KB and 2 MB, MaxHtSize 64 MB).
    // On a B200 a table that fits the 126 MB L2 resolves its atomics there: below l2_bytes it may always grow, up to ten times that
    // it must reduce by more than 1.1, beyond by more than 2; max_ht_bytes bounds the table FORCE_PREAGG / PREAGG may build.
    size_t l2_bytes = (size_t)96 << 20;
    size_t max_ht_bytes = (size_t)8 << 30;

    StreamAutoState state = STREAM_INIT_PREAGG;
    size_t adjust_count = 0, pass_through_count = 0, preagg_count = 0, selective_preagg_count = 0, continuous_limit = 100;

    void update_continuous_limit() { continuous_limit = continuous_limit * 2 > ContinuousUpperLimit ? ContinuousUpperLimit : continuous_limit * 2; }

    // Aggregator::should_expand_preagg_hash_tables: keep growing only while the rows aggregated so far reduce by more than the
    // factor the table's size asks for
    bool should_expand(uint64_t groups, size_t table_bytes, int64_t rows_before, int64_t rows_passed_through) const {
        if (groups == 0) return true;
        const double min_reduction = table_bytes < l2_bytes ? 0.0 : (table_bytes < l2_bytes * 10 ? 1.1 : 2.0);
        const int64_t aggregated = rows_before - rows_passed_through;
        if (aggregated <= 0) return true;
        return (double)aggregated / (double)groups > min_reduction;
    }

    // how the next unit of n rows meets the table; groups / capacity / table_bytes describe the table now, rows_before the input rows
    // handed to the sink before this unit, rows_passed_through those of them that were not aggregated
    StreamRowMode next(int64_t n, uint64_t groups, uint64_t capacity, size_t table_bytes, int64_t rows_before, int64_t rows_passed_through) {
        if (state == STREAM_INIT_PREAGG) {
            const bool needs_expansion = (groups + (uint64_t)n) * 2 > capacity;
            if (!needs_expansion || should_expand(groups, table_bytes, rows_before, rows_passed_through)) return STREAM_ROWS_AGGREGATE;
            state = STREAM_ADJUST;
            adjust_count = 0;
        }
        switch (state) {
        case STREAM_ADJUST:
        case STREAM_SELECTIVE_PREAGG: return STREAM_ROWS_SELECTIVE;
        case STREAM_PASS_THROUGH: return STREAM_ROWS_PASS;
        default: return STREAM_ROWS_AGGREGATE; // FORCE_PREAGG, PREAGG
        }
    }

    // after the unit ran: `hits` of its `rows` selected rows found their group already in the table
    void observe(int64_t hits, int64_t rows, size_t table_bytes) {
        switch (state) {
        case STREAM_INIT_PREAGG: break;
        case STREAM_ADJUST: {
            adjust_count++;
            if (adjust_count < continuous_limit && (double)hits <= LowReduction * (double)rows) {
                pass_through_count++;
                preagg_count = selective_preagg_count = 0;
                if (pass_through_count == (size_t)StableLimit) state = STREAM_PASS_THROUGH;
            } else if (adjust_count < continuous_limit && (double)hits >= HighReduction * (double)rows && table_bytes < max_ht_bytes) {
                preagg_count++;
                pass_through_count = selective_preagg_count = 0;
                if (preagg_count == (size_t)StableLimit) {
                    state = STREAM_PREAGG;
                    preagg_count = 0;
                }
            } else {
                selective_preagg_count++;
                pass_through_count = preagg_count = 0;
                if (selective_preagg_count == (size_t)StableLimit) state = STREAM_SELECTIVE_PREAGG;
            }
            break;
        }
        case STREAM_PASS_THROUGH:
            pass_through_count++;
            if (pass_through_count > continuous_limit) {
                state = table_bytes < max_ht_bytes ? STREAM_FORCE_PREAGG : STREAM_ADJUST;
                pass_through_count = preagg_count = adjust_count = 0;
                update_continuous_limit();
            }
            break;
        case STREAM_FORCE_PREAGG:
        case STREAM_PREAGG: {
            preagg_count++;
            const size_t limit = state == STREAM_FORCE_PREAGG ? ForcePreaggLimit : PreaggLimit;
            if (preagg_count > limit) {
                state = STREAM_ADJUST;
                preagg_count = adjust_count = 0;
            }
            break;
        }
        case STREAM_SELECTIVE_PREAGG:
            selective_preagg_count++;
            if (selective_preagg_count > continuous_limit) {
                state = STREAM_ADJUST;
                selective_preagg_count = adjust_count = 0;
                update_continuous_limit();
            }
            break;
        }
    }
};

} // namespace srb
