// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_arith.h header).
// Hash aggregation restating:
//   Aggregator::compute_single_agg_state / compute_batch_agg_states   be/src/exec/aggregator.cpp:833-880
//   Aggregator::build_hash_map / convert_hash_map_to_chunk            be/src/exec/aggregator.cpp:1357-1514
//   AggHashMapWithOneNumberKeyWithNullable::compute_agg_states        be/src/exec/aggregate/agg_hash_map.h:228-339
//   SUM / AVG / COUNT / MIN / MAX states                             be/src/exprs/agg/{sum.h:41-117,avg.h:64-245,count.h:36-120,maxmin.h:30-122}
//   NullableAggregateFunctionUnary (skip NULL rows, NULL if none)     be/src/exprs/agg/nullable_aggregate.h:340-400
#pragma once
#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include "orc_engine.h"

namespace orc {

// One accumulator per aggregate function per group. `n` counts non-null inputs (the nullable wrapper's
// "is_null" flag is n == 0); COUNT uses n as its value.
struct AggState {
    int128_t isum = 0; // SUM/AVG over integer-like values: wrapping int128 (sum.h:61 `sum += x`)
    double dsum = 0;   // SUM/AVG over DOUBLE, and AVG over plain integers (ImmediateAvgResultLT = DOUBLE, avg.h:49-53)
    int64_t n = 0;
    int128_t imm = 0; // MIN/MAX
    double dmm = 0;
};

// result types of the finalize phase
inline void agg_result_type(int fn, int arg_type, int arg_scale, int* type, int* precision, int* scale) {
    *precision = 0;
    *scale = 0;
    switch (fn) {
    case SRB_AGG_COUNT: *type = SRB_TYPE_BIGINT; return;
    case SRB_AGG_SUM:
        if (is_decimal(arg_type)) { // DecimalSumAggregateFunction: any decimal -> DECIMAL128 (sum.h:176-181)
            *type = SRB_TYPE_DECIMAL128;
            *precision = 38;
            *scale = arg_scale;
        } else if (arg_type == SRB_TYPE_DOUBLE)
            *type = SRB_TYPE_DOUBLE;
        else if (arg_type == SRB_TYPE_LARGEINT)
            *type = SRB_TYPE_LARGEINT;
        else
            *type = SRB_TYPE_BIGINT; // SumBigIntLTGuard (sum.h:30-31)
        return;
    case SRB_AGG_AVG:
        if (is_decimal(arg_type)) { // avg.h:36-39, scale from decimal_div_integer (arithmetic_operation.h:409-422)
            *type = SRB_TYPE_DECIMAL128;
            *precision = 38;
            *scale = div_result_scale(arg_scale);
        } else if (arg_type == SRB_TYPE_DATE)
            *type = SRB_TYPE_DATE;
        else
            *type = SRB_TYPE_DOUBLE;
        return;
    default: // MIN / MAX keep the argument type (maxmin.h)
        *type = arg_type;
        *scale = arg_scale;
        return;
    }
}

struct OutCol {
    int type = 0, precision = 0, scale = 0;
    std::vector<uint8_t> data; // fixed-width values, or bytes for binary
    std::vector<uint8_t> nulls;
    std::vector<uint32_t> offsets; // binary
    bool has_nulls = false;
    void push_int(int128_t x) {
        int w = type_width(type);
        size_t o = data.size();
        data.resize(o + w);
        memcpy(data.data() + o, &x, w); // little-endian truncation
    }
    void push_double(double x) {
        size_t o = data.size();
        data.resize(o + 8);
        memcpy(data.data() + o, &x, 8);
    }
    void push_bytes(const void* p, size_t n) {
        if (offsets.empty()) offsets.push_back(0);
        size_t o = data.size();
        data.resize(o + n);
        memcpy(data.data() + o, p, n);
        offsets.push_back((uint32_t)data.size());
    }
};

struct Result {
    int64_t n_rows = 0;
    std::vector<int32_t> slot_ids;
    std::vector<OutCol> cols;
    int64_t rows_in = 0, rows_selected = 0;
    // C view
    std::vector<srb_column> cview;
    srb_chunk chunk{};
    void finish() {
        cview.resize(cols.size());
        for (size_t i = 0; i < cols.size(); i++) {
            OutCol& c = cols[i];
            if (is_binary(c.type) && c.offsets.empty()) c.offsets.push_back(0);
            srb_column& o = cview[i];
            o.type = c.type;
            o.precision = c.precision;
            o.scale = c.scale;
            o.mem_kind = SRB_MEM_HOST;
            o.data = c.data.data();
            o.nulls = c.has_nulls ? c.nulls.data() : nullptr;
            o.offsets = is_binary(c.type) ? c.offsets.data() : nullptr;
            o.n_bytes = (int64_t)c.data.size();
        }
        chunk.n_rows = n_rows;
        chunk.n_cols = (int32_t)cols.size();
        chunk.slot_ids = slot_ids.data();
        chunk.cols = cview.data();
    }
};

// group key = serialized bytes, per key column: [null byte if the column is nullable][value bytes | u32 len + bytes]
// (AggHashMapWithSerializedKey, agg_hash_map.h:746-748; BinaryColumn::serialize, binary_column.h:228-237).
struct KeyColMeta {
    int type, precision, scale;
    bool nullable;
};

struct GroupTable {
    // open addressing over the reference's hash of the serialized key; internals are not observable
    std::vector<int64_t> slots; // group index + 1
    std::vector<uint64_t> hashes;
    std::vector<std::string> keys;
    std::vector<uint8_t> streamed; // 1: a pass-through row of the streaming sink (never findable), sized lazily
    size_t n_indexed = 0;
    size_t mask = 0;
    int seed;
    explicit GroupTable(int seed_) : seed(seed_) { rehash(1024); }
    bool is_streamed(size_t g) const { return g < streamed.size() && streamed[g]; }
    void rehash(size_t cap) {
        slots.assign(cap, 0);
        mask = cap - 1;
        for (size_t g = 0; g < keys.size(); g++) {
            if (is_streamed(g)) continue;
            size_t p = hashes[g] & mask;
            while (slots[p]) p = (p + 1) & mask;
            slots[p] = (int64_t)g + 1;
        }
    }
    uint64_t hash_key(const std::string& k) const {
        // single <=8-byte key: StdHashWithSeed (util/hash.h:211-216); 16-byte: Hash128WithSeed (column_hash.h:143-147);
        // otherwise the slice hash crc_hash_64 (util/hash.h:183-204)
        if (k.size() == 4) {
            int32_t x;
            memcpy(&x, k.data(), 4);
            return std_hash_with_seed(x, seed);
        }
        if (k.size() == 8) {
            int64_t x;
            memcpy(&x, k.data(), 8);
            return std_hash_with_seed(x, seed);
        }
        if (k.size() == 16) {
            int128_t x;
            memcpy(&x, k.data(), 16);
            return hash128_with_seed(x, seed);
        }
        return crc_hash_64(k.data(), (int32_t)k.size(), seed == 1 ? CRC_HASH_SEED1 : CRC_HASH_SEED2);
    }
    // returns group index, creating it when absent (lazy_emplace + AllocateState, aggregator.h:620-641)
    size_t emplace(const std::string& k, bool* created) {
        uint64_t h = hash_key(k);
        size_t p = h & mask;
        while (slots[p]) {
            size_t g = (size_t)slots[p] - 1;
            if (hashes[g] == h && keys[g] == k) {
                *created = false;
                return g;
            }
            p = (p + 1) & mask;
        }
        size_t g = keys.size();
        keys.push_back(k);
        hashes.push_back(h);
        slots[p] = (int64_t)g + 1;
        *created = true;
        n_indexed++;
        if (n_indexed * 2 > slots.size()) rehash(slots.size() * 2);
        return g;
    }
    // streaming sink: lookup without insertion (build_hash_map_with_selection, agg_hash_map.h:298-339: not-found rows are selected
    // for pass-through); returns false when absent
    bool find(const std::string& k, size_t* g_out) const {
        uint64_t h = hash_key(k);
        size_t p = h & mask;
        while (slots[p]) {
            size_t g = (size_t)slots[p] - 1;
            if (hashes[g] == h && keys[g] == k) {
                *g_out = g;
                return true;
            }
            p = (p + 1) & mask;
        }
        return false;
    }
    // a pass-through row: gets a state block of its own and an output position, never enters the hash slots
    size_t append_streamed(const std::string& k) {
        size_t g = keys.size();
        keys.push_back(k);
        hashes.push_back(0);
        streamed.resize(g + 1, 0);
        streamed[g] = 1;
        return g;
    }
    // hash_map_variant().need_expand(n): inserting n new keys would grow the table
    bool need_expand(size_t n) const { return (n_indexed + n) * 2 > slots.size(); }
};

// how a chunk's rows meet the table (AggregateStreamingSinkOperator::_push_chunk_by_*): every row aggregated with insertion
// (force_preaggregation), every row passed through in intermediate format (force_streaming), or found rows aggregated and the
// others passed through (selective_preaggregation)
enum RowMode { ROWS_AGGREGATE = 0, ROWS_STREAM = 1, ROWS_SELECTIVE = 2 };

struct Aggregator {
    const srb_scan_agg_plan* plan;
    std::vector<KeyColMeta> key_meta;
    bool key_meta_known = false;
    GroupTable table;
    std::vector<AggState> states; // group-major: states[g * n_aggs + a]
    std::vector<int> arg_type, arg_scale, arg_prec;
    bool arg_meta_known = false;
    int64_t rows_in = 0, rows_selected = 0;
    int n_aggs;
    bool merge_input;

    explicit Aggregator(const srb_scan_agg_plan* p)
            : plan(p), table(p->phase >= SRB_AGG_MERGE_FINALIZE ? 2 : 1), n_aggs(p->n_aggs),
              merge_input(p->phase >= SRB_AGG_MERGE_FINALIZE) {
        arg_type.assign(n_aggs, 0);
        arg_scale.assign(n_aggs, 0);
        arg_prec.assign(n_aggs, 0);
        if (plan->n_keys == 0) states.resize(n_aggs);
    }

    static void update_one(int fn, AggState& st, const Col* c, size_t i, bool avg_as_double) {
        if (fn == SRB_AGG_COUNT) {
            if (!c || !c->is_null(i)) st.n++;
            return;
        }
        if (c->is_null(i)) return;
        bool dbl = c->type == SRB_TYPE_DOUBLE;
        switch (fn) {
        case SRB_AGG_SUM:
            if (dbl)
                st.dsum += c->d[i];
            else
                st.isum = (int128_t)((uint128_t)st.isum + (uint128_t)c->v[i]);
            st.n++;
            break;
        case SRB_AGG_AVG:
            if (dbl)
                st.dsum += c->d[i];
            else if (avg_as_double)
                st.dsum += (double)c->v[i];
            else
                st.isum = (int128_t)((uint128_t)st.isum + (uint128_t)c->v[i]);
            st.n++;
            break;
        case SRB_AGG_MIN:
            if (dbl) {
                if (st.n == 0 || c->d[i] < st.dmm) st.dmm = c->d[i];
            } else if (st.n == 0 || c->v[i] < st.imm)
                st.imm = c->v[i];
            st.n++;
            break;
        case SRB_AGG_MAX:
            if (dbl) {
                if (st.n == 0 || c->d[i] > st.dmm) st.dmm = c->d[i];
            } else if (st.n == 0 || c->v[i] > st.imm)
                st.imm = c->v[i];
            st.n++;
            break;
        }
    }

    static bool avg_double(int t) { return !(is_decimal(t) || t == SRB_TYPE_DATE || t == SRB_TYPE_LARGEINT); }

    // one <=chunk_size slice of raw rows: predicates -> selection -> exprs -> table -> states
    void serialize_key(const std::vector<const Col*>& keys, size_t i, std::string& kb) const {
        kb.clear();
        for (int k = 0; k < plan->n_keys; k++) {
            const Col& c = *keys[k];
            bool isnull = c.is_null(i);
            if (key_meta[k].nullable) kb.push_back((char)isnull);
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
    }

    // build_hash_map_with_selection (aggregator.h, agg_hash_map.h:298-339) as a dry run: how many selected rows of the chunk
    // already have their group in the table. Nothing is inserted or updated.
    size_t count_hits(const srb_chunk& ch, int64_t base, size_t rows) {
        std::vector<uint8_t> sel;
        eval_predicates(plan->preds, plan->n_preds, ch, base, rows, sel);
        std::vector<uint32_t> idx;
        for (size_t i = 0; i < rows; i++)
            if (sel[i]) idx.push_back((uint32_t)i);
        if (idx.empty() || !key_meta_known) return 0;
        ExprEval ev(plan->nodes, plan->n_nodes, &ch, base, &idx);
        std::vector<const Col*> keys(plan->n_keys);
        for (int k = 0; k < plan->n_keys; k++) keys[k] = &ev.eval(plan->key_nodes[k]);
        std::string kb;
        size_t hits = 0, g;
        for (size_t i = 0; i < idx.size(); i++) {
            serialize_key(keys, i, kb);
            hits += table.find(kb, &g) ? 1 : 0;
        }
        return hits;
    }

    // build_hash_map_with_selection / build_hash_map_with_selection_and_allocation (agg_hash_map.h:298-339, aggregator.h:620-641):
    // not_found[i] = 1 when row i's key is absent; with `allocate` an absent key is inserted as it is met, so that a later row of
    // the same chunk with that key is found
    void lookup_selection(const srb_chunk& ch, bool allocate, std::vector<uint8_t>& not_found) {
        std::vector<uint32_t> idx((size_t)ch.n_rows);
        for (size_t i = 0; i < idx.size(); i++) idx[i] = (uint32_t)i;
        not_found.assign(idx.size(), 0);
        if (idx.empty()) return;
        ExprEval ev(plan->nodes, plan->n_nodes, &ch, 0, &idx);
        std::vector<const Col*> keys(plan->n_keys);
        for (int k = 0; k < plan->n_keys; k++) keys[k] = &ev.eval(plan->key_nodes[k]);
        if (!key_meta_known) {
            for (int k = 0; k < plan->n_keys; k++) key_meta.push_back({keys[k]->type, keys[k]->precision, keys[k]->scale, keys[k]->nullable});
            key_meta_known = true;
        }
        std::string kb;
        size_t g;
        for (size_t i = 0; i < idx.size(); i++) {
            serialize_key(keys, i, kb);
            if (table.find(kb, &g)) continue;
            not_found[i] = 1;
            if (allocate) {
                bool created;
                table.emplace(kb, &created);
                if (created) states.resize(states.size() + n_aggs);
            }
        }
    }

    int64_t rows_streamed = 0;
    // returns the number of rows whose group already existed (the streaming sink's hit count)
    size_t push_rows(const srb_chunk& ch, int64_t base, size_t rows, RowMode mode = ROWS_AGGREGATE) {
        size_t hits = 0;
        rows_in += (int64_t)rows;
        std::vector<uint8_t> sel;
        eval_predicates(plan->preds, plan->n_preds, ch, base, rows, sel);
        std::vector<uint32_t> idx;
        idx.reserve(rows);
        for (size_t i = 0; i < rows; i++)
            if (sel[i]) idx.push_back((uint32_t)i);
        rows_selected += (int64_t)idx.size();
        if (idx.empty()) return 0; // empty chunks are dropped by the driver (pipeline_driver.cpp:340-342)
        ExprEval ev(plan->nodes, plan->n_nodes, &ch, base, &idx);
        size_t n = idx.size();
        std::vector<const Col*> keys(plan->n_keys), args(n_aggs, nullptr);
        for (int k = 0; k < plan->n_keys; k++) keys[k] = &ev.eval(plan->key_nodes[k]);
        for (int a = 0; a < n_aggs; a++)
            if (plan->aggs[a].arg >= 0) args[a] = &ev.eval(plan->aggs[a].arg);
        if (!arg_meta_known) {
            for (int a = 0; a < n_aggs; a++)
                if (args[a]) {
                    arg_type[a] = args[a]->type;
                    arg_scale[a] = args[a]->scale;
                    arg_prec[a] = args[a]->precision;
                }
            arg_meta_known = true;
        }
        if (!key_meta_known) {
            for (int k = 0; k < plan->n_keys; k++)
                key_meta.push_back({keys[k]->type, keys[k]->precision, keys[k]->scale, keys[k]->nullable});
            key_meta_known = true;
        }
        std::vector<size_t> gid(n, 0);
        if (plan->n_keys > 0) {
            std::string kb;
            for (size_t i = 0; i < n; i++) {
                serialize_key(keys, i, kb);
                bool created = false;
                if (mode == ROWS_AGGREGATE) {
                    gid[i] = table.emplace(kb, &created);
                    hits += created ? 0 : 1;
                } else if (mode == ROWS_SELECTIVE && table.find(kb, &gid[i])) {
                    hits++;
                } else { // convert_to_serialize_format (aggregate.h:337): the row's own one-row state, emitted as it stands
                    gid[i] = table.append_streamed(kb);
                    created = true;
                    rows_streamed++;
                }
                if (created) states.resize(states.size() + n_aggs);
            }
        }
        // per function, per row indirect update (AggregateFunctionBatchHelper::update_batch, aggregate.h:403-408)
        for (int a = 0; a < n_aggs; a++) {
            int fn = plan->aggs[a].fn;
            bool ad = args[a] ? avg_double(args[a]->type) : false;
            for (size_t i = 0; i < n; i++) update_one(fn, states[gid[i] * n_aggs + a], args[a], i, ad);
        }
        return hits;
    }

    // merge another aggregator's groups (phase-2 merge_batch semantics: sums add, counts add, min/max combine;
    // aggregator.cpp:871-875, sum.h:113-117, avg.h:161-166, count.h:92-96)
    void merge_from(const Aggregator& o) {
        rows_in += o.rows_in;
        rows_selected += o.rows_selected;
        if (!arg_meta_known && o.arg_meta_known) {
            arg_type = o.arg_type;
            arg_scale = o.arg_scale;
            arg_prec = o.arg_prec;
            arg_meta_known = true;
        }
        if (!key_meta_known && o.key_meta_known) {
            key_meta = o.key_meta;
            key_meta_known = true;
        }
        size_t ng = plan->n_keys ? o.table.keys.size() : 1;
        for (size_t g = 0; g < ng; g++) {
            size_t mine = 0;
            if (plan->n_keys) {
                bool created;
                mine = table.emplace(o.table.keys[g], &created);
                if (created) states.resize(states.size() + n_aggs);
            }
            for (int a = 0; a < n_aggs; a++) {
                const AggState& s = o.states[g * n_aggs + a];
                AggState& d = states[mine * n_aggs + a];
                int fn = plan->aggs[a].fn;
                if (fn == SRB_AGG_MIN || fn == SRB_AGG_MAX) {
                    if (s.n) {
                        bool less = fn == SRB_AGG_MIN;
                        if (d.n == 0 || (less ? s.imm < d.imm : s.imm > d.imm)) d.imm = s.imm;
                        if (d.n == 0 || (less ? s.dmm < d.dmm : s.dmm > d.dmm)) d.dmm = s.dmm;
                    }
                } else {
                    d.isum = (int128_t)((uint128_t)d.isum + (uint128_t)s.isum);
                    d.dsum += s.dsum;
                }
                d.n += s.n;
            }
        }
    }

    void emit(Result* r) const;
};

} // namespace orc
