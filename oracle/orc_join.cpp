// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_arith.h header).
// Chained hash join restating:
//   JoinHashTable::append_chunk (row 0 is a dummy; indices 1-based)   be/src/exec/join_hash_map.cpp:613-642
//   JoinBuildFunc::prepare / construct_hash_table                     be/src/exec/join_hash_map.tpp:29-91
//   JoinProbeFunc::lookup_init (first[bucket] gather)                 be/src/exec/join_hash_map.tpp:283-309
//   JoinHashMap::_do_probe_from_ht (inner-join chain walk)            be/src/exec/join_hash_map.tpp:1066-1143
//   serialize_fixed_size_key_column (multi-column fixed keys)         be/src/exec/join_hash_map.h:391-406
//   _probe_output / _build_output (append_selective by index)         be/src/exec/join_hash_map.tpp:534-550,624-640
//   outer / right-side joins: _probe_from_ht_for_left_outer_join, _for_right_outer_join, _for_full_outer_join, _for_right_semi_join,
//   _for_right_anti_join + _search_ht_remain (unmatched build rows, null-key rows included)   be/src/exec/join_hash_map.tpp:1187-1986;
//   a missing side's columns are NULL (append_default on nullable output columns, tpp:656-752)
#include <stdexcept>
#include <thread>

#include "orc_agg.h"

namespace orc {

struct PackedKeys {
    int width = 0;          // 4, 8 or 16 bytes
    std::vector<int128_t> k; // zero-extended packed key
    std::vector<uint8_t> null;
};

// packs the key columns of `ch` (rows 0..n) column by column at increasing byte offsets (join_hash_map.h:391-406)
// The reference runs the probe side with DOP pipeline drivers over ONE shared, read-only table (HashJoinProbeOperatorFactory clones a
// readable view per driver, be/src/exec/pipeline/hashjoin/hash_join_probe_operator.cpp:104); the build runs in one driver
// (hash_join_build_operator.cpp:85 -> JoinHashTable::build). n_threads stands for that DOP: contiguous probe row ranges, outputs
// concatenated in range order, so the result is the same rows in the same order as with one thread.
template <class F>
static void parallel_ranges(int64_t n, int n_threads, F&& f) {
    int T = (int)std::max<int64_t>(1, std::min<int64_t>(n_threads, n / 4096));
    if (T == 1) {
        f(0, 0, n);
        return;
    }
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) th.emplace_back([&, t] { f(t, n * t / T, n * (t + 1) / T); });
    for (auto& x : th) x.join();
}

static PackedKeys pack_keys(const srb_chunk& ch, const int32_t* slots, int n_keys, int n_threads = 1) {
    PackedKeys pk;
    int total = 0;
    std::vector<const srb_column*> cols(n_keys);
    for (int i = 0; i < n_keys; i++) {
        cols[i] = find_slot(ch, slots[i]);
        int w = type_width(cols[i]->type);
        if (w == 0) throw std::runtime_error("join: binary keys not supported");
        total += w;
    }
    pk.width = total <= 4 ? 4 : (total <= 8 ? 8 : 16);
    if (total > 16) throw std::runtime_error("join: packed key wider than 16 bytes");
    pk.k.assign(ch.n_rows, 0);
    pk.null.assign(ch.n_rows, 0);
    parallel_ranges(ch.n_rows, n_threads, [&](int, int64_t r0, int64_t r1) {
        for (int64_t r = r0; r < r1; r++) {
            uint8_t buf[16] = {0};
            int off = 0;
            for (int i = 0; i < n_keys; i++) {
                int w = type_width(cols[i]->type);
                memcpy(buf + off, (const uint8_t*)cols[i]->data + (size_t)w * r, w);
                off += w;
                if (cols[i]->nulls && cols[i]->nulls[r]) pk.null[r] = 1;
            }
            memcpy(&pk.k[r], buf, 16);
        }
    });
    return pk;
}

static inline uint32_t bucket_of(const PackedKeys& pk, int128_t key, uint32_t bucket_size, uint32_t log_bucket) {
    if (pk.width == 4) return join_key_hash32((uint32_t)key, log_bucket);
    if (pk.width == 8) return join_key_hash64((uint64_t)key, log_bucket);
    return join_key_hash_bytes(&key, 16, bucket_size); // JoinKeyHash<int128>: crc_hash_32 & (B-1) (join_hash_map.h:310-318)
}

struct JoinTable {
    PackedKeys keys; // index i-1 holds build row i
    std::vector<uint32_t> first, next;
    uint32_t bucket_size = 0, log_bucket = 0, row_count = 0;
};

static void build_table(JoinTable& t) {
    t.row_count = (uint32_t)t.keys.k.size();
    t.bucket_size = calc_bucket_size(t.row_count + 1);
    t.log_bucket = (uint32_t)__builtin_ctz(t.bucket_size);
    t.first.assign(t.bucket_size, 0);
    t.next.assign((size_t)t.row_count + 1, 0);
    for (uint32_t i = 1; i < t.row_count + 1; i++) {
        if (t.keys.null[i - 1]) continue; // null keys are never inserted (tpp:62-72)
        uint32_t b = bucket_of(t.keys, t.keys.k[i - 1], t.bucket_size, t.log_bucket);
        t.next[i] = t.first[b];
        t.first[b] = i;
    }
}

constexpr uint32_t NO_ROW = 0xFFFFFFFFu; // an output row whose side is missing: the column is NULL there

static void copy_rows(const srb_column& s, const std::vector<uint32_t>& rows, OutCol& o, int n_threads = 1, bool side_may_miss = false) {
    o.type = s.type;
    o.precision = s.precision;
    o.scale = s.scale;
    o.has_nulls = s.nulls != nullptr || side_may_miss;
    if (side_may_miss) { // (outer joins: row by row, fixed-width and binary alike)
        for (uint32_t row : rows) {
            const bool miss = row == NO_ROW;
            o.nulls.push_back(miss ? 1 : (s.nulls ? s.nulls[row] : 0));
            if (is_binary(s.type)) {
                if (miss) o.push_bytes("", 0);
                else o.push_bytes((const char*)s.data + s.offsets[row], s.offsets[row + 1] - s.offsets[row]);
            } else {
                const int w = type_width(s.type);
                const size_t off = o.data.size();
                o.data.resize(off + w, 0);
                if (!miss) memcpy(o.data.data() + off, (const uint8_t*)s.data + (size_t)w * row, w);
            }
        }
        if (is_binary(s.type) && o.offsets.empty()) o.offsets.push_back(0);
        return;
    }
    if (!is_binary(s.type) && n_threads > 1) {
        const int w = type_width(s.type);
        o.data.resize(rows.size() * (size_t)w);
        if (s.nulls) o.nulls.resize(rows.size());
        parallel_ranges((int64_t)rows.size(), n_threads, [&](int, int64_t a, int64_t b) {
            for (int64_t i = a; i < b; i++) {
                memcpy(o.data.data() + (size_t)i * w, (const uint8_t*)s.data + (size_t)w * rows[i], w);
                if (s.nulls) o.nulls[i] = s.nulls[rows[i]];
            }
        });
        return;
    }
    for (uint32_t row : rows) {
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
    if (is_binary(s.type) && o.offsets.empty()) o.offsets.push_back(0);
}

static void probe_range(const srb_join_plan* plan, const JoinTable& t, const PackedKeys& pk, int64_t i0, int64_t i1, std::vector<uint32_t>& probe_index,
                        std::vector<uint32_t>& build_index, std::vector<uint8_t>* matched) {
    const bool null_aware = plan->join_type == SRB_JOIN_NULL_AWARE_LEFT_ANTI;
    const bool anti = plan->join_type == SRB_JOIN_LEFT_ANTI || null_aware;
    const bool keep_left = plan->join_type == SRB_JOIN_LEFT_OUTER || plan->join_type == SRB_JOIN_FULL_OUTER; // unmatched probe rows stay, build index 0
    const bool right_only = plan->join_type == SRB_JOIN_RIGHT_SEMI || plan->join_type == SRB_JOIN_RIGHT_ANTI; // pairs are not emitted
    for (int64_t i = i0; i < i1; i++) {
        if (keep_left || matched) {
            bool any = false;
            if (!pk.null[i]) {
                for (uint32_t bi = t.first[bucket_of(pk, pk.k[i], t.bucket_size, t.log_bucket)]; bi != 0; bi = t.next[bi]) {
                    if (t.keys.k[bi - 1] != pk.k[i]) continue;
                    any = true;
                    if (matched) (*matched)[bi] = 1;
                    if (!right_only) {
                        probe_index.push_back((uint32_t)i);
                        build_index.push_back(bi);
                    }
                }
            }
            if (!any && keep_left) {
                probe_index.push_back((uint32_t)i);
                build_index.push_back(0);
            }
            continue;
        }
        if (pk.null[i]) {
            // lookup_init leaves next[i] = 0 for NULL probe keys: no match. LEFT ANTI emits such rows (join_hash_map.tpp:1380-1420,
            // "NULLs are not matches and therefore always returned", PlanNodes.thrift:710-713); every other type drops them; the NULL-aware
            // variant (NOT IN) skips them unless the build side is empty (join_hash_map.tpp:1356-1362, join_hash_map.cpp:676-690)
            if (anti && !(null_aware && t.row_count > 0)) {
                probe_index.push_back((uint32_t)i);
                build_index.push_back(1);
            }
            continue;
        }
        uint32_t bi = t.first[bucket_of(pk, pk.k[i], t.bucket_size, t.log_bucket)];
        bool matched = false;
        if (anti) {
            for (; bi != 0 && !matched; bi = t.next[bi]) matched = t.keys.k[bi - 1] == pk.k[i];
            if (!matched) {
                probe_index.push_back((uint32_t)i);
                build_index.push_back(1);
            }
            continue;
        }
        while (bi != 0) {
            if (t.keys.k[bi - 1] == pk.k[i]) {
                if (plan->join_type == SRB_JOIN_LEFT_SEMI) {
                    if (!matched) {
                        probe_index.push_back((uint32_t)i);
                        build_index.push_back(bi);
                    }
                    matched = true;
                    break;
                }
                probe_index.push_back((uint32_t)i);
                build_index.push_back(bi);
            }
            bi = t.next[bi];
        }
    }
}

Result* run_join(const srb_join_plan* plan, const srb_chunk* build, const srb_chunk* probe, int n_threads) {
    auto* r = new Result();
    JoinTable t;
    t.keys = pack_keys(*build, plan->build_key_slots, plan->n_keys, n_threads);
    build_table(t);
    PackedKeys pk = pack_keys(*probe, plan->probe_key_slots, plan->n_keys, n_threads);
    if (pk.width != t.keys.width) throw std::runtime_error("join: key width mismatch");
    std::vector<uint32_t> probe_index, build_index; // build_index is 1-based like the reference's
    const bool anti = plan->join_type == SRB_JOIN_LEFT_ANTI || plan->join_type == SRB_JOIN_NULL_AWARE_LEFT_ANTI;
    bool build_null = false; // NULL-aware LEFT ANTI: a NULL key on a non-empty build side ends the join (hash_joiner.h:386-396)
    if (plan->join_type == SRB_JOIN_NULL_AWARE_LEFT_ANTI)
        for (uint8_t x : t.keys.null) build_null = build_null || x != 0;
    const int jt = plan->join_type;
    const bool track_right = jt == SRB_JOIN_RIGHT_OUTER || jt == SRB_JOIN_FULL_OUTER || jt == SRB_JOIN_RIGHT_SEMI || jt == SRB_JOIN_RIGHT_ANTI;
    const bool left_may_miss = jt == SRB_JOIN_RIGHT_OUTER || jt == SRB_JOIN_FULL_OUTER;
    const bool right_may_miss = jt == SRB_JOIN_LEFT_OUTER || jt == SRB_JOIN_FULL_OUTER;
    std::vector<uint8_t> matched; // per build row (1-based), HashTableProbeState::build_match_index
    {
        const int T = std::max(1, n_threads);
        std::vector<std::vector<uint32_t>> pi(T), bi(T);
        std::vector<std::vector<uint8_t>> mt(track_right ? T : 0, std::vector<uint8_t>((size_t)t.row_count + 1, 0));
        if (!build_null)
            parallel_ranges(probe->n_rows, T, [&](int th, int64_t a, int64_t b) { probe_range(plan, t, pk, a, b, pi[th], bi[th], track_right ? &mt[th] : nullptr); });
        if (track_right) {
            matched.assign((size_t)t.row_count + 1, 0);
            for (auto& m : mt)
                for (size_t i = 0; i < m.size(); i++) matched[i] |= m[i];
        }
        size_t total = 0;
        for (int th = 0; th < T; th++) total += pi[th].size();
        probe_index.reserve(total);
        build_index.reserve(total);
        for (int th = 0; th < T; th++) {
            probe_index.insert(probe_index.end(), pi[th].begin(), pi[th].end());
            build_index.insert(build_index.end(), bi[th].begin(), bi[th].end());
        }
    }
    // _search_ht_remain: the build rows that never matched (RIGHT / FULL OUTER, RIGHT ANTI) or did (RIGHT SEMI), in build order; rows with
    // NULL keys were never inserted and therefore never match
    if (track_right) {
        const bool want = jt == SRB_JOIN_RIGHT_SEMI;
        for (uint32_t b = 1; b <= t.row_count; b++)
            if ((matched[b] != 0) == want) {
                probe_index.push_back(NO_ROW);
                build_index.push_back(b);
            }
    }
    // other-join-conjuncts of an INNER join filter the joined rows (HashJoiner::_process_other_conjunct, hash_joiner.cpp:488-507:
    // ExecNode::eval_conjuncts over the joined chunk; a comparison with a NULL is not true)
    if (plan->n_conjuncts > 0) {
        if (jt != SRB_JOIN_INNER) throw std::runtime_error("other-join-conjuncts: INNER joins only");
        auto side_of = [&](int slot, const srb_column** col) { // 0: probe column, 1: build column
            for (int i = 0; i < probe->n_cols; i++)
                if (probe->slot_ids[i] == slot) {
                    *col = &probe->cols[i];
                    return 0;
                }
            *col = find_slot(*build, slot);
            return 1;
        };
        size_t kept = 0;
        for (size_t j = 0; j < probe_index.size(); j++) {
            bool ok = true;
            for (int c = 0; c < plan->n_conjuncts && ok; c++) {
                const srb_join_conjunct& cj = plan->conjuncts[c];
                const srb_column *lc, *rc;
                const int ls = side_of(cj.left_slot, &lc), rs = side_of(cj.right_slot, &rc);
                const int64_t lrow = ls ? (int64_t)build_index[j] - 1 : (int64_t)probe_index[j], rrow = rs ? (int64_t)build_index[j] - 1 : (int64_t)probe_index[j];
                if ((lc->nulls && lc->nulls[lrow]) || (rc->nulls && rc->nulls[rrow])) {
                    ok = false;
                    break;
                }
                ok = cmp_apply(cj.op, load_int(*lc, lrow), load_int(*rc, rrow));
            }
            if (ok) {
                probe_index[kept] = probe_index[j];
                build_index[kept] = build_index[j];
                kept++;
            }
        }
        probe_index.resize(kept);
        build_index.resize(kept);
    }
    r->rows_in = probe->n_rows;
    r->rows_selected = (int64_t)probe_index.size();
    r->n_rows = (int64_t)probe_index.size();
    std::vector<uint32_t> build_rows(build_index.size());
    for (size_t i = 0; i < build_index.size(); i++) build_rows[i] = build_index[i] == 0 ? NO_ROW : build_index[i] - 1;
    const bool right_only = jt == SRB_JOIN_RIGHT_SEMI || jt == SRB_JOIN_RIGHT_ANTI;
    if (!right_only)
        for (int i = 0; i < plan->n_probe_out; i++) {
            r->cols.emplace_back();
            r->slot_ids.push_back(plan->probe_out_slots[i]);
            copy_rows(*find_slot(*probe, plan->probe_out_slots[i]), probe_index, r->cols.back(), n_threads, left_may_miss);
        }
    if (plan->join_type != SRB_JOIN_LEFT_SEMI && !anti)
        for (int i = 0; i < plan->n_build_out; i++) {
            r->cols.emplace_back();
            r->slot_ids.push_back(plan->build_out_slots[i]);
            copy_rows(*find_slot(*build, plan->build_out_slots[i]), build_rows, r->cols.back(), n_threads, right_may_miss);
        }
    if (plan->emit_indexes && !anti && !track_right && !right_may_miss) {
        for (int which = 0; which < 2; which++) {
            r->cols.emplace_back();
            OutCol& c = r->cols.back();
            c.type = SRB_TYPE_INT;
            r->slot_ids.push_back(which == 0 ? -4 : -3);
            for (uint32_t x : (which == 0 ? probe_index : build_index)) c.push_int((int128_t)x);
        }
    }
    r->finish();
    return r;
}

// packed fixed-size keys of a chunk, 16 bytes per row (CompileFixedSizeKeyColumn, join_hash_map_test.cpp:1041-1056)
int pack_keys_raw(const srb_chunk* ch, const int32_t* slots, int n_keys, void* out16) {
    PackedKeys pk = pack_keys(*ch, slots, n_keys);
    memcpy(out16, pk.k.data(), 16 * pk.k.size());
    return pk.width;
}

// raw table for the reference's unit vectors (be/test/exec/join_hash_map_test.cpp:1126-1478)
void join_build_raw(const void* keys, int width, uint32_t n, uint32_t* bucket_size, uint32_t* first, uint32_t* next) {
    JoinTable t;
    t.keys.width = width;
    t.keys.k.assign(n, 0);
    t.keys.null.assign(n, 0);
    for (uint32_t i = 0; i < n; i++) memcpy(&t.keys.k[i], (const uint8_t*)keys + (size_t)width * i, width);
    build_table(t);
    *bucket_size = t.bucket_size;
    if (first) memcpy(first, t.first.data(), sizeof(uint32_t) * t.bucket_size);
    if (next) memcpy(next, t.next.data(), sizeof(uint32_t) * ((size_t)n + 1));
}

} // namespace orc
