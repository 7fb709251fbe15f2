// srb_rowpipe.cuh — the per-row pipeline every scan/aggregate kernel is built from:
//   column load -> predicates -> expression nodes -> packed group key -> group-table slot -> state update.
//
// Reference semantics restated here (no reference code is used):
//   ColumnPredicateCmpBase::t_evaluate      be/src/storage/column_predicate_cmp.cpp:222-239  (NULL never satisfies a compare)
//   ConjunctivePredicates / OR groups       be/src/storage/conjunctive_predicates.cpp:35-63
//   VectorizedArithmeticExpr / DecimalBinaryFunction  be/src/exprs/arithmetic_expr.cpp:70-128, exprs/decimal_binary_function.h:47-213
//   AggHashMap*::compute_agg_states         be/src/exec/aggregate/agg_hash_map.h:228-339,676-818
//   SUM / AVG / COUNT / MIN / MAX update and merge   be/src/exprs/agg/{sum.h:58-117,avg.h:85-166,count.h:46-96,maxmin.h:30-122}
//   NullableAggregateFunctionUnary (skip NULL rows)   be/src/exprs/agg/nullable_aggregate.h:340-400
//
// Every function is templated on a program accessor P: RuntimeProg reads the DProg the host lowered; a static shape
// (srb_shapes.cuh) returns compile-time constants from the same accessors so that loops unroll and the switches fold.
#pragma once
#if !defined(__CUDACC_RTC__)
#include <cuda_runtime.h>
#endif

#include "srb_prog.h"

namespace srb {

// ---- program accessors -----------------------------------------------------------------------------
struct RuntimeProg {
    const DProg* p;
    static constexpr bool kStatic = false;
    static constexpr int kMaxNodes = MAX_NODES, kMaxAggs = MAX_AGGS, kMaxKeys = MAX_KEYS;
    __device__ __forceinline__ int n_nodes() const { return p->n_nodes; }
    __device__ __forceinline__ int n_aggs() const { return p->n_aggs; }
    __device__ __forceinline__ int n_keys() const { return p->n_keys; }
    __device__ __forceinline__ const DNode& node(int i) const { return p->nodes[i]; }
    __device__ __forceinline__ const DAgg& agg(int i) const { return p->aggs[i]; }
    __device__ __forceinline__ const DKey& key(int i) const { return p->keys[i]; }
};

// ---- relaxed / coherent global accesses used by the group table ------------------------------------------
SRB_D uint64_t ld_relaxed_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
SRB_D void st_relaxed_u64(uint64_t* p, uint64_t v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
SRB_D void st_release_u64(uint64_t* p, uint64_t v) { asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

// ---- column loads ------------------------------------------------------------------------------------
SRB_D Val load_col(const ColBind& c, int64_t row) {
    Val v;
    v.null = c.nulls != nullptr && c.nulls[row] != 0;
    switch (c.width) {
    case 1: {
        int8_t x = ((const int8_t*)c.data)[row];
        v.lo = (uint64_t)(int64_t)x;
        v.hi = x < 0 ? -1 : 0;
        break;
    }
    case 2: {
        int16_t x = ((const int16_t*)c.data)[row];
        v.lo = (uint64_t)(int64_t)x;
        v.hi = x < 0 ? -1 : 0;
        break;
    }
    case 4: {
        int32_t x = ((const int32_t*)c.data)[row];
        v.lo = (uint64_t)(int64_t)x;
        v.hi = x < 0 ? -1 : 0;
        break;
    }
    case 8: {
        int64_t x = ((const int64_t*)c.data)[row];
        v.lo = (uint64_t)x;
        v.hi = (x < 0 && c.kind != CK_DOUBLE) ? -1 : 0;
        break;
    }
    case 16: {
        const ulonglong2 x = ((const ulonglong2*)c.data)[row];
        v.lo = x.x;
        v.hi = (int64_t)x.y;
        break;
    }
    default: v.lo = 0; v.hi = 0;
    }
    return v;
}

SRB_D bool cmp_i128(int op, i128 a, i128 b) {
    switch (op) {
    case 1: return a == b;
    case 2: return a != b;
    case 3: return a > b;
    case 4: return a >= b;
    case 5: return a < b;
    default: return a <= b;
    }
}
SRB_D bool cmp_f64(int op, double a, double b) {
    switch (op) {
    case 1: return a == b;
    case 2: return a != b;
    case 3: return a > b;
    case 4: return a >= b;
    case 5: return a < b;
    default: return a <= b;
    }
}

// Slice::compare of a BinaryColumn value with a literal of <= 16 bytes: memcmp over the common length, then the lengths
SRB_D int cmp_bytes(const uint8_t* __restrict__ s, uint32_t n, uint64_t lit_lo, uint64_t lit_hi, uint32_t m) {
    const uint32_t k = n < m ? n : m;
    for (uint32_t i = 0; i < k; i++) {
        const uint8_t b = (uint8_t)((i < 8 ? lit_lo >> (8 * i) : lit_hi >> (8 * (i - 8))) & 0xFF);
        if (s[i] != b) return s[i] < b ? -1 : 1;
    }
    return n == m ? 0 : (n < m ? -1 : 1);
}

SRB_D bool eval_one_pred(const DProg& p, const DPred& pr, int64_t row) {
    const ColBind& c = p.cols[pr.col];
    if (c.kind == CK_BINARY) {
        const bool isnull = c.nulls && c.nulls[row];
        if (pr.op == 9) return isnull;
        if (pr.op == 10) return !isnull;
        if (isnull) return false;
        const uint32_t b = c.offsets[row], e = c.offsets[row + 1];
        const int r = cmp_bytes((const uint8_t*)c.data + b, e - b, pr.lit_lo, (uint64_t)pr.lit_hi, (uint32_t)pr.lit_len);
        switch (pr.op) {
        case 1: return r == 0;
        case 2: return r != 0;
        case 3: return r > 0;
        case 4: return r >= 0;
        case 5: return r < 0;
        default: return r <= 0;
        }
    }
    Val v = load_col(c, row);
    if (pr.op == 9) return v.null;
    if (pr.op == 10) return !v.null;
    bool r;
    if (c.kind == CK_DOUBLE)
        r = cmp_f64(pr.op, __longlong_as_double((long long)v.lo), __longlong_as_double((long long)pr.lit_lo));
    else
        r = cmp_i128(pr.op, make_i128(v.lo, v.hi), make_i128(pr.lit_lo, pr.lit_hi));
    return !v.null && r;
}

// AND of conjuncts and OR-groups (groups are contiguous after lowering)
SRB_D bool eval_preds(const DProg& p, int64_t row) {
    bool pass = true;
    int i = 0;
    const int n = p.n_preds;
    while (i < n) {
        if (p.preds[i].group == 0) {
            pass = pass && eval_one_pred(p, p.preds[i], row);
            i++;
        } else {
            int g = p.preds[i].group;
            bool any = false;
            while (i < n && p.preds[i].group == g) {
                any = any || eval_one_pred(p, p.preds[i], row);
                i++;
            }
            pass = pass && any;
        }
    }
    return pass;
}

// ---- expression nodes ---------------------------------------------------------------------------------
template <int W>
SRB_D bool scale_up_checked(i128 x, i128 f, i128* out) {
    return checked_binop<W>(6, x, f, out);
}

SRB_D bool checked_binop_w(int width, int op, i128 a, i128 b, i128* out) {
    if (width == 16) return checked_binop<16>(op, a, b, out);
    if (width == 8) return checked_binop<8>(op, a, b, out);
    return checked_binop<4>(op, a, b, out);
}
SRB_D i128 wrap_w(int width, i128 x) {
    switch (width) {
    case 1: return wrap_to<1>(x);
    case 2: return wrap_to<2>(x);
    case 4: return wrap_to<4>(x);
    case 8: return wrap_to<8>(x);
    }
    return x;
}

// evaluates node `nd` given already evaluated operands in v[]; `cols` are the bound columns
SRB_D Val eval_node(const DNode& nd, const Val* v, const ColBind* cols, int64_t row) {
    Val out;
    switch (nd.op) {
    case DOP_LOAD: out = load_col(cols[nd.col], row); break;
    case DOP_LIT:
        out.lo = nd.lit_lo;
        out.hi = nd.lit_hi;
        out.null = false;
        break;
    case DOP_CAST: {
        const Val a = v[nd.lhs];
        i128 x = make_i128(a.lo, a.hi), r = x;
        bool ov = false;
        i128 f = make_i128(nd.lit_lo, nd.lit_hi);
        if (nd.checked) {
            if (f != 1) ov = checked_binop_w(nd.width, 6, x, f, &r);
            if (nd.width == 8) ov = ov || !fits<8>(x);
            if (nd.width == 4) ov = ov || !fits<4>(x);
        }
        out.lo = lo64(r);
        out.hi = hi64(r);
        out.null = a.null || ov;
        break;
    }
    case DOP_MULFAST: {
        const Val a = v[nd.lhs], b = v[nd.rhs];
        i128 r = mul_i64_i64((int64_t)a.lo, (int64_t)b.lo);
        out.lo = lo64(r);
        out.hi = hi64(r);
        out.null = a.null || b.null;
        break;
    }
    default: { // ADD / SUB / MUL
        const Val a = v[nd.lhs], b = v[nd.rhs];
        i128 x = make_i128(a.lo, a.hi), y = make_i128(b.lo, b.hi), r = 0;
        bool ov = false;
        if (nd.checked) {
            if (nd.adj_l) ov |= checked_binop_w(nd.width, 6, x, make_i128(nd.fl_lo, nd.fl_hi), &x);
            if (nd.adj_r) ov |= checked_binop_w(nd.width, 6, y, make_i128(nd.fr_lo, nd.fr_hi), &y);
            if (!ov) ov = checked_binop_w(nd.width, nd.op, x, y, &r);
            if (ov) r = 0;
        } else {
            u128 ux = (u128)x, uy = (u128)y;
            u128 ur = nd.op == DOP_ADD ? ux + uy : (nd.op == DOP_SUB ? ux - uy : ux * uy);
            r = wrap_w(nd.width, (i128)ur);
        }
        out.lo = lo64(r);
        out.hi = hi64(r);
        out.null = a.null || b.null || ov;
        break;
    }
    }
    return out;
}

// ---- group table ---------------------------------------------------------------------------------------
// 16-byte accesses on a 16-byte aligned pair: one L2 transaction, so a reader sees either the empty pair or the complete one
SRB_D ulonglong2 ld_pair(const void* p) {
    ulonglong2 v;
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
SRB_D void st_pair(void* p, uint64_t a, uint64_t b) { asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory"); }
SRB_D ulonglong2 cas_pair(void* p, uint64_t c0, uint64_t c1, uint64_t n0, uint64_t n1) {
    ulonglong2 o;
    asm volatile(
            "{\n\t.reg .b128 c, n, o;\n\t"
            "mov.b128 c, {%2, %3};\n\t"
            "mov.b128 n, {%4, %5};\n\t"
            "atom.global.relaxed.gpu.cas.b128 o, [%6], c, n;\n\t"
            "mov.b128 {%0, %1}, o;\n\t}"
            : "=l"(o.x), "=l"(o.y)
            : "l"(c0), "l"(c1), "l"(n0), "l"(n1), "l"(p)
            : "memory");
    return o;
}

// Open addressing, linear probing. Returns the slot of (klo,khi), inserting it when absent (*inserted = true); nullptr when
// the probe ran past TABLE_MAX_PROBE slots (table too full: the row is deferred and replayed after the host grew the table).
SRB_D uint8_t* table_find_or_insert(const TableView& t, uint64_t klo, uint64_t khi, uint64_t hash, bool* inserted) {
    const uint64_t tagw = (hash & ~3ULL) | SLOT_READY;
    uint64_t idx = (hash >> 7) & t.mask;
    *inserted = false;
    for (int probe = 0; probe < TABLE_MAX_PROBE; probe++) {
        uint8_t* slot = t.slots + idx * (uint64_t)t.slot_bytes;
        ulonglong2 a = ld_pair(slot);
        if (a.x == SLOT_EMPTY) {
            a = cas_pair(slot, 0ULL, 0ULL, tagw, klo);
            if (a.x == SLOT_EMPTY) { // claimed: tag and first key word are already visible as one unit
                if (t.key_words > 1) st_pair(slot + 16, khi, 1ULL);
                *inserted = true;
                return slot;
            }
        }
        if (a.x == tagw && a.y == klo) {
            if (t.key_words == 1) return slot;
            ulonglong2 b = ld_pair(slot + 16);
            while (b.y == 0) { // the claimer's second pair is in flight
                __nanosleep(20);
                b = ld_pair(slot + 16);
            }
            if (b.x == khi) return slot;
        }
        idx = (idx + 1) & t.mask;
    }
    return nullptr;
}

// Lookup without insertion (the streaming sink's build_hash_map_with_selection, agg_hash_map.h:298-339): the slot of (klo,khi) or
// nullptr when the key is absent (or further than TABLE_MAX_PROBE slots from home, where no insert would have put it).
SRB_D uint8_t* table_find(const TableView& t, uint64_t klo, uint64_t khi, uint64_t hash) {
    const uint64_t tagw = (hash & ~3ULL) | SLOT_READY;
    uint64_t idx = (hash >> 7) & t.mask;
    for (int probe = 0; probe < TABLE_MAX_PROBE; probe++) {
        uint8_t* slot = t.slots + idx * (uint64_t)t.slot_bytes;
        const ulonglong2 a = ld_pair(slot);
        if (a.x == SLOT_EMPTY) return nullptr;
        if (a.x == tagw && a.y == klo) {
            if (t.key_words == 1) return slot;
            ulonglong2 b = ld_pair(slot + 16);
            while (b.y == 0) { // the claimer's second pair is in flight
                __nanosleep(20);
                b = ld_pair(slot + 16);
            }
            if (b.x == khi) return slot;
        }
        idx = (idx + 1) & t.mask;
    }
    return nullptr;
}

// ---- packed group key ----------------------------------------------------------------------------------
// per key column: [null byte if nullable][value bytes | u32 len + bytes] (AggHashMapWithSerializedKey,
// agg_hash_map.h:746-748; BinaryColumn::serialize, column/binary_column.h:228-237). Up to 16 bytes.
struct PackedKey {
    uint64_t lo, hi;
    int len;
    bool too_long;
};
SRB_D void key_put(PackedKey& k, uint64_t bits, int nbytes) {
    // append the low `nbytes` (<= 8) bytes of `bits`
    if (nbytes < 8) bits &= (1ULL << (8 * nbytes)) - 1;
    const int off = k.len;
    if (off + nbytes > 16) {
        k.too_long = true;
        return;
    }
    if (off < 8) {
        k.lo |= bits << (8 * off);
        if (off + nbytes > 8 && off > 0) k.hi |= bits >> (8 * (8 - off));
    } else if (nbytes) {
        k.hi |= bits << (8 * (off - 8));
    }
    k.len = off + nbytes;
}
// [u32 length][bytes] of a BinaryColumn value (BinaryColumn::serialize, column/binary_column.h:228-237); a key beyond 16 bytes sets too_long.
// (Gathering the bytes into words first and appending those was measured slower on the static kernel: 1.98 vs 1.76 ms per 134 M rows of Q1
// with VARCHAR(1) keys, and so was a branch-free 128-bit shift in key_put.)
SRB_D void key_put_binary(PackedKey& k, const uint8_t* first, uint32_t len) {
    key_put(k, len, 4);
    for (uint32_t j = 0; j < len && !k.too_long; j++) key_put(k, first[j], 1);
}

SRB_D PackedKey pack_key(const DProg& p, const Val* v, int64_t row) {
    PackedKey k{0, 0, 0, false};
    for (int i = 0; i < p.n_keys; i++) {
        const DKey& kd = p.keys[i];
        const DNode& nd = p.nodes[kd.node];
        const ColBind& c = p.cols[nd.col];
        bool isnull = c.nulls != nullptr && c.nulls[row] != 0;
        if (kd.nullable) key_put(k, isnull ? 1 : 0, 1);
        if (kd.width == 0) { // binary
            uint32_t b = c.offsets[row], e = c.offsets[row + 1];
            uint32_t len = isnull ? 0 : e - b;
            key_put_binary(k, (const uint8_t*)c.data + b, len);
        } else {
            const Val& x = v[kd.node];
            uint64_t lo = isnull ? 0 : x.lo, hi = isnull ? 0 : (uint64_t)x.hi;
            if (kd.width <= 8)
                key_put(k, lo, kd.width);
            else {
                key_put(k, lo, 8);
                key_put(k, hi, 8);
            }
        }
    }
    return k;
}

// ---- wide keys: 17 to 64 packed bytes (three BIGINT keys, a DECIMAL128 beside a date, a VARCHAR(40)) -----------------------------------
// Same serialisation (AggHashMapWithSerializedKey, agg_hash_map.h:652-898: the key columns' bytes one after the other), more words. The
// slot keeps word 0 in pair A and word 1 + the ready flag in pair B as a 16-byte key does; words 2.. follow pair B and are written BEFORE
// pair B is published, so a reader that has seen the flag (and fenced) sees them. Only the interpreter takes these keys.
struct WideKey {
    uint64_t w[MAX_KEY_WORDS];
    int len;
    bool too_long;
};
SRB_D void wide_put(WideKey& k, uint64_t bits, int nbytes, int cap_bytes) { // append the low `nbytes` (<= 8) bytes of `bits`
    if (nbytes < 8) bits &= (1ULL << (8 * nbytes)) - 1;
    const int off = k.len;
    if (off + nbytes > cap_bytes) {
        k.too_long = true;
        return;
    }
    const int word = off >> 3, sh = (off & 7) * 8;
    k.w[word] |= bits << sh;
    if (sh && (off & 7) + nbytes > 8) k.w[word + 1] |= bits >> (64 - sh);
    k.len = off + nbytes;
}
SRB_D WideKey pack_key_wide(const DProg& p, const Val* v, int64_t row) {
    WideKey k;
#pragma unroll
    for (int i = 0; i < MAX_KEY_WORDS; i++) k.w[i] = 0;
    k.len = 0;
    k.too_long = false;
    const int cap = p.key_words * 8;
    for (int i = 0; i < p.n_keys; i++) {
        const DKey& kd = p.keys[i];
        const DNode& nd = p.nodes[kd.node];
        const ColBind& c = p.cols[nd.col];
        const bool isnull = c.nulls != nullptr && c.nulls[row] != 0;
        if (kd.nullable) wide_put(k, isnull ? 1 : 0, 1, cap);
        if (kd.width == 0) { // binary: [u32 length][bytes]
            const uint32_t b = c.offsets[row], e = c.offsets[row + 1];
            const uint32_t len = isnull ? 0 : e - b;
            wide_put(k, len, 4, cap);
            const uint8_t* s = (const uint8_t*)c.data + b;
            for (uint32_t j = 0; j < len && !k.too_long; j++) wide_put(k, s[j], 1, cap);
        } else {
            const Val& x = v[kd.node];
            const uint64_t lo = isnull ? 0 : x.lo, hi = isnull ? 0 : (uint64_t)x.hi;
            if (kd.width <= 8) {
                wide_put(k, lo, kd.width, cap);
            } else {
                wide_put(k, lo, 8, cap);
                wide_put(k, hi, 8, cap);
            }
        }
    }
    return k;
}
SRB_D uint64_t hash_wide_key(const WideKey& k, int seed) { // SliceHashWithSeed over the serialized bytes (crc_hash_64, util/hash.h:139-170)
    const uint64_t h = crc_hash_64((const uint8_t*)k.w, k.len, seed == 1 ? 0x811C9DC5ULL : 0x811C9DD7ULL);
    return h * 0x9E3779B97F4A7C15ULL; // crc values are 32 bits wide: spread them over tag and bucket index
}
SRB_D bool wide_rest_equal(const uint8_t* slot, const WideKey& k, int key_words) {
    bool eq = true;
    for (int i = 2; i < key_words; i++) eq = eq && ld_relaxed_u64((const uint64_t*)(slot + 32) + (i - 2)) == k.w[i];
    return eq;
}
SRB_D uint8_t* table_find_or_insert_wide(const TableView& t, const WideKey& k, uint64_t hash, bool* inserted) {
    const uint64_t tagw = (hash & ~3ULL) | SLOT_READY;
    uint64_t idx = (hash >> 7) & t.mask;
    *inserted = false;
    for (int probe = 0; probe < TABLE_MAX_PROBE; probe++) {
        uint8_t* slot = t.slots + idx * (uint64_t)t.slot_bytes;
        ulonglong2 a = ld_pair(slot);
        if (a.x == SLOT_EMPTY) {
            a = cas_pair(slot, 0ULL, 0ULL, tagw, k.w[0]);
            if (a.x == SLOT_EMPTY) {
                for (int i = 2; i < t.key_words; i++) st_relaxed_u64((uint64_t*)(slot + 32) + (i - 2), k.w[i]);
                __threadfence(); // words 2.. are visible before the flag of pair B is
                st_pair(slot + 16, k.w[1], 1ULL);
                *inserted = true;
                return slot;
            }
        }
        if (a.x == tagw && a.y == k.w[0]) {
            ulonglong2 b = ld_pair(slot + 16);
            while (b.y == 0) { // the claimer's further words are in flight
                __nanosleep(20);
                b = ld_pair(slot + 16);
            }
            __threadfence();
            if (b.x == k.w[1] && wide_rest_equal(slot, k, t.key_words)) return slot;
        }
        idx = (idx + 1) & t.mask;
    }
    return nullptr;
}
SRB_D uint8_t* table_find_wide(const TableView& t, const WideKey& k, uint64_t hash) {
    const uint64_t tagw = (hash & ~3ULL) | SLOT_READY;
    uint64_t idx = (hash >> 7) & t.mask;
    for (int probe = 0; probe < TABLE_MAX_PROBE; probe++) {
        uint8_t* slot = t.slots + idx * (uint64_t)t.slot_bytes;
        const ulonglong2 a = ld_pair(slot);
        if (a.x == SLOT_EMPTY) return nullptr;
        if (a.x == tagw && a.y == k.w[0]) {
            ulonglong2 b = ld_pair(slot + 16);
            while (b.y == 0) {
                __nanosleep(20);
                b = ld_pair(slot + 16);
            }
            __threadfence();
            if (b.x == k.w[1] && wide_rest_equal(slot, k, t.key_words)) return slot;
        }
        idx = (idx + 1) & t.mask;
    }
    return nullptr;
}

// hash of the packed key with the reference's functions (agg_hash_map.h:58-98, column/hash_set.h:86-159):
// 4/8-byte keys as one integer through phmap_mix, 16-byte through Hash128WithSeed, anything else crc_hash_64.
SRB_D uint64_t hash_packed_key(const PackedKey& k, int seed) {
    if (k.len == 4) return phmap_mix8((uint64_t)(int64_t)(int32_t)(uint32_t)k.lo, seed);
    if (k.len == 8) return phmap_mix8(k.lo, seed);
    if (k.len == 16) return hash128_with_seed(k.lo, k.hi, seed);
    uint8_t b[16];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        b[i] = (uint8_t)(k.lo >> (8 * i));
        b[8 + i] = (uint8_t)(k.hi >> (8 * i));
    }
    uint64_t h = crc_hash_64(b, k.len, seed == 1 ? 0x811C9DC5ULL : 0x811C9DD7ULL);
    // crc values are 32 bits wide; spread them so that the tag and the bucket index do not alias
    return h * 0x9E3779B97F4A7C15ULL;
}

// ---- state updates on a table slot (global atomics) --------------------------------------------------------
SRB_D void atomic_add128(uint8_t* slot, int off_lo, int off_hi, uint64_t lo, uint64_t hi) {
    if (off_hi < 0) {
        atomicAdd((unsigned long long*)(slot + off_lo), (unsigned long long)lo);
        return;
    }
    unsigned long long old = atomicAdd((unsigned long long*)(slot + off_lo), (unsigned long long)lo);
    unsigned long long carry = (old + lo < old) ? 1ULL : 0ULL;
    unsigned long long add_hi = hi + carry;
    if (add_hi) atomicAdd((unsigned long long*)(slot + off_hi), add_hi);
}

// MIN / MAX states hold order-preserving unsigned encodings whose identity is all-zero bits, so a zero-filled slot is a
// valid empty state: MAX keeps enc(x) = x ^ signbit, MIN keeps ~enc(x); both are folded with an unsigned atomic max.
SRB_HD uint64_t enc_i64(int64_t x, bool is_min) {
    uint64_t u = (uint64_t)x ^ 0x8000000000000000ULL;
    return is_min ? ~u : u;
}
SRB_HD int64_t dec_i64(uint64_t u, bool is_min) { return (int64_t)((is_min ? ~u : u) ^ 0x8000000000000000ULL); }
SRB_HD uint64_t enc_f64(uint64_t bits, bool is_min) {
    uint64_t u = bits ^ ((bits >> 63) ? ~0ULL : 0x8000000000000000ULL);
    return is_min ? ~u : u;
}
SRB_HD uint64_t dec_f64(uint64_t u, bool is_min) {
    if (is_min) u = ~u;
    return u ^ ((u >> 63) ? 0x8000000000000000ULL : ~0ULL);
}
SRB_HD u128 enc_i128(i128 x, bool is_min) {
    u128 u = (u128)x ^ ((u128)1 << 127);
    return is_min ? ~u : u;
}
SRB_HD i128 dec_i128(u128 u, bool is_min) { return (i128)((is_min ? ~u : u) ^ ((u128)1 << 127)); }

SRB_D void atomic_umax128(uint8_t* slot, int off_lo, u128 x) {
    // 16-byte aligned (lo,hi) pair, compare-and-swap loop on the whole pair
    ulonglong2 cur = ld_pair(slot + off_lo);
    for (;;) {
        u128 c = ((u128)cur.y << 64) | cur.x;
        if (!(x > c)) return;
        ulonglong2 o = cas_pair(slot + off_lo, cur.x, cur.y, (uint64_t)x, (uint64_t)(x >> 64));
        if (o.x == cur.x && o.y == cur.y) return;
        cur = o;
    }
}

SRB_D void update_agg_global(const DProg& p, const DAgg& g, uint8_t* slot, const Val* v, int64_t row) {
    if (g.fn == FN_COUNT) {
        if (g.merge) {
            const Val& x = v[g.arg];
            if (!x.null) atomicAdd((unsigned long long*)(slot + g.off_n), (unsigned long long)x.lo);
        } else if (g.arg < 0 || !v[g.arg].null) {
            atomicAdd((unsigned long long*)(slot + g.off_n), 1ULL);
        }
        return;
    }
    if (g.merge && g.fn == FN_AVG) { // {sum,count} blob (avg.h:168-181)
        const ColBind& c = p.cols[g.blob_col];
        if (c.nulls && c.nulls[row]) return;
        uint32_t b = c.offsets[row], e = c.offsets[row + 1];
        const uint8_t* s = (const uint8_t*)c.data + b;
        uint64_t w[3] = {0, 0, 0};
        int nw = (e - b) / 8;
        for (int i = 0; i < nw && i < 3; i++) w[i] = load_u64_le(s + 8 * i);
        if (e - b == 24) {
            atomic_add128(slot, g.off_lo, g.off_hi, w[0], w[1]);
            atomicAdd((unsigned long long*)(slot + g.off_n), (unsigned long long)w[2]);
        } else if (e - b == 16) {
            atomicAdd((double*)(slot + g.off_lo), __longlong_as_double((long long)w[0]));
            atomicAdd((unsigned long long*)(slot + g.off_n), (unsigned long long)w[1]);
        }
        return;
    }
    const Val& x = v[g.arg];
    if (x.null) return;
    switch (g.fn) {
    case FN_SUM:
    case FN_AVG:
        if (g.is_double)
            atomicAdd((double*)(slot + g.off_lo), __longlong_as_double((long long)x.lo));
        else if (g.avg_as_double)
            atomicAdd((double*)(slot + g.off_lo), (double)(int64_t)x.lo);
        else
            atomic_add128(slot, g.off_lo, g.off_hi, x.lo, (uint64_t)x.hi);
        break;
    default: {
        const bool mn = g.fn == FN_MIN;
        if (g.is_double)
            atomicMax((unsigned long long*)(slot + g.off_lo), (unsigned long long)enc_f64(x.lo, mn));
        else if (g.wide)
            atomic_umax128(slot, g.off_lo, enc_i128(make_i128(x.lo, x.hi), mn));
        else
            atomicMax((unsigned long long*)(slot + g.off_lo), (unsigned long long)enc_i64((int64_t)x.lo, mn));
    }
    }
    if (g.off_n >= 0) atomicAdd((unsigned long long*)(slot + g.off_n), 1ULL);
}

} // namespace srb
