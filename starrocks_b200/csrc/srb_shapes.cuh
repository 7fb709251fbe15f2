// srb_shapes.cuh — statically specialised scan/aggregate kernels ("shapes") with TMA-staged column tiles.
//
// A shape is a compile-time copy of the STRUCTURE of a lowered program (srb_prog.h): expression ops, widths, operand
// links, aggregate list, key widths. Column bindings, literals, scale factors, predicates and slot offsets stay
// runtime data, so one shape serves every plan with that structure. The row pipeline is the same as the generic
// kernel's (same primitives from srb_dev.cuh), but because the structure is a template constant the interpreter
// folds into straight-line code and every intermediate lives in registers.
//
// Kernel design (per SM: one persistent CTA):
//   * every referenced column is streamed through a STAGES-deep shared-memory ring of TILE-row tiles filled by
//     cp.async.bulk (TMA, SASS UBLKCP) completing on an mbarrier — loads never occupy registers or LSU issue slots,
//     and memory-level parallelism is decoupled from occupancy (the kernel is register heavy);
//   * rows are read back with conflict-free LDS (lane = row), predicates and expressions evaluated in registers;
//   * GMAX > 0: low-cardinality GROUP BY — a per-CTA group directory (<= GMAX groups) in shared memory maps the
//     packed key to a register accumulator bank; adds are predicated per group (no dynamic register indexing, no
//     atomics per row); banks are warp-reduced and flushed with a handful of global atomics when the CTA retires.
//     Rows of groups that do not fit the directory take the global-table path, so the kernel is correct for any input.
//   * GMAX == 0: high-cardinality GROUP BY — every row probes the HBM table (same table code as the generic kernel).
// Reference loops replaced: ColumnPredicate evaluation (storage/column_predicate_cmp.cpp:222-239), DecimalBinaryFunction
// (exprs/decimal_binary_function.h:47-213), AggHashMap compute_agg_states + update_batch (exec/aggregate/agg_hash_map.h:228-339,
// exprs/agg/aggregate.h:403-408).
#pragma once
#include <utility>

#include "srb_rowpipe.cuh"

namespace srb {

struct Counters;

struct SNode {
    int op, width, lhs, rhs, checked, adj_l, adj_r;
};
struct SAgg {
    int fn, arg, wide, has_n; // wide: state keeps the high word (off_hi >= 0); has_n: state keeps its own count
};
struct SKey {
    int node, width;
};

// ---- the registered shapes ---------------------------------------------------------------------------------------
// TPC-H Q6 pipeline (fe/fe-core/src/test/resources/sql/external/hive/tpch/q6.sql:8,24):
//   sum(cast(l_extendedprice as D128) * cast(l_discount as D128)), no GROUP BY
struct ShapeSumMulFast {
    static constexpr const char* name = "sum(d64*d64) no-group";
    static constexpr int N_NODES = 5, N_AGGS = 1, N_KEYS = 0, GMAX = 1, THREADS = 512;
    static constexpr SNode node(int i) {
        constexpr SNode t[N_NODES] = {{DOP_LOAD, 8, -1, -1, 0, 0, 0},
                                      {DOP_LOAD, 8, -1, -1, 0, 0, 0},
                                      {DOP_CAST, 16, 0, -1, 1, 0, 0},
                                      {DOP_CAST, 16, 1, -1, 1, 0, 0},
                                      {DOP_MULFAST, 16, 0, 1, 0, 0, 0}};
        return t[i];
    }
    static constexpr SAgg agg(int) { return SAgg{FN_SUM, 4, 1, 1}; }
    static constexpr SKey key(int) { return SKey{0, 0}; }
};

// TPC-H Q1 pipeline with dictionary-coded keys (typed plan: .../tpch/q1.sql:71-102):
//   keys (l_returnflag, l_linestatus) as two INT codes;
//   sum(qty), sum(ext), sum(ext*(1-disc)), sum(ext*(1-disc)*(1+tax)), avg(qty), avg(ext), avg(disc), count(*)
struct ShapeQ1Dict {
    static constexpr const char* name = "q1 (2 int keys, 8 decimal aggregates)";
    static constexpr int N_NODES = 14, N_AGGS = 8, N_KEYS = 2, GMAX = 4, THREADS = 384;
    static constexpr SNode node(int i) {
        constexpr SNode t[N_NODES] = {
                {DOP_LOAD, 8, -1, -1, 0, 0, 0},    // 0 l_quantity
                {DOP_LOAD, 8, -1, -1, 0, 0, 0},    // 1 l_extendedprice
                {DOP_LOAD, 8, -1, -1, 0, 0, 0},    // 2 l_discount
                {DOP_LOAD, 8, -1, -1, 0, 0, 0},    // 3 l_tax
                {DOP_LOAD, 4, -1, -1, 0, 0, 0},    // 4 l_returnflag code
                {DOP_LOAD, 4, -1, -1, 0, 0, 0},    // 5 l_linestatus code
                {DOP_LIT, 8, -1, -1, 0, 0, 0},     // 6 literal 1 as DECIMAL64(16,2)
                {DOP_SUB, 8, 6, 2, 1, 0, 0},       // 7 1 - l_discount
                {DOP_CAST, 16, 1, -1, 1, 0, 0},    // 8
                {DOP_CAST, 16, 7, -1, 1, 0, 0},    // 9
                {DOP_MULFAST, 16, 1, 7, 0, 0, 0},  // 10 ext * (1 - disc)            DECIMAL128(31,4)
                {DOP_ADD, 8, 6, 3, 1, 0, 0},       // 11 1 + l_tax
                {DOP_CAST, 16, 11, -1, 1, 0, 0},   // 12
                {DOP_MUL, 16, 10, 12, 1, 0, 0}};   // 13 ext*(1-disc)*(1+tax)        DECIMAL128(38,6), overflow -> NULL
        return t[i];
    }
    static constexpr SAgg agg(int i) {
        constexpr SAgg t[N_AGGS] = {{FN_SUM, 0, 1, 0}, {FN_SUM, 1, 1, 0}, {FN_SUM, 10, 1, 1}, {FN_SUM, 13, 1, 1},
                                    {FN_AVG, 0, 1, 1}, {FN_AVG, 1, 1, 1}, {FN_AVG, 2, 1, 1},  {FN_COUNT, -1, 0, 1}};
        return t[i];
    }
    static constexpr SKey key(int i) { return i == 0 ? SKey{4, 4} : SKey{5, 4}; }
};

// High-cardinality GROUP BY: select k, sum(v), count(*) group by k   (k, v BIGINT)
struct ShapeHighCard {
    static constexpr const char* name = "group by bigint: sum(bigint), count(*)";
    static constexpr int N_NODES = 2, N_AGGS = 2, N_KEYS = 1, GMAX = 0, THREADS = 1024;
    static constexpr SNode node(int) { return SNode{DOP_LOAD, 8, -1, -1, 0, 0, 0}; }
    static constexpr SAgg agg(int i) { return i == 0 ? SAgg{FN_SUM, 1, 0, 0} : SAgg{FN_COUNT, -1, 0, 1}; }
    static constexpr SKey key(int) { return SKey{0, 8}; }
};

// ---- matching a lowered program against a shape ---------------------------------------------------------------------
template <class S>
inline bool shape_matches(const DProg& p) {
    if (p.n_nodes != S::N_NODES || p.n_aggs != S::N_AGGS || p.n_keys != S::N_KEYS) return false;
    for (int i = 0; i < S::N_NODES; i++) {
        const SNode s = S::node(i);
        const DNode& d = p.nodes[i];
        if (d.op != s.op || d.width != s.width || d.kind != CK_INT) return false;
        if (d.op == DOP_LOAD || d.op == DOP_LIT) continue;
        if (d.lhs != s.lhs || d.checked != s.checked) return false;
        if (d.op != DOP_CAST && (d.rhs != s.rhs || d.adj_l != s.adj_l || d.adj_r != s.adj_r)) return false;
    }
    for (int a = 0; a < S::N_AGGS; a++) {
        const SAgg s = S::agg(a);
        const DAgg& d = p.aggs[a];
        if (d.fn != s.fn || d.arg != s.arg || d.is_double || d.avg_as_double || d.merge) return false;
        if ((d.off_hi >= 0) != (s.wide != 0) || (d.off_n >= 0) != (s.has_n != 0)) return false;
    }
    for (int k = 0; k < S::N_KEYS; k++) {
        const SKey s = S::key(k);
        if (p.keys[k].node != s.node || p.keys[k].width != s.width || p.keys[k].nullable) return false;
    }
    for (int c = 0; c < p.n_cols; c++)
        if (p.cols[c].kind != CK_INT) return false;
    return true;
}

// ---- TMA / mbarrier primitives (inline PTX; SASS: UBLKCP, SYNCS) ------------------------------------------------------
SRB_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
SRB_D void mbar_init(uint64_t* bar, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
SRB_D void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
SRB_D void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "WAIT_%=:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
            "@p bra DONE_%=;\n\t"
            "bra WAIT_%=;\n\t"
            "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
            "r"(parity)
            : "memory");
}
SRB_D void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
SRB_D void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
SRB_D void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// ---- static row pipeline ---------------------------------------------------------------------------------------------------
constexpr int ST_TILE = 1024;  // rows per tile
constexpr int ST_STAGES = 4;

struct StaticArgs { // host-computed tile layout: column c sits at col_off[c] inside a stage of stage_bytes
    int32_t col_off[MAX_COLS];
    int32_t stage_bytes;
    int32_t _pad;
};
struct SmemCols {
    const uint8_t* stage; // current stage
    const int32_t* col_off;
    __device__ __forceinline__ const uint8_t* base(int c) const { return stage + col_off[c]; }
};

template <int W>
SRB_D Val lds_val(const uint8_t* col, int r) {
    Val v;
    v.null = false;
    if (W == 1) {
        int8_t x = ((const int8_t*)col)[r];
        v.lo = (uint64_t)(int64_t)x;
        v.hi = x < 0 ? -1 : 0;
    } else if (W == 2) {
        int16_t x = ((const int16_t*)col)[r];
        v.lo = (uint64_t)(int64_t)x;
        v.hi = x < 0 ? -1 : 0;
    } else if (W == 4) {
        int32_t x = ((const int32_t*)col)[r];
        v.lo = (uint64_t)(int64_t)x;
        v.hi = x < 0 ? -1 : 0;
    } else if (W == 8) {
        int64_t x = ((const int64_t*)col)[r];
        v.lo = (uint64_t)x;
        v.hi = x < 0 ? -1 : 0;
    } else {
        ulonglong2 x = ((const ulonglong2*)col)[r];
        v.lo = x.x;
        v.hi = (int64_t)x.y;
    }
    return v;
}
SRB_D Val lds_val_rt(const uint8_t* col, int width, int r) {
    switch (width) {
    case 1: return lds_val<1>(col, r);
    case 2: return lds_val<2>(col, r);
    case 4: return lds_val<4>(col, r);
    case 8: return lds_val<8>(col, r);
    default: return lds_val<16>(col, r);
    }
}

// predicates over the shared-memory tile (runtime list; same semantics as eval_preds)
SRB_D bool eval_preds_smem(const DProg& p, const SmemCols& sc, int r) {
    bool pass = true;
    int i = 0;
    const int n = p.n_preds;
    while (i < n) {
        const int g = p.preds[i].group;
        bool any = false;
        do {
            const DPred& pr = p.preds[i];
            Val v = lds_val_rt(sc.base(pr.col), p.cols[pr.col].width, r);
            bool res = pr.op == 9 ? false : (pr.op == 10 ? true : cmp_i128(pr.op, make_i128(v.lo, v.hi), make_i128(pr.lit_lo, pr.lit_hi)));
            any = any || res;
            i++;
        } while (g != 0 && i < n && p.preds[i].group == g);
        pass = pass && any;
    }
    return pass;
}

template <class S, int I>
SRB_D void eval_static_node(const DProg& p, Val* v, const SmemCols& sc, int r) {
    constexpr SNode nd = S::node(I);
    const DNode& rt = p.nodes[I];
    if constexpr (nd.op == DOP_LOAD) {
        v[I] = lds_val<nd.width>(sc.base(rt.col), r);
    } else if constexpr (nd.op == DOP_LIT) {
        v[I].lo = rt.lit_lo;
        v[I].hi = rt.lit_hi;
        v[I].null = false;
    } else if constexpr (nd.op == DOP_CAST) {
        const Val a = v[nd.lhs];
        i128 x = make_i128(a.lo, a.hi), res = x;
        bool ov = false;
        if constexpr (nd.checked != 0) {
            const i128 f = make_i128(rt.lit_lo, rt.lit_hi);
            if (f != 1) ov = checked_binop<nd.width>(6, x, f, &res);
            if constexpr (nd.width == 8) ov = ov || !fits<8>(x);
            if constexpr (nd.width == 4) ov = ov || !fits<4>(x);
        }
        v[I].lo = lo64(res);
        v[I].hi = hi64(res);
        v[I].null = a.null || ov;
    } else if constexpr (nd.op == DOP_MULFAST) {
        const Val a = v[nd.lhs], b = v[nd.rhs];
        i128 res = mul_i64_i64((int64_t)a.lo, (int64_t)b.lo);
        v[I].lo = lo64(res);
        v[I].hi = hi64(res);
        v[I].null = a.null || b.null;
    } else {
        const Val a = v[nd.lhs], b = v[nd.rhs];
        i128 x = make_i128(a.lo, a.hi), y = make_i128(b.lo, b.hi), res = 0;
        bool ov = false;
        if constexpr (nd.checked != 0) {
            if constexpr (nd.adj_l != 0) ov |= checked_binop<nd.width>(6, x, make_i128(rt.fl_lo, rt.fl_hi), &x);
            if constexpr (nd.adj_r != 0) ov |= checked_binop<nd.width>(6, y, make_i128(rt.fr_lo, rt.fr_hi), &y);
            if (!ov) ov = checked_binop<nd.width>(nd.op, x, y, &res);
            if (ov) res = 0;
        } else {
            u128 ux = (u128)x, uy = (u128)y;
            u128 ur = nd.op == DOP_ADD ? ux + uy : (nd.op == DOP_SUB ? ux - uy : ux * uy);
            res = wrap_to<nd.width>((i128)ur);
        }
        v[I].lo = lo64(res);
        v[I].hi = hi64(res);
        v[I].null = a.null || b.null || ov;
    }
}
template <class S, int... I>
SRB_D void eval_static_all(const DProg& p, Val* v, const SmemCols& sc, int r, std::integer_sequence<int, I...>) {
    (eval_static_node<S, I>(p, v, sc, r), ...);
}

// aggregate a is a "representative" when no earlier SUM/AVG aggregate has the same argument (their partial sums are identical)
template <class S>
constexpr int agg_rep(int a) {
    const SAgg g = S::agg(a);
    if (g.fn != FN_SUM && g.fn != FN_AVG) return a;
    for (int b = 0; b < a; b++) {
        const SAgg h = S::agg(b);
        if ((h.fn == FN_SUM || h.fn == FN_AVG) && h.arg == g.arg) return b;
    }
    return a;
}
// can the argument of aggregate a be NULL? (static shapes bind NOT NULL columns, so only checked ops produce NULLs)
template <class S>
constexpr bool node_can_be_null(int n) {
    if (n < 0) return false;
    const SNode nd = S::node(n);
    if (nd.checked) return true;
    return node_can_be_null<S>(nd.lhs) || node_can_be_null<S>(nd.rhs);
}

// register accumulator bank of one group
template <class S>
struct Bank {
    uint64_t lo[S::N_AGGS], hi[S::N_AGGS]; // only representatives are touched; the rest is dead and removed
    uint32_t n[S::N_AGGS];
    uint32_t rows;
};

template <class S, int A>
SRB_D void bank_update(Bank<S>& b, const Val* v, bool on) {
    constexpr SAgg g = S::agg(A);
    if constexpr (g.fn == FN_COUNT) {
        if constexpr (g.arg >= 0 && node_can_be_null<S>(g.arg)) {
            if (on && !v[g.arg].null) b.n[A]++;
        }
    } else if constexpr (agg_rep<S>(A) == A) {
        const Val& x = v[g.arg];
        bool take = on;
        if constexpr (node_can_be_null<S>(g.arg)) take = take && !x.null;
        const uint64_t alo = take ? x.lo : 0ULL, ahi = take ? (uint64_t)x.hi : 0ULL;
        const uint64_t nlo = b.lo[A] + alo;
        b.hi[A] += ahi + (nlo < alo ? 1ULL : 0ULL);
        b.lo[A] = nlo;
        if constexpr (node_can_be_null<S>(g.arg)) b.n[A] += take ? 1u : 0u;
    }
}
template <class S, int... A>
SRB_D void bank_update_all(Bank<S>& b, const Val* v, bool on, std::integer_sequence<int, A...>) {
    (bank_update<S, A>(b, v, on), ...);
    b.rows += on ? 1u : 0u;
}
template <class S, int... A>
SRB_D void bank_clear(Bank<S>& b, std::integer_sequence<int, A...>) {
    ((b.lo[A] = 0, b.hi[A] = 0, b.n[A] = 0), ...);
    b.rows = 0;
}

SRB_D uint64_t warp_sum_u64(uint64_t x) {
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    return x;
}
// 128-bit warp reduction: the low words are summed as three 32-bit-safe pieces so that carries are exact
SRB_D void warp_sum_u128(uint64_t& lo, uint64_t& hi) {
    for (int o = 16; o > 0; o >>= 1) {
        uint64_t olo = __shfl_down_sync(0xffffffffu, lo, o), ohi = __shfl_down_sync(0xffffffffu, hi, o);
        uint64_t nlo = lo + olo;
        hi += ohi + (nlo < olo ? 1ULL : 0ULL);
        lo = nlo;
    }
}

// flush one bank into table slot `slot` (lane 0 of each warp issues the atomics)
template <class S, int A>
SRB_D void bank_flush_one(const DProg& p, Bank<S>& b, uint8_t* slot, uint64_t rows_total) {
    constexpr SAgg g = S::agg(A);
    const DAgg& rt = p.aggs[A];
    const int lane = threadIdx.x & 31;
    if constexpr (g.fn == FN_COUNT) {
        uint64_t n = rows_total;
        if constexpr (g.arg >= 0 && node_can_be_null<S>(g.arg)) n = warp_sum_u64(b.n[A]);
        if (lane == 0 && n) atomicAdd((unsigned long long*)(slot + rt.off_n), (unsigned long long)n);
    } else {
        constexpr int R = agg_rep<S>(A);
        // the representative's registers are reduced once per representative and re-used by its followers
        uint64_t lo = b.lo[R], hi = b.hi[R];
        if constexpr (R == A) {
            warp_sum_u128(lo, hi);
            b.lo[R] = lo; // lane 0 now holds the warp total
            b.hi[R] = hi;
            if constexpr (node_can_be_null<S>(g.arg)) b.n[R] = (uint32_t)warp_sum_u64(b.n[R]);
        }
        uint64_t n = rows_total;
        if constexpr (node_can_be_null<S>(g.arg)) n = b.n[R];
        if (lane == 0 && n) {
            atomic_add128(slot, rt.off_lo, rt.off_hi, lo, hi);
            if (rt.off_n >= 0) atomicAdd((unsigned long long*)(slot + rt.off_n), (unsigned long long)n);
        }
    }
}
template <class S, int... A>
SRB_D void bank_flush(const DProg& p, Bank<S>& b, uint8_t* slot, std::integer_sequence<int, A...>) {
    const uint64_t rows_total = warp_sum_u64(b.rows);
    (bank_flush_one<S, A>(p, b, slot, rows_total), ...);
}

// per-row update of a table slot with compile-time operand indexes (static shapes carry SUM / AVG / COUNT over integers)
template <class S, int A>
SRB_D void row_update_global_one(const DProg& p, uint8_t* slot, const Val* v) {
    constexpr SAgg g = S::agg(A);
    const DAgg& rt = p.aggs[A];
    if constexpr (g.fn == FN_COUNT) {
        bool on = true;
        if constexpr (g.arg >= 0) on = !v[g.arg].null;
        if (on) atomicAdd((unsigned long long*)(slot + rt.off_n), 1ULL);
    } else {
        const Val& x = v[g.arg];
        if (!x.null) {
            atomic_add128(slot, rt.off_lo, rt.off_hi, x.lo, (uint64_t)x.hi);
            if (rt.off_n >= 0) atomicAdd((unsigned long long*)(slot + rt.off_n), 1ULL);
        }
    }
}
template <class S, int... A>
SRB_D void row_update_global(const DProg& p, uint8_t* slot, const Val* v, std::integer_sequence<int, A...>) {
    (row_update_global_one<S, A>(p, slot, v), ...);
}

template <class S>
constexpr int static_key_bytes() {
    int n = 0;
    for (int k = 0; k < S::N_KEYS; k++) n += S::key(k).width;
    return n;
}
// hash of a packed key whose length is a compile-time constant (same functions as hash_packed_key)
template <int LEN>
SRB_D uint64_t hash_packed_key_static(const PackedKey& k, int seed) {
    if constexpr (LEN == 4) return phmap_mix8((uint64_t)(int64_t)(int32_t)(uint32_t)k.lo, seed);
    if constexpr (LEN == 8) return phmap_mix8(k.lo, seed);
    if constexpr (LEN == 16) return hash128_with_seed(k.lo, k.hi, seed);
    return hash_packed_key(k, seed);
}

template <class S>
SRB_D PackedKey pack_key_static(const Val* v) {
    PackedKey k{0, 0, 0, false};
    if constexpr (S::N_KEYS >= 1) key_put(k, v[S::key(0).node].lo, S::key(0).width);
    if constexpr (S::N_KEYS >= 2) key_put(k, v[S::key(1).node].lo, S::key(1).width);
    if constexpr (S::N_KEYS >= 3) key_put(k, v[S::key(2).node].lo, S::key(2).width);
    if constexpr (S::N_KEYS >= 4) key_put(k, v[S::key(3).node].lo, S::key(3).width);
    return k;
}

struct StaticCtr {
    unsigned long long* rows_selected;
    unsigned long long* deferred;
};

template <class S>
__global__ void __launch_bounds__(S::THREADS, 1)
        k_scan_agg_static(const __grid_constant__ DProg prog, const __grid_constant__ StaticArgs sa, const TableView tv, const int64_t base,
                          const int64_t n_tiles, uint32_t* __restrict__ defer_out, StaticCtr ctr) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t full_bar[ST_STAGES];
    __shared__ uint64_t dir_lo[S::GMAX > 0 ? S::GMAX : 1], dir_hi[S::GMAX > 0 ? S::GMAX : 1];
    __shared__ uint8_t* dir_slot[S::GMAX > 0 ? S::GMAX : 1];
    __shared__ int dir_cnt, dir_lock;
    const DProg& p = prog;
    using NodeSeq = std::make_integer_sequence<int, S::N_NODES>;
    using AggSeq = std::make_integer_sequence<int, S::N_AGGS>;

    // column c occupies [col_off[c], col_off[c] + TILE*width) inside a stage (TILE*width keeps 16-byte alignment)
    const int stage_bytes = sa.stage_bytes;
    const int tid = threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < ST_STAGES; s++) mbar_init(&full_bar[s], 1);
        dir_cnt = 0;
        dir_lock = 0;
        fence_barrier_init();
    }
    __syncthreads();

    auto issue = [&](int64_t tile, int s) { // one elected thread: one bulk copy per column, all completing on full_bar[s]
        mbar_expect_tx(&full_bar[s], (uint32_t)stage_bytes);
        uint8_t* dst = smem + (size_t)s * stage_bytes;
        for (int c = 0; c < p.n_cols; c++) {
            const int w = p.cols[c].width;
            tma_load_1d(dst + sa.col_off[c], (const uint8_t*)p.cols[c].data + (size_t)(base + tile * ST_TILE) * w, (uint32_t)(ST_TILE * w),
                        &full_bar[s]);
        }
    };
    const int64_t first = blockIdx.x, step = gridDim.x;
    if (tid == 0)
        for (int s = 0; s < ST_STAGES; s++)
            if (first + s * step < n_tiles) issue(first + s * step, s);

    Bank<S> bank[S::GMAX > 0 ? S::GMAX : 1];
#pragma unroll
    for (int g = 0; g < (S::GMAX > 0 ? S::GMAX : 1); g++) bank_clear<S>(bank[g], AggSeq{});
    unsigned long long selected = 0, deferred = 0;

    int it = 0;
    for (int64_t tile = first; tile < n_tiles; tile += step, it++) {
        const int s = it % ST_STAGES;
        mbar_wait(&full_bar[s], (uint32_t)((it / ST_STAGES) & 1));
        const SmemCols sc{smem + (size_t)s * stage_bytes, sa.col_off};
        for (int r = tid; r < ST_TILE; r += S::THREADS) {
            const bool pass = eval_preds_smem(p, sc, r);
            selected += pass ? 1 : 0;
            if (!__any_sync(__activemask(), pass)) continue;
            Val v[S::N_NODES];
            eval_static_all<S>(p, v, sc, r, NodeSeq{});
            if constexpr (S::N_KEYS == 0) {
                bank_update_all<S>(bank[0], v, pass, AggSeq{});
            } else {
                const PackedKey k = pack_key_static<S>(v);
                int gid = -1;
                if constexpr (S::GMAX > 0) {
                    const int cnt = *(volatile int*)&dir_cnt;
#pragma unroll
                    for (int g = 0; g < S::GMAX; g++)
                        if (g < cnt && ((volatile uint64_t*)dir_lo)[g] == k.lo && (p.key_words == 1 || ((volatile uint64_t*)dir_hi)[g] == k.hi)) gid = g;
#pragma unroll
                    for (int g = 0; g < S::GMAX; g++) bank_update_all<S>(bank[g], v, pass && gid == g, AggSeq{});
                }
                if (pass && gid < 0) {
                    // global-table path: the group is not (yet) in this CTA's directory
                    uint8_t* slot = table_find_or_insert(tv, p, k.lo, k.hi, hash_packed_key_static<static_key_bytes<S>()>(k, p.phase_seed));
                    if (!slot) {
                        const int64_t i = tile * ST_TILE + r;
                        atomicOr(&defer_out[i >> 5], 1u << (i & 31));
                        deferred++;
                    } else {
                        row_update_global<S>(p, slot, v, AggSeq{});
                        if constexpr (S::GMAX > 0) {
                            // best-effort publication into the directory (try-lock; duplicates are excluded under the lock)
                            if (*(volatile int*)&dir_cnt < S::GMAX && atomicCAS(&dir_lock, 0, 1) == 0) {
                                const int c = *(volatile int*)&dir_cnt;
                                bool present = false;
                                for (int g = 0; g < c; g++)
                                    present = present || (((volatile uint64_t*)dir_lo)[g] == k.lo && ((volatile uint64_t*)dir_hi)[g] == k.hi);
                                if (!present && c < S::GMAX) {
                                    ((volatile uint64_t*)dir_lo)[c] = k.lo;
                                    ((volatile uint64_t*)dir_hi)[c] = k.hi;
                                    ((uint8_t* volatile*)dir_slot)[c] = slot;
                                    __threadfence_block();
                                    *(volatile int*)&dir_cnt = c + 1;
                                }
                                __threadfence_block();
                                atomicExch(&dir_lock, 0);
                            }
                        }
                    }
                }
            }
        }
        __syncthreads(); // every thread is done with stage s
        if (tid == 0 && tile + (int64_t)ST_STAGES * step < n_tiles) {
            fence_proxy_async();
            issue(tile + (int64_t)ST_STAGES * step, s);
        }
    }
    __syncthreads();
    // retire: warp-reduce every bank and fold it into the table
    if constexpr (S::N_KEYS == 0) {
        bank_flush<S>(p, bank[0], tv.slots, AggSeq{});
    } else if constexpr (S::GMAX > 0) {
        const int cnt = dir_cnt;
#pragma unroll
        for (int g = 0; g < S::GMAX; g++)
            if (g < cnt) bank_flush<S>(p, bank[g], dir_slot[g], AggSeq{});
    }
    selected = warp_sum_u64(selected);
    deferred = warp_sum_u64(deferred);
    if ((tid & 31) == 0) {
        if (selected) atomicAdd(ctr.rows_selected, selected);
        if (deferred) atomicAdd(ctr.deferred, deferred);
    }
}

// ---- registry ---------------------------------------------------------------------------------------------------------------
inline int match_static_shape(const DProg& p) {
    if (shape_matches<ShapeSumMulFast>(p)) return 0;
    if (shape_matches<ShapeQ1Dict>(p)) return 1;
    if (shape_matches<ShapeHighCard>(p)) return 2;
    return -1;
}
inline const char* static_shape_name(int id) {
    switch (id) {
    case 0: return ShapeSumMulFast::name;
    case 1: return ShapeQ1Dict::name;
    case 2: return ShapeHighCard::name;
    }
    return "generic";
}

template <class S>
inline int64_t launch_shape(const DProg& p, const TableView& tv, int64_t base, int64_t n, uint32_t* defer_out, unsigned long long* rows_selected,
                            unsigned long long* deferred, int n_sms, cudaStream_t stream) {
    int row_bytes = 0;
    StaticArgs sa{};
    for (int c = 0; c < p.n_cols; c++) {
        sa.col_off[c] = ST_TILE * row_bytes;
        row_bytes += p.cols[c].width;
        // TMA bulk copies need 16-byte aligned sources: tile starts are multiples of 1024 rows, so the base pointers decide
        if (p.cols[c].nulls || (((uintptr_t)p.cols[c].data + (size_t)base * p.cols[c].width) & 15)) return 0;
    }
    const int64_t n_tiles = n / ST_TILE;
    if (n_tiles == 0 || row_bytes == 0) return 0;
    sa.stage_bytes = ST_TILE * row_bytes;
    const size_t smem = (size_t)ST_STAGES * ST_TILE * row_bytes;
    if (smem > 200 * 1024) return 0;
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(k_scan_agg_static<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        configured = true;
    }
    const int grid = (int)(n_tiles < n_sms ? n_tiles : n_sms);
    StaticCtr ctr{rows_selected, deferred};
    k_scan_agg_static<S><<<grid, S::THREADS, smem, stream>>>(p, sa, tv, base, n_tiles, defer_out, ctr);
    return n_tiles * ST_TILE;
}

// Launches the static kernel of `shape` over the leading whole tiles of [base, base+n); returns the number of rows it
// covered (0 when the preconditions — no null maps, 16-byte aligned columns, at least one tile — do not hold). The caller
// runs the generic kernel over the remaining rows.
inline int64_t launch_static_shape(int shape, const DProg& p, const TableView& tv, int64_t base, int64_t n, uint32_t* defer_out,
                                   unsigned long long* rows_selected, unsigned long long* deferred, int n_sms, cudaStream_t stream) {
    switch (shape) {
    case 0: return launch_shape<ShapeSumMulFast>(p, tv, base, n, defer_out, rows_selected, deferred, n_sms, stream);
    case 1: return launch_shape<ShapeQ1Dict>(p, tv, base, n, defer_out, rows_selected, deferred, n_sms, stream);
    case 2: return launch_shape<ShapeHighCard>(p, tv, base, n, defer_out, rows_selected, deferred, n_sms, stream);
    }
    return 0;
}

} // namespace srb
