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
    static constexpr int N_NODES = 5, N_AGGS = 1, N_KEYS = 0, GMAX = 1, THREADS = 1024;
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
    static constexpr int N_NODES = 14, N_AGGS = 8, N_KEYS = 2, GMAX = 4, THREADS = 512;
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
constexpr int ST_TILE = 256;       // rows per tile (one consumer group of 128 threads takes 2 rows per thread)
constexpr int ST_MAX_STAGES = 4;   // ring depth (the host picks the deepest ring that fits beside the accumulators)
constexpr int ST_SMEM_BUDGET = 220 * 1024;

// predicate lowered for the tile layout: column offset inside a stage, width, and a 3-bit truth table over
// {v > lit, v == lit, v < lit} (bit 0, 1, 2) so that one compare pair serves all six operators without branching
struct SPred {
    int32_t col_off;
    int32_t width;
    uint32_t opmask;
    int32_t group;
    uint64_t lit_lo;
    int64_t lit_hi;
};
struct StaticArgs { // host-computed tile layout: column c sits at col_off[c] inside a stage of stage_bytes
    int32_t col_off[MAX_COLS];
    int32_t stage_bytes;
    int32_t n_stages;
    int32_t n_preds;
    int32_t acc_off; // byte offset of the per-thread accumulators inside dynamic shared memory
    SPred preds[MAX_PREDS];
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

// one predicate on one row of the tile (NULL-free columns): NULL never satisfies a compare, IS_NULL is constant false
SRB_D bool spred_eval(const SPred& pr, const uint8_t* stage, int r) {
    const uint8_t* col = stage + pr.col_off;
    int idx;
    if (pr.width == 16) {
        ulonglong2 x = ((const ulonglong2*)col)[r];
        i128 v = make_i128(x.x, (int64_t)x.y), lit = make_i128(pr.lit_lo, pr.lit_hi);
        idx = v < lit ? 2 : (v == lit ? 1 : 0);
    } else {
        int64_t v;
        if (pr.width == 8)
            v = ((const int64_t*)col)[r];
        else if (pr.width == 4)
            v = ((const int32_t*)col)[r];
        else if (pr.width == 2)
            v = ((const int16_t*)col)[r];
        else
            v = ((const int8_t*)col)[r];
        const int64_t lit = (int64_t)pr.lit_lo;
        idx = v < lit ? 2 : (v == lit ? 1 : 0);
    }
    return (pr.opmask >> idx) & 1u;
}
// AND of conjuncts and OR-groups over R rows at once (predicate loop outside, rows inside: uniform control flow is amortised)
template <int R>
SRB_D void eval_preds_rows(const StaticArgs& sa, const uint8_t* stage, const int (&rows)[R], bool (&pass)[R]) {
#pragma unroll
    for (int j = 0; j < R; j++) pass[j] = true;
    int i = 0;
    const int n = sa.n_preds;
    while (i < n) {
        const int g = sa.preds[i].group;
        bool any[R];
#pragma unroll
        for (int j = 0; j < R; j++) any[j] = false;
        do {
            const SPred& pr = sa.preds[i];
#pragma unroll
            for (int j = 0; j < R; j++) any[j] = any[j] || spred_eval(pr, stage, rows[j]);
            i++;
        } while (g != 0 && i < n && sa.preds[i].group == g);
#pragma unroll
        for (int j = 0; j < R; j++) pass[j] = pass[j] && any[j];
    }
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
// dense index of a representative among the representatives (its accumulator plane), and their number
template <class S>
constexpr int rep_index(int a) {
    int n = 0;
    for (int b = 0; b < a; b++) {
        const SAgg h = S::agg(b);
        if ((h.fn == FN_SUM || h.fn == FN_AVG) && agg_rep<S>(b) == b) n++;
    }
    return n;
}
template <class S>
constexpr int n_reps() {
    return rep_index<S>(S::N_AGGS);
}
// can the value of node n be NULL? (static shapes bind NOT NULL columns, so only checked ops produce NULLs)
template <class S>
constexpr bool node_can_be_null(int n) {
    if (n < 0) return false;
    const SNode nd = S::node(n);
    if (nd.checked) return true;
    return node_can_be_null<S>(nd.lhs) || node_can_be_null<S>(nd.rhs);
}

// Per-thread accumulators of the low-cardinality path. The LOW 64 bits of every partial sum live in shared memory, one
// private word per (representative, group, thread): lanes of a warp hit distinct banks whatever their groups are, so the
// update is a plain LDS / add / STS with the group as a runtime address — no atomics and no per-group predication. The
// HIGH part (carries + sign words, almost always zero per row) stays in registers and is only touched when a warp vote
// finds a non-zero delta.
template <class S>
struct HiRegs {
    static constexpr int G = S::GMAX > 0 ? S::GMAX : 1;
    static constexpr int NR = n_reps<S>() > 0 ? n_reps<S>() : 1;
    uint32_t w2[G][NR], w3[G][NR]; // bits 64..95 and 96..127 of each partial sum (w3 only for 16-byte arguments)
    uint32_t n[G][NR];             // non-NULL rows (only for arguments that can be NULL)
    uint32_t rows[G];
};

template <class S, int A, int THREADS>
SRB_D void acc_update(HiRegs<S>& h, uint8_t* acc_tid, const Val* v, bool on, int gid) {
    constexpr SAgg g = S::agg(A);
    constexpr int G = HiRegs<S>::G;
    if constexpr (g.fn == FN_COUNT) {
        // count(*) and count of a NOT NULL argument are the group's row count
    } else if constexpr (agg_rep<S>(A) == A) {
        constexpr int RI = rep_index<S>(A);
        constexpr bool wide = S::node(g.arg).width == 16;
        constexpr bool nullable = node_can_be_null<S>(g.arg);
        const Val& x = v[g.arg];
        bool take = on;
        if constexpr (nullable) take = take && !x.null;
        uint32_t d2 = 0, d3 = 0;
        if (take) {
            uint64_t* slot = (uint64_t*)(acc_tid + (size_t)(RI * G + gid) * THREADS * 8);
            const uint64_t old = *slot, nw = old + x.lo;
            *slot = nw;
            const uint64_t dh = (uint64_t)x.hi + (nw < old ? 1ULL : 0ULL);
            d2 = (uint32_t)dh;
            d3 = (uint32_t)(dh >> 32);
        }
        const bool any_hi = wide ? ((d2 | d3) != 0) : (d2 != 0);
        if (__any_sync(0xffffffffu, any_hi)) {
#pragma unroll
            for (int q = 0; q < G; q++) {
                const bool mine = gid == q;
                if constexpr (wide) {
                    const uint32_t a2 = mine ? d2 : 0u, a3 = mine ? d3 : 0u;
                    const uint32_t s2 = h.w2[q][RI] + a2;
                    h.w3[q][RI] += a3 + (s2 < a2 ? 1u : 0u);
                    h.w2[q][RI] = s2;
                } else {
                    h.w2[q][RI] += mine ? d2 : 0u;
                }
            }
        }
        if constexpr (nullable) {
#pragma unroll
            for (int q = 0; q < G; q++) h.n[q][RI] += (take && gid == q) ? 1u : 0u;
        }
    }
}
template <class S, int THREADS, int... A>
SRB_D void acc_update_all(HiRegs<S>& h, uint8_t* acc_tid, const Val* v, bool on, int gid, std::integer_sequence<int, A...>) {
    (acc_update<S, A, THREADS>(h, acc_tid, v, on, gid), ...);
#pragma unroll
    for (int q = 0; q < HiRegs<S>::G; q++) h.rows[q] += (on && gid == q) ? 1u : 0u;
}

SRB_D uint64_t warp_sum_u64(uint64_t x) {
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    return x;
}
SRB_D void warp_sum_u128(uint64_t& lo, uint64_t& hi) {
    for (int o = 16; o > 0; o >>= 1) {
        uint64_t olo = __shfl_down_sync(0xffffffffu, lo, o), ohi = __shfl_down_sync(0xffffffffu, hi, o);
        uint64_t nlo = lo + olo;
        hi += ohi + (nlo < olo ? 1ULL : 0ULL);
        lo = nlo;
    }
}

// fold group q of this thread's accumulators into table slot `slot`: warp reduction, then lane 0 issues the atomics
template <class S, int THREADS>
struct Flusher {
    static constexpr int G = HiRegs<S>::G;
    static constexpr int NR = HiRegs<S>::NR;
    uint64_t red_lo[NR], red_hi[NR], red_n[NR];

    template <int A>
    __device__ __forceinline__ void reduce_rep(const HiRegs<S>& h, const uint8_t* acc_tid, int q) {
        constexpr SAgg g = S::agg(A);
        if constexpr ((g.fn == FN_SUM || g.fn == FN_AVG) && agg_rep<S>(A) == A) {
            constexpr int RI = rep_index<S>(A);
            constexpr bool wide = S::node(g.arg).width == 16;
            uint64_t lo = *(const uint64_t*)(acc_tid + (size_t)(RI * G + q) * THREADS * 8);
            // static q is needed for the register arrays: select by unrolled compare
            uint32_t w2 = 0, w3 = 0, n = 0;
#pragma unroll
            for (int k = 0; k < G; k++)
                if (k == q) {
                    w2 = h.w2[k][RI];
                    w3 = h.w3[k][RI];
                    n = h.n[k][RI];
                }
            uint64_t hi = wide ? (((uint64_t)w3 << 32) | w2) : (uint64_t)(int64_t)(int32_t)w2;
            warp_sum_u128(lo, hi);
            red_lo[RI] = lo;
            red_hi[RI] = hi;
            if constexpr (node_can_be_null<S>(g.arg)) red_n[RI] = warp_sum_u64(n);
        }
    }
    template <int A>
    __device__ __forceinline__ void emit(const DProg& p, uint8_t* slot, uint64_t rows_total) {
        constexpr SAgg g = S::agg(A);
        const DAgg& rt = p.aggs[A];
        const int lane = threadIdx.x & 31;
        if constexpr (g.fn == FN_COUNT) {
            if (lane == 0 && rows_total) atomicAdd((unsigned long long*)(slot + rt.off_n), (unsigned long long)rows_total);
        } else {
            constexpr int RI = rep_index<S>(agg_rep<S>(A));
            uint64_t n = rows_total;
            if constexpr (node_can_be_null<S>(g.arg)) n = red_n[RI];
            if (lane == 0 && n) {
                atomic_add128(slot, rt.off_lo, rt.off_hi, red_lo[RI], red_hi[RI]);
                if (rt.off_n >= 0) atomicAdd((unsigned long long*)(slot + rt.off_n), (unsigned long long)n);
            }
        }
    }
    template <int... A>
    __device__ __forceinline__ void run(const DProg& p, const HiRegs<S>& h, const uint8_t* acc_tid, int q, uint8_t* slot,
                                        std::integer_sequence<int, A...>) {
        uint32_t rows = 0;
#pragma unroll
        for (int k = 0; k < G; k++)
            if (k == q) rows = h.rows[k];
        const uint64_t rows_total = warp_sum_u64(rows);
        (reduce_rep<A>(h, acc_tid, q), ...);
        (emit<A>(p, slot, rows_total), ...);
    }
};

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
    unsigned long long* inserted;
};

constexpr int ST_GROUP = 128;                  // threads per consumer group
constexpr int ST_MAX_GROUPS = 8;
SRB_D void group_barrier(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(ST_GROUP) : "memory"); }

// The CTA is split into consumer groups of 128 threads. Every group streams its OWN tiles through its OWN ring of stages and
// synchronises only with itself (named barrier), so the groups drift apart and one group's barrier / latency bubbles are
// filled by the others — the block never stalls as a whole, and there is no block-wide barrier in the steady state.
template <class S>
__global__ void __launch_bounds__(S::THREADS, 1)
        k_scan_agg_static(const __grid_constant__ DProg prog, const __grid_constant__ StaticArgs sa, const TableView tv, const int64_t base,
                          const int64_t n_tiles, uint32_t* __restrict__ defer_out, StaticCtr ctr) {
    constexpr int THREADS = S::THREADS;
    constexpr int NG = THREADS / ST_GROUP;
    constexpr int R = ST_TILE / ST_GROUP; // rows per thread per tile, strided by the group size (conflict-free LDS)
    constexpr int G = HiRegs<S>::G;
    static_assert(THREADS % ST_GROUP == 0 && NG <= ST_MAX_GROUPS && ST_TILE % ST_GROUP == 0, "block / tile geometry");
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t full_bar[ST_MAX_GROUPS][ST_MAX_STAGES];
    __shared__ uint64_t dir_lo[G], dir_hi[G];
    __shared__ uint8_t* dir_slot[G];
    __shared__ int dir_cnt, dir_lock;
    const DProg& p = prog;
    using NodeSeq = std::make_integer_sequence<int, S::N_NODES>;
    using AggSeq = std::make_integer_sequence<int, S::N_AGGS>;

    const int stage_bytes = sa.stage_bytes, n_stages = sa.n_stages;
    const int tid = threadIdx.x, grp = tid / ST_GROUP, gtid = tid % ST_GROUP;
    if (tid == 0) {
        for (int g = 0; g < NG; g++)
            for (int s = 0; s < n_stages; s++) mbar_init(&full_bar[g][s], 1);
        dir_cnt = 0;
        dir_lock = 0;
        fence_barrier_init();
    }
    // this thread's private accumulator words start zeroed
    uint8_t* acc_tid = smem + sa.acc_off + tid * 8;
    if constexpr (S::GMAX > 0) {
#pragma unroll
        for (int k = 0; k < HiRegs<S>::NR * G; k++) *(uint64_t*)(acc_tid + (size_t)k * THREADS * 8) = 0;
    }
    __syncthreads();

    uint8_t* ring = smem + (size_t)grp * n_stages * stage_bytes; // this group's stages
    auto issue = [&](int64_t tile, int s) { // the group's elected thread: one bulk copy per column, all completing on full_bar[grp][s]
        mbar_expect_tx(&full_bar[grp][s], (uint32_t)stage_bytes);
        uint8_t* dst = ring + (size_t)s * stage_bytes;
        for (int c = 0; c < p.n_cols; c++) {
            const int w = p.cols[c].width;
            tma_load_1d(dst + sa.col_off[c], (const uint8_t*)p.cols[c].data + (size_t)(base + tile * ST_TILE) * w, (uint32_t)(ST_TILE * w),
                        &full_bar[grp][s]);
        }
    };
    const int64_t first = (int64_t)blockIdx.x * NG + grp, step = (int64_t)gridDim.x * NG;
    if (gtid == 0)
        for (int s = 0; s < n_stages; s++)
            if (first + s * step < n_tiles) issue(first + s * step, s);

    HiRegs<S> h;
#pragma unroll
    for (int q = 0; q < G; q++) {
        h.rows[q] = 0;
#pragma unroll
        for (int k = 0; k < HiRegs<S>::NR; k++) h.w2[q][k] = h.w3[q][k] = h.n[q][k] = 0;
    }
    unsigned long long selected = 0, deferred = 0, inserted = 0;

    int s = 0;
    uint32_t parity = 0;
    for (int64_t tile = first; tile < n_tiles; tile += step) {
        mbar_wait(&full_bar[grp][s], parity);
        const uint8_t* stage = ring + (size_t)s * stage_bytes;
        const SmemCols sc{stage, sa.col_off};
        int rows[R];
        bool pass[R];
#pragma unroll
        for (int j = 0; j < R; j++) rows[j] = gtid + j * ST_GROUP;
        eval_preds_rows<R>(sa, stage, rows, pass);
#pragma unroll
        for (int j = 0; j < R; j++) {
            selected += pass[j] ? 1 : 0;
            if (!__any_sync(0xffffffffu, pass[j])) continue;
            const int r = rows[j];
            Val v[S::N_NODES];
            eval_static_all<S>(p, v, sc, r, NodeSeq{});
            if constexpr (S::N_KEYS == 0) {
                acc_update_all<S, THREADS>(h, acc_tid, v, pass[j], 0, AggSeq{});
            } else {
                const PackedKey k = pack_key_static<S>(v);
                int gid = -1;
                if constexpr (S::GMAX > 0) {
                    const int cnt = *(volatile int*)&dir_cnt;
#pragma unroll
                    for (int q = 0; q < G; q++)
                        if (q < cnt && ((volatile uint64_t*)dir_lo)[q] == k.lo && (p.key_words == 1 || ((volatile uint64_t*)dir_hi)[q] == k.hi)) gid = q;
                    acc_update_all<S, THREADS>(h, acc_tid, v, pass[j] && gid >= 0, gid < 0 ? 0 : gid, AggSeq{});
                }
                if (pass[j] && gid < 0) {
                    // global-table path: the group is not (yet) in this CTA's directory
                    bool fresh;
                    uint8_t* slot = table_find_or_insert(tv, k.lo, k.hi, hash_packed_key_static<static_key_bytes<S>()>(k, p.phase_seed), &fresh);
                    inserted += fresh ? 1 : 0;
                    if (!slot) {
                        const int64_t i = tile * ST_TILE + r;
                        atomicOr(&defer_out[i >> 5], 1u << (i & 31));
                        deferred++;
                    } else {
                        row_update_global<S>(p, slot, v, AggSeq{});
                        if constexpr (S::GMAX > 0) {
                            // best-effort publication into the directory (try-lock; duplicates are excluded under the lock)
                            if (*(volatile int*)&dir_cnt < G && atomicCAS(&dir_lock, 0, 1) == 0) {
                                const int c = *(volatile int*)&dir_cnt;
                                bool present = false;
                                for (int q = 0; q < c; q++)
                                    present = present || (((volatile uint64_t*)dir_lo)[q] == k.lo && ((volatile uint64_t*)dir_hi)[q] == k.hi);
                                if (!present && c < G) {
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
        group_barrier(1 + grp); // every thread of the group is done with stage s
        if (gtid == 0 && tile + (int64_t)n_stages * step < n_tiles) {
            fence_proxy_async();
            issue(tile + (int64_t)n_stages * step, s);
        }
        if (++s == n_stages) {
            s = 0;
            parity ^= 1u;
        }
    }
    __syncthreads(); // the directory is final
    // retire: warp-reduce every group's accumulators and fold them into the table
    if constexpr (S::GMAX > 0) {
        Flusher<S, THREADS> f;
        if constexpr (S::N_KEYS == 0) {
            f.run(p, h, acc_tid, 0, tv.slots, AggSeq{});
        } else {
            const int cnt = dir_cnt;
            for (int q = 0; q < cnt; q++) f.run(p, h, acc_tid, q, dir_slot[q], AggSeq{});
        }
    }
    selected = warp_sum_u64(selected);
    deferred = warp_sum_u64(deferred);
    inserted = warp_sum_u64(inserted);
    if ((tid & 31) == 0) {
        if (selected) atomicAdd(ctr.rows_selected, selected);
        if (deferred) atomicAdd(ctr.deferred, deferred);
        if (inserted) atomicAdd(ctr.inserted, inserted);
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

inline uint32_t pred_opmask(int op) {
    switch (op) {
    case 1: return 0b010; // EQ
    case 2: return 0b101; // NE
    case 3: return 0b001; // GT
    case 4: return 0b011; // GE
    case 5: return 0b100; // LT
    case 6: return 0b110; // LE
    case 10: return 0b111; // NOT NULL on a NULL-free column
    default: return 0;     // IS NULL
    }
}

template <class S>
inline int64_t launch_shape(const DProg& p, const TableView& tv, int64_t base, int64_t n, uint32_t* defer_out, unsigned long long* rows_selected,
                            unsigned long long* deferred, unsigned long long* inserted, int n_sms, cudaStream_t stream) {
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
    const int acc_bytes = S::GMAX > 0 ? HiRegs<S>::NR * HiRegs<S>::G * S::THREADS * 8 : 0;
    constexpr int NG = S::THREADS / ST_GROUP;
    int stages = (ST_SMEM_BUDGET - acc_bytes) / (sa.stage_bytes * NG);
    if (stages > ST_MAX_STAGES) stages = ST_MAX_STAGES;
    if (stages < 2) return 0;
    sa.n_stages = stages;
    sa.acc_off = NG * stages * sa.stage_bytes;
    sa.n_preds = p.n_preds;
    for (int i = 0; i < p.n_preds; i++) {
        const DPred& d = p.preds[i];
        sa.preds[i] = SPred{sa.col_off[d.col], p.cols[d.col].width, pred_opmask(d.op), d.group, d.lit_lo, d.lit_hi};
    }
    const size_t smem = (size_t)sa.acc_off + acc_bytes;
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(k_scan_agg_static<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST_SMEM_BUDGET);
        configured = true;
    }
    const int64_t want = (n_tiles + NG - 1) / NG;
    const int grid = (int)(want < n_sms ? want : n_sms);
    StaticCtr ctr{rows_selected, deferred, inserted};
    k_scan_agg_static<S><<<grid, S::THREADS, smem, stream>>>(p, sa, tv, base, n_tiles, defer_out, ctr);
    return n_tiles * ST_TILE;
}

// Launches the static kernel of `shape` over the leading whole tiles of [base, base+n); returns the number of rows it
// covered (0 when the preconditions — no null maps, 16-byte aligned columns, at least one tile — do not hold). The caller
// runs the generic kernel over the remaining rows.
inline int64_t launch_static_shape(int shape, const DProg& p, const TableView& tv, int64_t base, int64_t n, uint32_t* defer_out,
                                   unsigned long long* rows_selected, unsigned long long* deferred, unsigned long long* inserted, int n_sms,
                                   cudaStream_t stream) {
    switch (shape) {
    case 0: return launch_shape<ShapeSumMulFast>(p, tv, base, n, defer_out, rows_selected, deferred, inserted, n_sms, stream);
    case 1: return launch_shape<ShapeQ1Dict>(p, tv, base, n, defer_out, rows_selected, deferred, inserted, n_sms, stream);
    case 2: return launch_shape<ShapeHighCard>(p, tv, base, n, defer_out, rows_selected, deferred, inserted, n_sms, stream);
    }
    return 0;
}

} // namespace srb
