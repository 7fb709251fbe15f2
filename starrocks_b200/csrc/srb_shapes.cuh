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
#include "srb_rowpipe.cuh"

namespace srb {

// compile-time index packs (what std::integer_sequence gives; spelled out so that NVRTC needs no host headers)
template <int... I>
struct iseq {};
template <int N, int... I>
struct make_iseq_impl : make_iseq_impl<N - 1, N - 1, I...> {};
template <int... I>
struct make_iseq_impl<0, I...> {
    typedef iseq<I...> type;
};
template <int N>
using make_iseq = typename make_iseq_impl<N>::type;

struct Counters;

struct SNode {
    int op, width, lhs, rhs, checked, adj_l, adj_r;
    int nullable = 0; // LOAD: the column arrives with a null map (one byte per row in the stage)
};
struct SAgg {
    int fn, arg, wide, has_n; // wide: state keeps the high word (off_hi >= 0); has_n: state keeps its own count
    int merge = 0;            // phase-2 aggregate over an intermediate column (COUNT adds the partial count)
    int dbl = 0;              // 1: the argument is a DOUBLE (the state holds a double), 2: an integer argument accumulated in double (AVG over plain
                              // integers, avg.h:49-53)
};
struct SKey {
    int node, width;
    int nullable = 0; // the key area carries a leading null byte for this column (DKey::nullable)
};

// ---- the registered shapes ---------------------------------------------------------------------------------------
// TPC-H Q6 pipeline (fe/fe-core/src/test/resources/sql/external/hive/tpch/q6.sql:8,24):
//   sum(cast(l_extendedprice as D128) * cast(l_discount as D128)), no GROUP BY
struct ShapeSumMulFast {
    static constexpr const char* name = "sum(d64*d64) no-group";
    static constexpr int N_NODES = 5, N_AGGS = 1, N_KEYS = 0, GMAX = 1, THREADS = 512, ROWS = 4;
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
    static constexpr int N_NODES = 14, N_AGGS = 8, N_KEYS = 2, GMAX = 4, THREADS = 512, ROWS = 2; // 256 x 4 measured slower (IPC 1.35 vs 2.1)
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
    static constexpr int N_NODES = 2, N_AGGS = 2, N_KEYS = 1, GMAX = 0, THREADS = 992, ROWS = 2; // + the producer warp = 1024
    static constexpr SNode node(int) { return SNode{DOP_LOAD, 8, -1, -1, 0, 0, 0}; }
    static constexpr SAgg agg(int i) { return i == 0 ? SAgg{FN_SUM, 1, 0, 0} : SAgg{FN_COUNT, -1, 0, 1}; }
    static constexpr SKey key(int) { return SKey{0, 8}; }
};

// TPC-H Q3's aggregation over the join output (typed plan: fe/fe-core/src/test/resources/sql/external/hive/tpch/q3.sql:34-36):
//   group by l_orderkey, o_orderdate, o_shippriority (INT, DATE, INT: a 12-byte packed key); sum(l_extendedprice * (1 - l_discount))
// ~14 M groups at SF100: every row probes the HBM table (GMAX = 0), like the high-cardinality shape, but the expression and the
// three-column key are compiled in and the five columns arrive through the TMA ring.
struct ShapeQ3Agg {
    static constexpr const char* name = "q3 (3 int keys, sum(d64 * (1 - d64)))";
    static constexpr int N_NODES = 10, N_AGGS = 1, N_KEYS = 3, GMAX = 0, THREADS = 992, ROWS = 2;
    static constexpr SNode node(int i) {
        constexpr SNode t[N_NODES] = {{DOP_LOAD, 4, -1, -1, 0, 0, 0},   // 0 l_orderkey
                                      {DOP_LOAD, 4, -1, -1, 0, 0, 0},   // 1 o_orderdate
                                      {DOP_LOAD, 4, -1, -1, 0, 0, 0},   // 2 o_shippriority
                                      {DOP_LOAD, 8, -1, -1, 0, 0, 0},   // 3 l_extendedprice
                                      {DOP_LOAD, 8, -1, -1, 0, 0, 0},   // 4 l_discount
                                      {DOP_LIT, 8, -1, -1, 0, 0, 0},    // 5 literal 1 as DECIMAL64(16,2)
                                      {DOP_SUB, 8, 5, 4, 1, 0, 0},      // 6 1 - l_discount
                                      {DOP_CAST, 16, 3, -1, 1, 0, 0},   // 7
                                      {DOP_CAST, 16, 6, -1, 1, 0, 0},   // 8
                                      {DOP_MULFAST, 16, 3, 6, 0, 0, 0}}; // 9 ext * (1 - disc)   DECIMAL128(31,4)
        return t[i];
    }
    static constexpr SAgg agg(int) { return SAgg{FN_SUM, 9, 1, 1}; }
    static constexpr SKey key(int i) { return SKey{i, 4}; }
};

// Phase 2 of the two-phase high-cardinality GROUP BY (aggregator.cpp:871-875): rows are (k, partial sum, partial count) tuples
// received from the shuffle; sum and count are merged into the final table.
struct ShapeHighCardMerge {
    static constexpr const char* name = "merge by bigint: sum(bigint), count";
    static constexpr int N_NODES = 3, N_AGGS = 2, N_KEYS = 1, GMAX = 0, THREADS = 992, ROWS = 2;
    static constexpr SNode node(int) { return SNode{DOP_LOAD, 8, -1, -1, 0, 0, 0}; }
    static constexpr SAgg agg(int i) { return i == 0 ? SAgg{FN_SUM, 1, 0, 1, 1} : SAgg{FN_COUNT, 2, 0, 1, 1}; }
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
        if (d.fn != s.fn || d.arg != s.arg || d.is_double || d.avg_as_double || (d.merge != 0) != (s.merge != 0)) return false;
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
SRB_D void mbar_wait_a(uint32_t bar, uint32_t parity) { // bar: shared-window address
    asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "WAIT_%=:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
            "@p bra DONE_%=;\n\t"
            "bra WAIT_%=;\n\t"
            "DONE_%=:\n\t}" ::"r"(bar),
            "r"(parity)
            : "memory");
}
SRB_D void mbar_arrive_a(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
SRB_D void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
SRB_D void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
SRB_D void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// ---- static row pipeline ---------------------------------------------------------------------------------------------------
constexpr int ST_MAX_STAGES = 6;   // ring depth (the host picks the deepest ring that fits beside the accumulators)
constexpr int ST_SMEM_BUDGET = 226 * 1024; // of the 227 KB a CTA may take; the kernel's static shared data is < 1 KB
constexpr int ST_MAX_RANGES = 6;   // conjunct predicates after lowering to per-column ranges (more: the generic predicate loop)

// rows of one tile: every consumer thread takes S::ROWS rows, strided by the thread count (conflict-free LDS)
template <class S>
constexpr int tile_rows() {
    return S::THREADS * S::ROWS;
}
// LOAD nodes own the leading columns of a stage, in node order: their offsets are compile-time constants, so a row's
// operands are fetched with LDS [row_base + immediate]
// A BinaryColumn (LOAD of width 0; group keys only) is staged as its u32 offsets — tile + 1 of them, padded to tile + 4 for the 16-byte
// granularity of bulk copies — and a window of its bytes (bin_window_bytes) whose position the producer decides per tile.
constexpr int bin_window_bytes(int tile) { return tile * 8 + 32; }
template <class S>
constexpr int load_bytes(int node) {
    return S::node(node).width == 0 ? 4 * tile_rows<S>() + 16 : S::node(node).width * tile_rows<S>();
}
template <class S>
constexpr int load_off(int node) {
    int off = 0;
    for (int j = 0; j < node; j++)
        if (S::node(j).op == DOP_LOAD) off += load_bytes<S>(j);
    return off;
}
template <class S>
constexpr bool shape_has_binary() {
    for (int i = 0; i < S::N_NODES; i++)
        if (S::node(i).op == DOP_LOAD && S::node(i).width == 0) return true;
    return false;
}
// where the bytes of the binary key columns of the current tile are: base[k] + (a value's offset in the column) is its first byte,
// whether the window was staged or (it did not fit) the bytes are read from the column itself
struct TileBin {
    const uint8_t* base[4];
};

// A conjunct (or several on one column) lowered to "lo <= v <= lo + span" (unsigned compare after the subtraction),
// optionally negated: one subtract and one compare per row whatever the operator was.
struct SRange {
    int32_t col_off;
    int32_t width; // 4 or 8
    int32_t neg;
    int32_t _pad;
    int64_t lo;
    uint64_t span;
};

// predicate lowered for the tile layout: column offset inside a stage, width, and a 3-bit truth table over
// {v > lit, v == lit, v < lit} (bit 0, 1, 2) so that one compare pair serves all six operators without branching
struct SPred {
    int32_t col_off;
    int32_t width;
    uint32_t opmask; // bit 3: the result for a NULL value (IS NULL)
    int32_t group;
    int32_t null_off; // the column's null bytes inside a stage, -1: NULL-free
    int32_t _pad;
    uint64_t lit_lo;
    int64_t lit_hi;
};
struct StaticArgs { // host-computed tile layout: column c sits at col_off[c] inside a stage of stage_bytes
    int32_t col_off[MAX_COLS];
    int32_t null_off[MAX_COLS];       // its null map (one byte per row), -1: none
    int32_t node_null_off[MAX_NODES]; // the same offset by LOAD node (a compile-time index in the kernel)
    int32_t bin_win_off[MAX_COLS];    // binary column: its byte window inside a stage, -1: not binary
    int32_t bin_hdr_off[MAX_COLS];    // and the 16-byte header the producer writes per tile: {u32 window start in the column, u32 overflow}
    int32_t stage_bytes;
    int32_t n_stages;
    int32_t n_preds;
    int32_t acc_off; // byte offset of the per-thread accumulators inside dynamic shared memory
    int32_t n_ranges; // >= 0: the predicates are the conjunction of rng[0..n_ranges); -1: evaluate preds[] generically
    int32_t narrow_ok; // the plan's literals and casts allow the narrow tile body (narrow_literals_ok)
    SRange rng[ST_MAX_RANGES];
    SPred preds[MAX_PREDS];
};

// the conjunction of the lowered ranges over R rows (range loop outside, rows inside; every field of a range is a
// fixed kernel-parameter address, so nothing is fetched through an index)
template <int R, int STRIDE, int I>
SRB_D void eval_ranges_from(const StaticArgs& sa, const uint8_t* stage, int r0, bool (&pass)[R]) {
    if constexpr (I < ST_MAX_RANGES) {
        if (I < sa.n_ranges) { // nested, not chained: a plan with one range leaves after one test
            const SRange& g = sa.rng[I];
            const uint8_t* col = stage + g.col_off;
            const bool neg = g.neg != 0;
            if (g.width == 8) {
#pragma unroll
                for (int j = 0; j < R; j++) {
                    const uint64_t d = (uint64_t)((const int64_t*)col)[r0 + j * STRIDE] - (uint64_t)g.lo;
                    pass[j] = pass[j] && ((d <= g.span) != neg);
                }
            } else {
                const uint32_t lo = (uint32_t)g.lo, span = (uint32_t)g.span;
#pragma unroll
                for (int j = 0; j < R; j++) {
                    const uint32_t d = (uint32_t)((const int32_t*)col)[r0 + j * STRIDE] - lo;
                    pass[j] = pass[j] && ((d <= span) != neg);
                }
            }
            eval_ranges_from<R, STRIDE, I + 1>(sa, stage, r0, pass);
        }
    }
}
template <int R, int STRIDE>
SRB_D void eval_ranges_rows(const StaticArgs& sa, const uint8_t* stage, int r0, bool (&pass)[R]) {
#pragma unroll
    for (int j = 0; j < R; j++) pass[j] = true;
    eval_ranges_from<R, STRIDE, 0>(sa, stage, r0, pass);
}

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
    if (pr.null_off >= 0 && stage[pr.null_off + r]) return (pr.opmask >> 3) & 1u; // NULL: only IS NULL holds
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

// the general checked 128-bit product, kept out of line: the hot loop only carries the call
struct Mul128Res {
    uint64_t lo;
    int64_t hi;
    int ov;
};
static __device__ __noinline__ Mul128Res mul_overflow128_cold(uint64_t xl, int64_t xh, uint64_t yl, int64_t yh) {
    i128 r;
    const bool ov = mul_overflow128(make_i128(xl, xh), make_i128(yl, yh), &r);
    return Mul128Res{lo64(r), hi64(r), ov ? 1 : 0};
}

// One node of the shape over the R rows a thread owns in a tile (row j = r0 + j * S::THREADS). Working node-by-node across
// the rows gives the scheduler R independent dependency chains and lets every warp vote cover all R rows at once.
// Narrow (<= 8 byte) checked ADD / SUB stay in 64-bit arithmetic; products first try the warp-uniform small-operand path
// (every lane, every row in [0, 2^32): money amounts times small factors — one 32x32->64 multiply is the exact product
// and no overflow flag can be set); everything else is the generic primitive of srb_dev.cuh, so results are the same bits
// as the interpreter's. Must be called by all 32 lanes (votes inside).
template <class S, int I, int R>
SRB_D void eval_static_node(const DProg& p, const StaticArgs& sa, Val (&v)[R][S::N_NODES], const uint8_t* stage, int r0) {
    constexpr SNode nd = S::node(I);
    const DNode& rt = p.nodes[I];
    if constexpr (nd.op == DOP_LOAD && nd.width == 0) { // a BinaryColumn value: (offset of its first byte, length)
#pragma unroll
        for (int j = 0; j < R; j++) {
            const uint32_t* o = (const uint32_t*)(stage + load_off<S>(I)) + r0 + j * S::THREADS;
            const uint32_t b = o[0], e = o[1];
            v[j][I].lo = b;
            v[j][I].hi = (int64_t)(e - b);
            v[j][I].null = false;
            if constexpr (nd.nullable != 0) v[j][I].null = stage[sa.node_null_off[I] + r0 + j * S::THREADS] != 0;
        }
    } else if constexpr (nd.op == DOP_LOAD) {
#pragma unroll
        for (int j = 0; j < R; j++) {
            v[j][I] = lds_val<nd.width>(stage + load_off<S>(I), r0 + j * S::THREADS);
            if constexpr (nd.nullable != 0) v[j][I].null = stage[sa.node_null_off[I] + r0 + j * S::THREADS] != 0;
        }
    } else if constexpr (nd.op == DOP_LIT) {
#pragma unroll
        for (int j = 0; j < R; j++) {
            v[j][I].lo = rt.lit_lo;
            v[j][I].hi = rt.lit_hi;
            v[j][I].null = false;
        }
    } else if constexpr (nd.op == DOP_CAST) {
        const bool rescale = !(rt.lit_lo == 1 && rt.lit_hi == 0); // uniform: a cast between equal scales multiplies by 1
#pragma unroll
        for (int j = 0; j < R; j++) {
            const Val a = v[j][nd.lhs];
            i128 x = make_i128(a.lo, a.hi), res = x;
            bool ov = false;
            if constexpr (nd.checked != 0) {
                if (rescale) {
                    if constexpr (nd.width == 16) {
                        const Mul128Res m = mul_overflow128_cold(a.lo, a.hi, rt.lit_lo, rt.lit_hi);
                        res = make_i128(m.lo, m.hi);
                        ov = m.ov != 0;
                    } else {
                        ov = checked_binop<nd.width>(6, x, make_i128(rt.lit_lo, rt.lit_hi), &res);
                    }
                }
                if constexpr (nd.width == 8) ov = ov || !fits<8>(x);
                if constexpr (nd.width == 4) ov = ov || !fits<4>(x);
            }
            v[j][I].lo = lo64(res);
            v[j][I].hi = hi64(res);
            v[j][I].null = a.null || ov;
        }
    } else if constexpr (nd.op == DOP_MULFAST) {
        bool small = true;
#pragma unroll
        for (int j = 0; j < R; j++) small = small && (((v[j][nd.lhs].lo | v[j][nd.rhs].lo) >> 32) == 0);
        const bool fast = __all_sync(0xffffffffu, small);
#pragma unroll
        for (int j = 0; j < R; j++) {
            const Val a = v[j][nd.lhs], b = v[j][nd.rhs];
            if (fast) {
                v[j][I].lo = (uint64_t)(uint32_t)a.lo * (uint64_t)(uint32_t)b.lo;
                v[j][I].hi = 0;
            } else {
                const i128 r = mul_i64_i64((int64_t)a.lo, (int64_t)b.lo);
                v[j][I].lo = lo64(r);
                v[j][I].hi = hi64(r);
            }
            v[j][I].null = a.null || b.null;
        }
    } else if constexpr (nd.checked != 0 && nd.width == 8 && nd.adj_l == 0 && nd.adj_r == 0 && (nd.op == DOP_ADD || nd.op == DOP_SUB)) {
#pragma unroll
        for (int j = 0; j < R; j++) {
            const Val a = v[j][nd.lhs], b = v[j][nd.rhs];
            const uint64_t x = a.lo, y = b.lo;
            const uint64_t s = nd.op == DOP_ADD ? x + y : x - y;
            // signed overflow: ADD when the operands agree in sign and the sum does not; SUB when they differ and the result left x's sign
            const uint64_t o = nd.op == DOP_ADD ? ((x ^ s) & (y ^ s)) : ((x ^ y) & (x ^ s));
            v[j][I].lo = s; // the value of a NULL is never read: consumers test .null first
            v[j][I].hi = (int64_t)s >> 63;
            v[j][I].null = a.null || b.null || ((int64_t)o < 0);
        }
    } else if constexpr (nd.checked != 0 && nd.op == DOP_MUL && nd.width == 16 && nd.adj_l == 0 && nd.adj_r == 0) {
        bool small = true;
#pragma unroll
        for (int j = 0; j < R; j++) {
            const Val a = v[j][nd.lhs], b = v[j][nd.rhs];
            small = small && (((a.lo | b.lo) >> 32) == 0) && ((a.hi | b.hi) == 0);
        }
        const bool fast = __all_sync(0xffffffffu, small);
#pragma unroll
        for (int j = 0; j < R; j++) {
            const Val a = v[j][nd.lhs], b = v[j][nd.rhs];
            bool ov = false;
            if (fast) {
                v[j][I].lo = (uint64_t)(uint32_t)a.lo * (uint64_t)(uint32_t)b.lo;
                v[j][I].hi = 0;
            } else {
                const Mul128Res m = mul_overflow128_cold(a.lo, a.hi, b.lo, b.hi);
                v[j][I].lo = m.lo;
                v[j][I].hi = m.hi;
                ov = m.ov != 0;
            }
            v[j][I].null = a.null || b.null || ov;
        }
    } else {
#pragma unroll
        for (int j = 0; j < R; j++) {
            const Val a = v[j][nd.lhs], b = v[j][nd.rhs];
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
            v[j][I].lo = lo64(res);
            v[j][I].hi = hi64(res);
            v[j][I].null = a.null || b.null || ov;
        }
    }
}
template <class S, int R, int... I>
SRB_D void eval_static_all(const DProg& p, const StaticArgs& sa, Val (&v)[R][S::N_NODES], const uint8_t* stage, int r0, iseq<I...>) {
    (eval_static_node<S, I, R>(p, sa, v, stage, r0), ...);
}


// can the value of node n be NULL? (a column with a null map, a checked operation that overflowed)
template <class S>
constexpr bool node_can_be_null(int n) {
    if (n < 0) return false;
    const SNode nd = S::node(n);
    if (nd.op == DOP_LOAD) return nd.nullable != 0;
    if (nd.op == DOP_LIT) return false;
    if (nd.checked) return true;
    return node_can_be_null<S>(nd.lhs) || node_can_be_null<S>(nd.rhs);
}

// Accumulator planes of the low-cardinality path. Kind 1 (sum plane): SUM / AVG, and COUNT(x) over an argument that can be NULL (it needs the
// plane's NULL counter) — aggregates over the same argument share one plane, the first of them is its "representative". Kind 2 (max plane):
// MIN / MAX over a value of up to 8 bytes, one plane each, holding the table's own order-preserving encoding (enc_i64: identity zero).
// Kind 0: COUNT(*) and COUNT over an argument that is never NULL — the group's row count.
template <class S>
constexpr int plane_kind(int a) {
    const SAgg g = S::agg(a);
    if (g.fn == FN_SUM || g.fn == FN_AVG) return g.dbl != 0 ? 3 : 1; // kind 3: a sum kept as a double (no spill word, no carries)
    if (g.fn == FN_COUNT) return (g.merge == 0 && g.arg >= 0 && node_can_be_null<S>(g.arg)) ? 1 : 0;
    return 2;
}
template <class S>
constexpr int agg_rep(int a) {
    const int kind = plane_kind<S>(a);
    if (kind != 1 && kind != 3) return a;
    for (int b = 0; b < a; b++)
        if (plane_kind<S>(b) == kind && S::agg(b).arg == S::agg(a).arg && (kind == 1 || S::agg(b).dbl == S::agg(a).dbl)) return b;
    return a;
}
// dense index of a representative among the representatives (its accumulator plane), and their number
template <class S>
constexpr int rep_index(int a) {
    int n = 0;
    for (int b = 0; b < a; b++)
        if (plane_kind<S>(b) != 0 && agg_rep<S>(b) == b) n++;
    return n;
}
template <class S>
constexpr int n_reps() {
    return rep_index<S>(S::N_AGGS);
}
// Planes whose argument reads a column with a null map see NULLs on ordinary rows, not once in a blue moon: they keep a per-thread count
// of the rows they TOOK in shared memory (one more u32 per (plane, group, thread)) instead of sending every NULL down the rare path.
template <class S>
constexpr bool node_reads_nullable(int n) {
    if (n < 0) return false;
    const SNode nd = S::node(n);
    if (nd.op == DOP_LOAD) return nd.nullable != 0;
    if (nd.op == DOP_LIT) return false;
    return node_reads_nullable<S>(nd.lhs) || node_reads_nullable<S>(nd.rhs);
}
template <class S>
constexpr bool plane_counts(int a) { // a: a representative
    return plane_kind<S>(a) != 0 && node_reads_nullable<S>(S::agg(a).arg);
}
template <class S>
constexpr int cnt_index(int a) {
    int n = 0;
    for (int b = 0; b < a; b++)
        if (plane_kind<S>(b) != 0 && agg_rep<S>(b) == b && plane_counts<S>(b)) n++;
    return n;
}
template <class S>
constexpr int n_cnt() {
    return cnt_index<S>(S::N_AGGS);
}
// a double-sum plane?
template <class S>
constexpr bool plane_is_dbl(int ri) {
    for (int a = 0; a < S::N_AGGS; a++)
        if (plane_kind<S>(a) == 3 && agg_rep<S>(a) == a && rep_index<S>(a) == ri) return true;
    return false;
}
// is plane ri a max plane? (fold() walks the planes by index)
template <class S>
constexpr bool plane_is_max(int ri) {
    for (int a = 0; a < S::N_AGGS; a++)
        if (plane_kind<S>(a) == 2 && rep_index<S>(a) == ri) return true;
    return false;
}
// Per-thread accumulators of the low-cardinality path. The LOW 64 bits of every partial sum and the row count live in
// shared memory, one private word per (representative, group, thread): lanes of a warp hit distinct banks whatever their
// groups are, so an update is a plain LDS / add / STS with the group as a runtime address — no atomics, no per-group
// predication. What is almost never non-zero per row — the carry / sign word above bit 63 and NULL arguments — is
// collected into ONE warp vote per row; only when it fires are the rare-path counters (thread-local memory, indexed by
// group) touched.
template <class S>
struct AccGeom {
    static constexpr int G = S::GMAX > 0 ? S::GMAX : 1;
    static constexpr int NR = n_reps<S>() > 0 ? n_reps<S>() : 1;
    static constexpr int LO_BYTES = NR * G * S::THREADS * 8; // [rep][group][thread] u64
    static constexpr int ROWS_BYTES = G * S::THREADS * 4;    // [group][thread] u32
    static constexpr int TRASH_BYTES = S::THREADS * 8;       // [thread] u64: where rows that do not count are steered
    static constexpr int NC = n_cnt<S>();
    static constexpr int CNT_BYTES = NC * G * S::THREADS * 4; // [counting plane][group][thread] u32: rows the plane took
    static constexpr int BYTES = S::GMAX > 0 ? LO_BYTES + ROWS_BYTES + TRASH_BYTES + CNT_BYTES : 0;
};
template <class S>
struct RareAcc {
    uint64_t lo[AccGeom<S>::G * AccGeom<S>::NR];    // what the narrow path moved out of a shared-memory word that reached 2^63
    uint64_t hi[AccGeom<S>::G * AccGeom<S>::NR];    // bits 64..127 of each partial sum (sign-extended for 8-byte arguments)
    uint32_t nulls[AccGeom<S>::G * AccGeom<S>::NR]; // selected rows whose argument was NULL
};

// old (64-bit accumulator word) + x (128-bit addend): the new word and what spills above bit 63, as one carry chain
SRB_D void add_word_spill(uint64_t old, uint64_t xlo, uint64_t xhi, uint64_t& nw, uint64_t& spill) {
    uint32_t n0, n1, d0, d1;
    asm("add.cc.u32 %0, %4, %6;\n\t"
        "addc.cc.u32 %1, %5, %7;\n\t"
        "addc.cc.u32 %2, %8, 0;\n\t"
        "addc.u32 %3, %9, 0;"
        : "=r"(n0), "=r"(n1), "=r"(d0), "=r"(d1)
        : "r"((uint32_t)old), "r"((uint32_t)(old >> 32)), "r"((uint32_t)xlo), "r"((uint32_t)(xlo >> 32)), "r"((uint32_t)xhi), "r"((uint32_t)(xhi >> 32)));
    nw = ((uint64_t)n1 << 32) | n0;
    spill = ((uint64_t)d1 << 32) | d0;
}

// One representative, one row. Branch-free: a row that does not count (filtered out, NULL argument) is steered to the
// thread's trash word instead of being skipped, so the instruction stream is the same for every lane.
template <class S, int A>
SRB_D void acc_one(uint8_t* a, uint8_t* c, uint8_t* trash, const Val* v, bool on, uint64_t* dh, uint64_t& row_spill, bool& any_null) {
    constexpr SAgg g = S::agg(A);
    if constexpr (plane_kind<S>(A) == 1 && agg_rep<S>(A) == A) {
        constexpr int RI = rep_index<S>(A);
        const Val& x = v[g.arg];
        bool take = on;
        if constexpr (node_can_be_null<S>(g.arg)) {
            take = on && !x.null;
            if constexpr (!plane_counts<S>(A)) any_null = any_null || (on && x.null);
        }
        uint64_t* slot = (uint64_t*)(take ? a + (size_t)RI * AccGeom<S>::G * S::THREADS * 8 : trash);
        uint64_t nw;
        add_word_spill(*slot, x.lo, (uint64_t)x.hi, nw, dh[RI]);
        *slot = nw;
        if constexpr (plane_counts<S>(A)) {
            row_spill |= take ? dh[RI] : 0ULL;
            uint32_t* tc = (uint32_t*)(take ? c + (size_t)cnt_index<S>(A) * AccGeom<S>::G * S::THREADS * 4 : trash);
            *tc += 1u;
        } else {
            row_spill |= dh[RI]; // meaningless for a trashed row: the caller masks it (or any_null already forces the rare path)
        }
    } else if constexpr (plane_kind<S>(A) == 2) {
        constexpr int RI = rep_index<S>(A);
        const Val& x = v[g.arg];
        bool take = on;
        if constexpr (node_can_be_null<S>(g.arg)) {
            take = on && !x.null;
            if constexpr (!plane_counts<S>(A)) any_null = any_null || (on && x.null);
        }
        uint64_t* slot = (uint64_t*)(take ? a + (size_t)RI * AccGeom<S>::G * S::THREADS * 8 : trash);
        const uint64_t e = g.dbl != 0 ? enc_f64(x.lo, g.fn == FN_MIN) : enc_i64((int64_t)x.lo, g.fn == FN_MIN), cur = *slot;
        *slot = e > cur ? e : cur;
        dh[RI] = 0;
        if constexpr (plane_counts<S>(A)) {
            uint32_t* tc = (uint32_t*)(take ? c + (size_t)cnt_index<S>(A) * AccGeom<S>::G * S::THREADS * 4 : trash);
            *tc += 1u;
        }
    } else if constexpr (plane_kind<S>(A) == 3 && agg_rep<S>(A) == A) {
        constexpr int RI = rep_index<S>(A);
        const Val& x = v[g.arg];
        bool take = on;
        if constexpr (node_can_be_null<S>(g.arg)) {
            take = on && !x.null;
            if constexpr (!plane_counts<S>(A)) any_null = any_null || (on && x.null);
        }
        uint64_t* slot = (uint64_t*)(take ? a + (size_t)RI * AccGeom<S>::G * S::THREADS * 8 : trash);
        const double val = g.dbl == 1 ? __longlong_as_double((long long)x.lo) : (double)(int64_t)x.lo;
        *slot = (uint64_t)__double_as_longlong(__longlong_as_double((long long)*slot) + val);
        dh[RI] = 0;
        if constexpr (plane_counts<S>(A)) {
            uint32_t* tc = (uint32_t*)(take ? c + (size_t)cnt_index<S>(A) * AccGeom<S>::G * S::THREADS * 4 : trash);
            *tc += 1u;
        }
    }
}
template <class S, int A>
SRB_D void acc_rare_one(RareAcc<S>& ra, int gid, const Val* v, bool on, const uint64_t* dh) {
    constexpr SAgg g = S::agg(A);
    if constexpr (plane_kind<S>(A) != 0 && agg_rep<S>(A) == A) {
        constexpr int RI = rep_index<S>(A);
        bool take = on;
        if constexpr (node_can_be_null<S>(g.arg)) {
            take = on && !v[g.arg].null;
            if constexpr (!plane_counts<S>(A)) ra.nulls[gid * AccGeom<S>::NR + RI] += (on && v[g.arg].null) ? 1u : 0u;
        }
        if constexpr (plane_kind<S>(A) == 1) ra.hi[gid * AccGeom<S>::NR + RI] += take ? dh[RI] : 0ULL;
    }
}
// the R rows of a tile into their groups (0 <= gid[j] < G; on[j] false: the row does not count). One vote covers the
// spills and NULLs of all R rows. All 32 lanes must call.
template <class S, int R, int... A>
SRB_D void acc_update_rows(RareAcc<S>& ra, uint8_t* acc_tid, uint8_t* rows_tid, uint8_t* trash, const Val (&v)[R][S::N_NODES], const bool (&on)[R],
                           const int (&gid)[R], iseq<A...>) {
    uint64_t dh[R][AccGeom<S>::NR];
    uint64_t any_spill = 0;
    bool any_null = false;
#pragma unroll
    for (int j = 0; j < R; j++) {
        uint8_t* a = acc_tid + (size_t)gid[j] * (S::THREADS * 8);
        uint8_t* c = rows_tid + AccGeom<S>::ROWS_BYTES + AccGeom<S>::TRASH_BYTES + (size_t)gid[j] * (S::THREADS * 4); // the counting planes of this group
        uint64_t row_spill = 0;
        (acc_one<S, A>(a, c, trash, v[j], on[j], dh[j], row_spill, any_null), ...);
        any_spill |= on[j] ? row_spill : 0ULL;
        uint32_t* rc = (uint32_t*)(on[j] ? rows_tid + (size_t)gid[j] * (S::THREADS * 4) : trash);
        *rc += 1u;
    }
    if (__any_sync(0xffffffffu, any_spill != 0 || any_null)) {
#pragma unroll
        for (int j = 0; j < R; j++) (acc_rare_one<S, A>(ra, gid[j], v[j], on[j], dh[j]), ...);
    }
}


// ---- the narrow tile body ---------------------------------------------------------------------------------------------------
// Money amounts, quantities, rates and their products are small non-negative integers. When EVERY value of a warp's rows is
// (every column value and every operand in [0, 2^32), products in [0, 2^64)), no checked operation can overflow, nothing is
// NULL, and the exact results are plain 32/64-bit unsigned arithmetic. The narrow body computes exactly that and ORs into
// `esc` every bit that would break the assumption; one warp vote on esc (together with "a group is missing from the
// directory") decides whether its results are used or the tile is redone by the general body. Values are identical either
// way; only the instruction count differs (no 128-bit products, no overflow flags, no sign words).
template <class S>
constexpr bool node_is_operand(int n) {
    for (int i = 0; i < S::N_NODES; i++) {
        const SNode nd = S::node(i);
        if (nd.op != DOP_LOAD && nd.op != DOP_LIT && (nd.lhs == n || nd.rhs == n)) return true;
    }
    return false;
}
template <class S>
constexpr bool node_is_key(int n) {
    for (int k = 0; k < S::N_KEYS; k++)
        if (S::key(k).node == n) return true;
    return false;
}
template <class S>
constexpr bool shape_has_narrow() {
    bool ok = S::GMAX > 0;
    for (int i = 0; ok && i < S::N_NODES; i++) {
        const SNode nd = S::node(i);
        const bool key = node_is_key<S>(i);
        if (key && (nd.op != DOP_LOAD || node_is_operand<S>(i))) return false; // keys: plain columns, bits taken as they are
        switch (nd.op) {
        case DOP_LOAD:
            if ((key && nd.nullable != 0) || (nd.width != 8 && !(key && (nd.width == 4 || nd.width == 0)))) return false;
            break; // (a nullable VALUE column is fine: the narrow body carries one NULL bit per such column and row)
        case DOP_LIT: break;
        case DOP_CAST:
            if (nd.width < 8) return false;
            break;
        case DOP_ADD:
        case DOP_SUB:
        case DOP_MUL:
        case DOP_MULFAST:
            if (nd.width < 8 || nd.adj_l != 0 || nd.adj_r != 0) return false;
            break;
        default: return false;
        }
    }
    for (int a = 0; a < S::N_AGGS; a++) {
        const SAgg g = S::agg(a);
        if (g.merge != 0 || g.dbl != 0) return false;
        // every function qualifies: under the narrow assumption nothing is NULL, so COUNT(x) is the row count, and MIN / MAX arguments are
        // 8-byte values in [0, 2^63)
        if ((g.fn == FN_MIN || g.fn == FN_MAX) && S::node(g.arg).width != 8) return false;
    }
    for (int k = 0; k < S::N_KEYS; k++)
        if (S::key(k).nullable != 0) return false;
    return true;
}
// NULL bits of the narrow body: nullable LOAD node I owns bit null_bit<S>(I) of a row's mask; a value is NULL iff the mask meets the bits of
// the nullable loads under it (null_dep). A NULL load contributes the value 0, so nothing derived from it can break the narrow assumption.
template <class S>
constexpr int null_bit(int n) {
    int b = 0;
    for (int i = 0; i < n; i++)
        if (S::node(i).op == DOP_LOAD && S::node(i).nullable != 0) b++;
    return b;
}
template <class S>
constexpr uint32_t null_dep(int n) {
    if (n < 0) return 0u;
    const SNode nd = S::node(n);
    if (nd.op == DOP_LOAD) return nd.nullable != 0 ? (1u << null_bit<S>(n)) : 0u;
    if (nd.op == DOP_LIT) return 0u;
    return null_dep<S>(nd.lhs) | null_dep<S>(nd.rhs);
}
template <class S, int I, int R>
SRB_D void eval_narrow_node(const DProg& p, const StaticArgs& sa, uint64_t (&nv)[R][S::N_NODES], uint32_t (&ln)[R], const uint8_t* stage, int r0, uint32_t& esc) {
    constexpr SNode nd = S::node(I);
    if constexpr (nd.op == DOP_LOAD && nd.width == 0) { // a binary key: (length << 32) | offset
#pragma unroll
        for (int j = 0; j < R; j++) {
            const uint32_t* o = (const uint32_t*)(stage + load_off<S>(I)) + r0 + j * S::THREADS;
            const uint32_t b = o[0], e = o[1];
            nv[j][I] = ((uint64_t)(e - b) << 32) | b;
        }
    } else if constexpr (nd.op == DOP_LOAD && nd.width == 4) {
#pragma unroll
        for (int j = 0; j < R; j++) nv[j][I] = ((const uint32_t*)(stage + load_off<S>(I)))[r0 + j * S::THREADS]; // key bits
    } else if constexpr (nd.op == DOP_LOAD && nd.nullable != 0) {
#pragma unroll
        for (int j = 0; j < R; j++) {
            const bool isnull = stage[sa.node_null_off[I] + r0 + j * S::THREADS] != 0;
            const uint64_t x = isnull ? 0ULL : ((const uint64_t*)(stage + load_off<S>(I)))[r0 + j * S::THREADS];
            ln[j] |= isnull ? (1u << null_bit<S>(I)) : 0u;
            esc |= (uint32_t)(x >> 32);
            nv[j][I] = (uint32_t)x;
        }
    } else if constexpr (nd.op == DOP_LOAD) {
#pragma unroll
        for (int j = 0; j < R; j++) {
            const uint64_t x = ((const uint64_t*)(stage + load_off<S>(I)))[r0 + j * S::THREADS];
            esc |= (uint32_t)(x >> 32);
            nv[j][I] = (uint32_t)x;
        }
    } else if constexpr (nd.op == DOP_LIT) {
#pragma unroll
        for (int j = 0; j < R; j++) nv[j][I] = (uint32_t)p.nodes[I].lit_lo;
    } else if constexpr (nd.op == DOP_CAST) {
#pragma unroll
        for (int j = 0; j < R; j++) nv[j][I] = nv[j][nd.lhs];
    } else if constexpr (nd.op == DOP_ADD || nd.op == DOP_SUB) {
#pragma unroll
        for (int j = 0; j < R; j++) {
            const uint64_t r = nd.op == DOP_ADD ? nv[j][nd.lhs] + nv[j][nd.rhs] : nv[j][nd.lhs] - nv[j][nd.rhs];
            esc |= (uint32_t)(r >> 32); // a negative difference sets every high bit
            nv[j][I] = (uint32_t)r;
        }
    } else { // MUL / MULFAST: operands are < 2^32, the product is exact in 64 bits
#pragma unroll
        for (int j = 0; j < R; j++) {
            const uint64_t r = (uint64_t)(uint32_t)nv[j][nd.lhs] * (uint64_t)(uint32_t)nv[j][nd.rhs];
            if constexpr (node_is_operand<S>(I)) {
                esc |= (uint32_t)(r >> 32);
                nv[j][I] = (uint32_t)r;
            } else {
                if constexpr (nd.width == 8) esc |= (uint32_t)(r >> 32) & 0x80000000u; // a checked 8-byte product must stay below 2^63
                nv[j][I] = r;
            }
        }
    }
}
template <class S, int R, int... I>
SRB_D void eval_narrow_all(const DProg& p, const StaticArgs& sa, uint64_t (&nv)[R][S::N_NODES], uint32_t (&ln)[R], const uint8_t* stage, int r0, uint32_t& esc,
                           iseq<I...>) {
#pragma unroll
    for (int j = 0; j < R; j++) ln[j] = 0;
    (eval_narrow_node<S, I, R>(p, sa, nv, ln, stage, r0, esc), ...);
}
template <class S, int KI>
SRB_D void key_put_narrow(PackedKey& k, const uint64_t* nv, const TileBin& tb) {
    if constexpr (KI < S::N_KEYS) {
        constexpr SKey kd = S::key(KI);
        if constexpr (kd.width == 0)
            key_put_binary(k, tb.base[KI] + (uint32_t)nv[kd.node], (uint32_t)(nv[kd.node] >> 32));
        else
            key_put(k, nv[kd.node], kd.width);
    }
}
template <class S>
SRB_D PackedKey pack_key_narrow(const uint64_t* nv, const TileBin& tb) {
    PackedKey k{0, 0, 0, false};
    key_put_narrow<S, 0>(k, nv, tb);
    key_put_narrow<S, 1>(k, nv, tb);
    key_put_narrow<S, 2>(k, nv, tb);
    key_put_narrow<S, 3>(k, nv, tb);
    return k;
}

// x (< 2^64, non-negative) into a shared-memory word: plain 64-bit add. No carry can be lost while the word and the
// addend are both below 2^63; `flag` collects the top bits of both, and when its sign bit is set after the tile the
// exact carries are recovered from the words themselves (acc_narrow_fixup).
template <class S, int A>
SRB_D void acc_narrow_one(uint8_t* a, uint8_t* c, uint8_t* trash, const uint64_t* nv, uint32_t ln, bool on, uint32_t& flag) {
    constexpr SAgg g = S::agg(A);
    if constexpr (plane_kind<S>(A) == 1 && agg_rep<S>(A) == A) {
        constexpr int RI = rep_index<S>(A);
        const uint64_t x = nv[g.arg];
        bool take = on;
        if constexpr (null_dep<S>(g.arg) != 0) take = on && (ln & null_dep<S>(g.arg)) == 0;
        uint64_t* slot = (uint64_t*)(take ? a + (size_t)RI * AccGeom<S>::G * S::THREADS * 8 : trash);
        const uint64_t old = *slot;
        *slot = old + x;
        flag |= (uint32_t)(old >> 32) | (uint32_t)(x >> 32);
        if constexpr (plane_counts<S>(A)) {
            uint32_t* tc = (uint32_t*)(take ? c + (size_t)cnt_index<S>(A) * AccGeom<S>::G * S::THREADS * 4 : trash);
            *tc += 1u;
        }
    } else if constexpr (plane_kind<S>(A) == 2 && null_dep<S>(S::agg(A).arg) != 0) {
        constexpr int RI = rep_index<S>(A);
        const bool take = on && (ln & null_dep<S>(g.arg)) == 0;
        uint64_t* slot = (uint64_t*)(a + (size_t)RI * AccGeom<S>::G * S::THREADS * 8);
        const uint64_t e = enc_i64((int64_t)nv[g.arg], g.fn == FN_MIN), cur = *slot;
        *slot = (take && e > cur) ? e : cur;
        uint32_t* tc = (uint32_t*)(take ? c + (size_t)cnt_index<S>(A) * AccGeom<S>::G * S::THREADS * 4 : trash);
        *tc += 1u;
    } else if constexpr (plane_kind<S>(A) == 2) {
        // a max plane: the word itself, whatever `on` says (an encoded value has its top bit set: it must stay away from the trash word,
        // whose top bit the sum planes read as "look for carries")
        constexpr int RI = rep_index<S>(A);
        uint64_t* slot = (uint64_t*)(a + (size_t)RI * AccGeom<S>::G * S::THREADS * 8);
        const uint64_t e = enc_i64((int64_t)nv[g.arg], g.fn == FN_MIN), cur = *slot;
        *slot = (on && e > cur) ? e : cur;
    }
}
// Rare: some word or addend of the warp had bit 63 set. Every add of the tile has already been applied modulo 2^64; the
// carry of row j's add is (word right after it) < x_j, and that word is the final one minus the later rows' addends to the
// same word. Words that reached 2^63 are then moved to the thread-local low/high pair so the next tiles stay quiet.
template <class S, int R>
struct NarrowFix { // by value into the out-of-line fix-up: the tile's arrays must not have their address taken
    uint64_t x[R][AccGeom<S>::NR];
    int gid[R];
    bool on[R];
};
template <class S, int R, int A>
SRB_D void narrow_fix_put(NarrowFix<S, R>& f, const uint64_t (&nv)[R][S::N_NODES], const uint32_t (&ln)[R]) {
    constexpr SAgg g = S::agg(A);
    if constexpr (plane_kind<S>(A) == 1 && agg_rep<S>(A) == A) {
#pragma unroll
        for (int j = 0; j < R; j++) // (a row whose argument was NULL added nothing to this plane: it carries nothing)
            f.x[j][rep_index<S>(A)] = (ln[j] & null_dep<S>(g.arg)) ? 0ULL : nv[j][g.arg];
    }
}
template <class S>
struct MaxPlaneMask { // bit ri: plane ri is a max plane (encodings, not sums: nothing to carry)
    static constexpr uint32_t value() {
        uint32_t m = 0;
        for (int ri = 0; ri < AccGeom<S>::NR; ri++) m |= plane_is_max<S>(ri) ? (1u << ri) : 0u;
        return m;
    }
};
template <class S, int R>
static __device__ __noinline__ void acc_narrow_fixup(RareAcc<S>* ra, uint8_t* acc_tid, uint8_t* trash, const NarrowFix<S, R> f) {
    constexpr int NR = AccGeom<S>::NR;
    constexpr uint32_t max_planes = MaxPlaneMask<S>::value();
    for (int ri = 0; ri < NR; ri++) {
        if ((max_planes >> ri) & 1u) continue;
        for (int j = R - 1; j >= 0; j--) {
            if (!f.on[j]) continue;
            const uint64_t* slot = (const uint64_t*)(acc_tid + (size_t)f.gid[j] * (S::THREADS * 8) + (size_t)ri * AccGeom<S>::G * S::THREADS * 8);
            uint64_t after = *slot;
            for (int k = j + 1; k < R; k++)
                if (f.on[k] && f.gid[k] == f.gid[j]) after -= f.x[k][ri];
            if (after < f.x[j][ri]) ra->hi[f.gid[j] * NR + ri] += 1;
        }
        for (int j = 0; j < R; j++) {
            if (!f.on[j]) continue;
            uint64_t* slot = (uint64_t*)(acc_tid + (size_t)f.gid[j] * (S::THREADS * 8) + (size_t)ri * AccGeom<S>::G * S::THREADS * 8);
            const uint64_t w = *slot;
            if (w >> 63) {
                const uint64_t lo = ra->lo[f.gid[j] * NR + ri] + w;
                ra->hi[f.gid[j] * NR + ri] += lo < w ? 1 : 0;
                ra->lo[f.gid[j] * NR + ri] = lo;
                *slot = 0;
            }
        }
    }
    if (*(uint64_t*)trash >> 63) *(uint64_t*)trash = 0; // the trash word only ever matters through its top bit
}
template <class S, int R, int... A>
SRB_D void acc_narrow_rows(RareAcc<S>& ra, uint8_t* acc_tid, uint8_t* rows_tid, uint8_t* trash, const uint64_t (&nv)[R][S::N_NODES], const uint32_t (&ln)[R],
                           const bool (&on)[R], const int (&gid)[R], iseq<A...>) {
    uint32_t flag = 0;
#pragma unroll
    for (int j = 0; j < R; j++) {
        uint8_t* a = acc_tid + (size_t)gid[j] * (S::THREADS * 8);
        uint8_t* c = rows_tid + AccGeom<S>::ROWS_BYTES + AccGeom<S>::TRASH_BYTES + (size_t)gid[j] * (S::THREADS * 4); // the counting planes of this group
        (acc_narrow_one<S, A>(a, c, trash, nv[j], ln[j], on[j], flag), ...);
        uint32_t* rc = (uint32_t*)(on[j] ? rows_tid + (size_t)gid[j] * (S::THREADS * 4) : trash);
        *rc += 1u;
    }
    if (__any_sync(0xffffffffu, (int32_t)flag < 0)) {
        NarrowFix<S, R> f;
        (narrow_fix_put<S, R, A>(f, nv, ln), ...);
#pragma unroll
        for (int j = 0; j < R; j++) {
            f.gid[j] = gid[j];
            f.on[j] = on[j];
        }
        acc_narrow_fixup<S, R>(&ra, acc_tid, trash, f);
    }
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

// Retiring a CTA's accumulators in two steps, so that the few table slots of a low-cardinality GROUP BY see one set of
// atomics per CTA rather than one per warp: (1) every warp reduces its lanes and stashes the totals in shared memory,
// (2) one thread per group adds the warps' totals and folds them into the group's table slot.
template <class S>
struct Flusher {
    static constexpr int G = AccGeom<S>::G;
    static constexpr int NR = AccGeom<S>::NR;
    static constexpr int WORDS = 3 * NR + 1; // per (warp, group): {lo, hi, nulls} per representative, then the row count
    static constexpr int STASH_BYTES = (S::THREADS / 32) * G * WORDS * 8;

    template <int A>
    static __device__ __forceinline__ void stash_rep(const RareAcc<S>& ra, const uint8_t* acc_tid, const uint8_t* rows_tid, int q, uint64_t* st) {
        constexpr SAgg g = S::agg(A);
        if constexpr (plane_kind<S>(A) != 0 && agg_rep<S>(A) == A) {
            constexpr int RI = rep_index<S>(A);
            const uint64_t w = *(const uint64_t*)(acc_tid + (size_t)(RI * G + q) * S::THREADS * 8);
            uint64_t lo, hi;
            if constexpr (plane_kind<S>(A) == 2) {
                lo = w;
                hi = 0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const uint64_t y = __shfl_xor_sync(0xffffffffu, lo, o);
                    lo = y > lo ? y : lo;
                }
            } else if constexpr (plane_kind<S>(A) == 3) {
                double sum = __longlong_as_double((long long)w);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                lo = (uint64_t)__double_as_longlong(sum);
                hi = 0;
            } else {
                lo = w + ra.lo[q * NR + RI];
                hi = ra.hi[q * NR + RI] + (lo < w ? 1ULL : 0ULL);
                warp_sum_u128(lo, hi);
            }
            uint64_t nulls = 0; // a counting plane: the rows it took; otherwise the selected rows whose argument was NULL
            if constexpr (plane_counts<S>(A))
                nulls = warp_sum_u64(*(const uint32_t*)(rows_tid + AccGeom<S>::ROWS_BYTES + AccGeom<S>::TRASH_BYTES +
                                                        (size_t)(cnt_index<S>(A) * G + q) * (S::THREADS * 4)));
            else if constexpr (node_can_be_null<S>(g.arg))
                nulls = warp_sum_u64(ra.nulls[q * NR + RI]);
            if ((threadIdx.x & 31) == 0) {
                st[RI * 3 + 0] = lo;
                st[RI * 3 + 1] = hi;
                st[RI * 3 + 2] = nulls;
            }
        }
    }
    // step 1, whole warps: group q of this warp's lanes -> stash
    template <int... A>
    static __device__ __forceinline__ void stash(const RareAcc<S>& ra, const uint8_t* acc_tid, const uint8_t* rows_tid, int q, uint64_t* stash_base,
                                                 iseq<A...>) {
        uint64_t* st = stash_base + (size_t)((threadIdx.x >> 5) * G + q) * WORDS;
        const uint64_t rows = warp_sum_u64(*(const uint32_t*)(rows_tid + (size_t)q * (S::THREADS * 4)));
        if ((threadIdx.x & 31) == 0) st[3 * NR] = rows;
        (stash_rep<A>(ra, acc_tid, rows_tid, q, st), ...);
    }
    template <int A>
    static __device__ __forceinline__ void emit(const DProg& p, uint8_t* slot, const uint64_t* tot, uint64_t rows_total) {
        constexpr SAgg g = S::agg(A);
        const DAgg& rt = p.aggs[A];
        if constexpr (plane_kind<S>(A) == 0) {
            if (rows_total) atomicAdd((unsigned long long*)(slot + rt.off_n), (unsigned long long)rows_total);
        } else {
            constexpr int RI = rep_index<S>(agg_rep<S>(A));
            constexpr bool counts = plane_counts<S>(agg_rep<S>(A));
            const uint64_t n = counts ? tot[RI * 3 + 2] : rows_total - tot[RI * 3 + 2];
            if (n) {
                if constexpr (g.fn == FN_COUNT) {
                    atomicAdd((unsigned long long*)(slot + rt.off_n), (unsigned long long)n);
                } else {
                    if constexpr (plane_kind<S>(A) == 2)
                        atomicMax((unsigned long long*)(slot + rt.off_lo), (unsigned long long)tot[RI * 3 + 0]);
                    else if constexpr (plane_kind<S>(A) == 3)
                        atomicAdd((double*)(slot + rt.off_lo), __longlong_as_double((long long)tot[RI * 3 + 0]));
                    else
                        atomic_add128(slot, rt.off_lo, rt.off_hi, tot[RI * 3 + 0], tot[RI * 3 + 1]);
                    if (rt.off_n >= 0) atomicAdd((unsigned long long*)(slot + rt.off_n), (unsigned long long)n);
                }
            }
        }
    }
    // step 2, one thread: totals of group q over the warps -> table slot
    // one warp's stash entry of plane RI into the running totals (which kind of plane it is must be a compile-time fact: evaluated per call,
    // the constexpr walks over the aggregate list cost more than the rest of the kernel's epilogue)
    template <int RI>
    static __device__ __forceinline__ void fold_plane(uint64_t* tot, const uint64_t* st) {
        if constexpr (plane_is_max<S>(RI)) {
            tot[RI * 3] = st[RI * 3] > tot[RI * 3] ? st[RI * 3] : tot[RI * 3];
        } else if constexpr (plane_is_dbl<S>(RI)) {
            tot[RI * 3] = (uint64_t)__double_as_longlong(__longlong_as_double((long long)tot[RI * 3]) + __longlong_as_double((long long)st[RI * 3]));
        } else {
            const uint64_t lo = tot[RI * 3] + st[RI * 3];
            tot[RI * 3 + 1] += st[RI * 3 + 1] + (lo < st[RI * 3] ? 1ULL : 0ULL);
            tot[RI * 3] = lo;
        }
        tot[RI * 3 + 2] += st[RI * 3 + 2];
    }
    template <int... RI>
    static __device__ __forceinline__ void fold_planes(uint64_t* tot, const uint64_t* st, iseq<RI...>) {
        (fold_plane<RI>(tot, st), ...);
    }
    template <int... A>
    static __device__ __forceinline__ void fold(const DProg& p, const uint64_t* stash_base, int q, uint8_t* slot, iseq<A...>) {
        uint64_t tot[WORDS];
#pragma unroll
        for (int k = 0; k < WORDS; k++) tot[k] = 0;
        for (int w = 0; w < S::THREADS / 32; w++) {
            const uint64_t* st = stash_base + (size_t)(w * G + q) * WORDS;
            fold_planes(tot, st, make_iseq<NR>{});
            tot[3 * NR] += st[3 * NR];
        }
        (emit<A>(p, slot, tot, tot[3 * NR]), ...);
    }
};


// per-row update of a table slot with compile-time operand indexes (static shapes carry SUM / AVG / COUNT over integers)
template <class S, int A>
SRB_D void row_update_global_one(const DProg& p, uint8_t* slot, const Val* v) {
    constexpr SAgg g = S::agg(A);
    const DAgg& rt = p.aggs[A];
    if constexpr (g.fn == FN_COUNT && g.merge != 0) {
        if (!v[g.arg].null) atomicAdd((unsigned long long*)(slot + rt.off_n), (unsigned long long)v[g.arg].lo);
    } else if constexpr (g.fn == FN_COUNT) {
        bool on = true;
        if constexpr (g.arg >= 0) on = !v[g.arg].null;
        if (on) atomicAdd((unsigned long long*)(slot + rt.off_n), 1ULL);
    } else {
        const Val& x = v[g.arg];
        if (!x.null) {
            if constexpr (g.fn == FN_MIN || g.fn == FN_MAX)
                atomicMax((unsigned long long*)(slot + rt.off_lo),
                          (unsigned long long)(g.dbl != 0 ? enc_f64(x.lo, g.fn == FN_MIN) : enc_i64((int64_t)x.lo, g.fn == FN_MIN)));
            else if constexpr (g.dbl != 0)
                atomicAdd((double*)(slot + rt.off_lo), g.dbl == 1 ? __longlong_as_double((long long)x.lo) : (double)(int64_t)x.lo);
            else
                atomic_add128(slot, rt.off_lo, rt.off_hi, x.lo, (uint64_t)x.hi);
            if (rt.off_n >= 0) atomicAdd((unsigned long long*)(slot + rt.off_n), 1ULL);
        }
    }
}
template <class S, int... A>
SRB_D void row_update_global(const DProg& p, uint8_t* slot, const Val* v, iseq<A...>) {
    (row_update_global_one<S, A>(p, slot, v), ...);
}

template <class S>
constexpr int static_key_bytes() {
    int n = 0;
    for (int k = 0; k < S::N_KEYS; k++) {
        if (S::key(k).width == 0) return 0; // a binary key: the packed length is a per-row fact (PackedKey::len)
        n += S::key(k).width + (S::key(k).nullable != 0 ? 1 : 0);
    }
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

// key column K of the packed key: [null byte if the column is nullable][value, zero when NULL] (pack_key of srb_rowpipe.cuh)
template <class S, int KI>
SRB_D void key_put_static(PackedKey& k, const Val* v, const TileBin& tb) {
    if constexpr (KI < S::N_KEYS) {
        constexpr SKey kd = S::key(KI);
        const Val& x = v[kd.node];
        if constexpr (kd.width == 0) {
            bool isnull = false;
            if constexpr (kd.nullable != 0) {
                isnull = x.null;
                key_put(k, isnull ? 1 : 0, 1);
            }
            key_put_binary(k, tb.base[KI] + (uint32_t)x.lo, isnull ? 0u : (uint32_t)x.hi);
        } else if constexpr (kd.nullable != 0) {
            key_put(k, x.null ? 1 : 0, 1);
            key_put(k, x.null ? 0ULL : x.lo, kd.width);
        } else {
            key_put(k, x.lo, kd.width);
        }
    }
}
template <class S>
SRB_D PackedKey pack_key_static(const Val* v, const TileBin& tb) {
    PackedKey k{0, 0, 0, false};
    key_put_static<S, 0>(k, v, tb);
    key_put_static<S, 1>(k, v, tb);
    key_put_static<S, 2>(k, v, tb);
    key_put_static<S, 3>(k, v, tb);
    return k;
}
// the current tile's byte sources of the binary key columns, from the headers the producer wrote into the stage
template <class S, int KI>
SRB_D void tile_bin_one(TileBin& tb, const DProg& p, const StaticArgs& sa, const uint8_t* stage) {
    if constexpr (KI < S::N_KEYS) {
        if constexpr (S::key(KI).width == 0) {
            const int c = p.nodes[S::key(KI).node].col;
            const uint2 hdr = *(const uint2*)(stage + sa.bin_hdr_off[c]);
            tb.base[KI] = hdr.y ? (const uint8_t*)p.cols[c].data
                                : (const uint8_t*)((uintptr_t)(stage + sa.bin_win_off[c]) - (uintptr_t)hdr.x); // + a value's offset >= hdr.x: inside the window
        }
    }
}
template <class S>
SRB_D TileBin tile_bin(const DProg& p, const StaticArgs& sa, const uint8_t* stage) {
    TileBin tb{};
    if constexpr (shape_has_binary<S>()) {
        tile_bin_one<S, 0>(tb, p, sa, stage);
        tile_bin_one<S, 1>(tb, p, sa, stage);
        tile_bin_one<S, 2>(tb, p, sa, stage);
        tile_bin_one<S, 3>(tb, p, sa, stage);
    }
    return tb;
}

struct StaticCtr {
    unsigned long long* rows_selected;
    unsigned long long* deferred;
    unsigned long long* inserted;
};

SRB_D void mbar_arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory"); }

// The CTA's group directory (shared memory): packed key -> accumulator bank + table slot. Entries below cnt are immutable.
template <int G>
struct GroupDir {
    uint64_t lo[G], hi[G];
    uint8_t* slot[G];
    int cnt, lock;
};
template <int G>
SRB_D bool dir_has(const GroupDir<G>* d, int c, uint64_t klo, uint64_t khi) {
    bool present = false;
    for (int q = 0; q < c; q++)
        present = present || (((const volatile uint64_t*)d->lo)[q] == klo && ((const volatile uint64_t*)d->hi)[q] == khi);
    return present;
}
// Rows of a warp whose group is not in the directory yet (need). One leader per distinct missing key of the warp takes
// the directory lock and — only if the group is still absent — claims its table slot and publishes it, so the HBM table
// sees one probe per (CTA, group) instead of one per row while the directory fills. Tries are bounded: a lane that gives
// up leaves its rows to the global-table path, which is always correct. All 32 lanes call; returns 1 for a fresh table
// insert by this lane. Out of line: it runs a handful of times per CTA.
template <int G, int KEYLEN>
static __device__ __noinline__ int dir_resolve(GroupDir<G>* d, uint8_t* slots, uint64_t mask, int slot_bytes, int key_words, uint64_t klo, uint64_t khi,
                                               int klen, int seed, bool need) {
    const unsigned need_mask = __ballot_sync(0xffffffffu, need);
    const unsigned peers = __match_any_sync(0xffffffffu, klo ^ (khi * 0x9e3779b97f4a7c15ULL)) & need_mask;
    int inserted = 0;
    if (need && (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0) {
        for (int t = 0; t < 256; t++) {
            const int c = *(volatile int*)&d->cnt;
            if (c >= G || dir_has(d, c, klo, khi)) break;
            if (atomicCAS(&d->lock, 0, 1) == 0) {
                const int c2 = *(volatile int*)&d->cnt;
                if (c2 < G && !dir_has(d, c2, klo, khi)) {
                    const TableView tv{slots, mask, slot_bytes, key_words};
                    PackedKey k{klo, khi, KEYLEN ? KEYLEN : klen, false}; // KEYLEN 0: binary keys, the length travels with the row
                    bool fresh;
                    uint8_t* sl = table_find_or_insert(tv, klo, khi, hash_packed_key_static<KEYLEN>(k, seed), &fresh);
                    inserted = fresh ? 1 : 0;
                    if (sl) {
                        ((volatile uint64_t*)d->lo)[c2] = klo;
                        ((volatile uint64_t*)d->hi)[c2] = khi;
                        ((uint8_t* volatile*)d->slot)[c2] = sl;
                        __threadfence_block();
                        *(volatile int*)&d->cnt = c2 + 1;
                    }
                }
                __threadfence_block();
                atomicExch(&d->lock, 0);
                break;
            }
            __nanosleep(64);
        }
    }
    __syncwarp();
    return inserted;
}

// The per-row global-table update of a low-cardinality shape, out of line (it only runs for groups that found no room in
// the directory). The aggregate arguments travel by value; returns bit 0 = fresh insert, bit 1 = no slot (defer the row).
template <class S>
struct AggArgs {
    uint64_t lo[S::N_AGGS];
    int64_t hi[S::N_AGGS];
    uint32_t nulls; // bit a: the argument of aggregate a is NULL
};
template <class S, int... A>
SRB_D AggArgs<S> pack_agg_args(const Val* v, iseq<A...>) {
    AggArgs<S> r;
    r.nulls = 0;
    ((r.lo[A] = S::agg(A).arg >= 0 ? v[S::agg(A).arg < 0 ? 0 : S::agg(A).arg].lo : 0,
      r.hi[A] = S::agg(A).arg >= 0 ? v[S::agg(A).arg < 0 ? 0 : S::agg(A).arg].hi : 0,
      r.nulls |= (S::agg(A).arg >= 0 && v[S::agg(A).arg < 0 ? 0 : S::agg(A).arg].null) ? (1u << A) : 0u),
     ...);
    return r;
}
template <class S, int A>
SRB_D void row_update_args_one(const DProg& p, uint8_t* slot, const AggArgs<S>& x) {
    constexpr SAgg g = S::agg(A);
    const DAgg& rt = p.aggs[A];
    const bool null = (x.nulls >> A) & 1u;
    if constexpr (g.fn == FN_COUNT) {
        if (!null) atomicAdd((unsigned long long*)(slot + rt.off_n), 1ULL);
    } else {
        if (!null) {
            if constexpr (g.fn == FN_MIN || g.fn == FN_MAX)
                atomicMax((unsigned long long*)(slot + rt.off_lo),
                          (unsigned long long)(g.dbl != 0 ? enc_f64(x.lo[A], g.fn == FN_MIN) : enc_i64((int64_t)x.lo[A], g.fn == FN_MIN)));
            else if constexpr (g.dbl != 0)
                atomicAdd((double*)(slot + rt.off_lo), g.dbl == 1 ? __longlong_as_double((long long)x.lo[A]) : (double)(int64_t)x.lo[A]);
            else
                atomic_add128(slot, rt.off_lo, rt.off_hi, x.lo[A], (uint64_t)x.hi[A]);
            if (rt.off_n >= 0) atomicAdd((unsigned long long*)(slot + rt.off_n), 1ULL);
        }
    }
}
template <class S, int... A>
static __device__ __noinline__ int global_row_cold(const DProg* pp, uint8_t* slots, uint64_t mask, int slot_bytes, int key_words, uint64_t klo, uint64_t khi,
                                                   int klen, AggArgs<S> x, iseq<A...>) {
    const TableView tv{slots, mask, slot_bytes, key_words};
    const PackedKey k{klo, khi, static_key_bytes<S>() ? static_key_bytes<S>() : klen, false};
    bool fresh;
    uint8_t* slot = table_find_or_insert(tv, klo, khi, hash_packed_key_static<static_key_bytes<S>()>(k, pp->phase_seed), &fresh);
    if (!slot) return 2;
    (row_update_args_one<S, A>(*pp, slot, x), ...);
    return fresh ? 1 : 0;
}

// One persistent CTA per SM: S::THREADS consumer threads plus one producer warp. The producer's elected lane keeps the
// ring full (one bulk copy per column and tile, completing on full_bar[s]) and only it ever waits for a stage to drain
// (empty_bar[s], one arrival per consumer warp). Consumer warps therefore never wait for each other: a warp that is
// done with its rows of a tile arrives and moves on to the next stage, so up to n_stages tiles are being consumed at once
// and there is no block-wide barrier in the steady state.
template <class S>
__global__ void __launch_bounds__(S::THREADS + 32, 1)
        k_scan_agg_static(const __grid_constant__ DProg prog, const __grid_constant__ StaticArgs sa, const TableView tv, const int64_t base,
                          const int64_t n_tiles, uint32_t* __restrict__ defer_out, StaticCtr ctr) {
    constexpr int THREADS = S::THREADS;
    constexpr int R = S::ROWS;
    constexpr int TILE = tile_rows<S>();
    constexpr int G = AccGeom<S>::G;
    constexpr int KW = (static_key_bytes<S>() > 8 || (S::N_KEYS > 0 && static_key_bytes<S>() == 0)) ? 2 : 1; // binary keys: always two words
    static_assert(THREADS % 32 == 0 && THREADS + 32 <= 1024 && (TILE * 1) % 16 == 0, "block / tile geometry");
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t full_bar[ST_MAX_STAGES], empty_bar[ST_MAX_STAGES];
    __shared__ GroupDir<G> dir;
    const DProg& p = prog;
    using NodeSeq = make_iseq<S::N_NODES>;
    using AggSeq = make_iseq<S::N_AGGS>;

    const int stage_bytes = sa.stage_bytes, n_stages = sa.n_stages;
    const int tid = threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < n_stages; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], THREADS / 32);
        }
        dir.cnt = 0;
        dir.lock = 0;
        fence_barrier_init();
    }
    // this thread's private accumulator words start zeroed
    uint8_t* acc_tid = smem + sa.acc_off + tid * 8;
    uint8_t* rows_tid = smem + sa.acc_off + AccGeom<S>::LO_BYTES + tid * 4;
    uint8_t* trash = smem + sa.acc_off + AccGeom<S>::LO_BYTES + AccGeom<S>::ROWS_BYTES + tid * 8;
    RareAcc<S> ra;
    if constexpr (S::GMAX > 0) {
        if (tid < THREADS) {
#pragma unroll
            for (int k = 0; k < AccGeom<S>::NR * G; k++) *(uint64_t*)(acc_tid + (size_t)k * THREADS * 8) = 0;
#pragma unroll
            for (int k = 0; k < G; k++) *(uint32_t*)(rows_tid + (size_t)k * THREADS * 4) = 0;
#pragma unroll
            for (int k = 0; k < AccGeom<S>::NC * G; k++)
                *(uint32_t*)(rows_tid + AccGeom<S>::ROWS_BYTES + AccGeom<S>::TRASH_BYTES + (size_t)k * THREADS * 4) = 0;
        }
#pragma unroll
        for (int k = 0; k < AccGeom<S>::NR * G; k++) {
            ra.lo[k] = 0;
            ra.hi[k] = 0;
            ra.nulls[k] = 0;
        }
    }
    __syncthreads();

    uint32_t selected32 = 0; // a thread sees well under 2^32 rows of one launch (a launch covers at most 2^30 rows)
    unsigned long long deferred = 0, inserted = 0;
    if (tid >= THREADS) {
        // ---- producer warp ----
        // Lane c issues column c's bulk copy, lane 16 + c the copy of its null map: a tile's copies leave in one instruction slot instead of
        // one after the other from a single thread (with 7-9 copies per 22 KB stage the lone thread could not keep the ring full).
        static_assert(MAX_COLS <= 16, "one producer lane per column and per null map");
        const int lane = tid - THREADS;
        const uint8_t* src = nullptr; // this lane's source at row `base`, nullptr: nothing to copy
        uint32_t tile_bytes = 0, tile_stride = 0; // bytes copied per tile, and how far the source moves from tile to tile
        int dst_off = 0;
        // a binary column: the lane copies tile + 4 of its offsets and, per tile, the 16-byte aligned window of its bytes that the tile's rows
        // reference (bin_window_bytes at most; a tile of longer strings is flagged and its consumers read the column itself)
        const uint32_t* goffs = nullptr;
        const uint8_t* gdata = nullptr;
        int win_off = 0, hdr_off = 0;
        if (lane < p.n_cols) {
            if (shape_has_binary<S>() && p.cols[lane].kind == CK_BINARY) {
                goffs = p.cols[lane].offsets + base;
                gdata = (const uint8_t*)p.cols[lane].data;
                src = (const uint8_t*)goffs;
                tile_bytes = (uint32_t)(TILE * 4 + 16);
                tile_stride = (uint32_t)(TILE * 4);
                win_off = sa.bin_win_off[lane];
                hdr_off = sa.bin_hdr_off[lane];
            } else {
                const int w = p.cols[lane].width;
                src = (const uint8_t*)p.cols[lane].data + (size_t)base * w;
                tile_bytes = tile_stride = (uint32_t)(TILE * w);
            }
            dst_off = sa.col_off[lane];
        } else if (lane >= 16 && lane - 16 < p.n_cols && sa.null_off[lane - 16] >= 0) {
            src = p.cols[lane - 16].nulls + (size_t)base;
            tile_bytes = tile_stride = (uint32_t)TILE;
            dst_off = sa.null_off[lane - 16];
        }
        uint32_t nb = 0, ne = 0; // first and one-past-last byte of the NEXT tile's strings (fetched one tile ahead: the load's latency is off the ring's path)
        if (goffs && (int64_t)blockIdx.x < n_tiles) {
            nb = goffs[(size_t)blockIdx.x * TILE];
            ne = goffs[(size_t)blockIdx.x * TILE + TILE];
        }
        int s = 0;
        uint32_t parity = 1; // a fresh mbarrier passes a wait on parity 1: the first lap needs no arrivals
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            uint32_t total = (uint32_t)stage_bytes, wb = 0, wlen = 0;
            bool ovf = false;
            if constexpr (shape_has_binary<S>()) {
                uint32_t mine = src ? tile_bytes : 0u;
                if (goffs) {
                    wb = nb & ~15u;
                    wlen = (ne - wb + 15u) & ~15u;
                    ovf = wlen > (uint32_t)bin_window_bytes(TILE);
                    mine += ovf ? 0u : wlen;
                }
                total = __reduce_add_sync(0xffffffffu, mine);
            }
            if (lane == 0) mbar_wait(&empty_bar[s], parity);
            __syncwarp(); // the stage is free
            uint8_t* dst = smem + (size_t)s * stage_bytes;
            if constexpr (shape_has_binary<S>()) {
                if (goffs) *(uint2*)(dst + hdr_off) = make_uint2(wb, ovf ? 1u : 0u);
                __syncwarp(); // the headers are written before lane 0's arrive releases them to the consumers
            }
            if (lane == 0) mbar_expect_tx(&full_bar[s], total);
            __syncwarp(); // the byte count is registered before any copy can complete
            if (src) tma_load_1d(dst + dst_off, src + (size_t)tile * tile_stride, tile_bytes, &full_bar[s]);
            if constexpr (shape_has_binary<S>()) {
                if (goffs) {
                    if (!ovf && wlen) tma_load_1d(dst + win_off, gdata + wb, wlen, &full_bar[s]);
                    const int64_t next = tile + gridDim.x;
                    if (next < n_tiles) {
                        nb = goffs[(size_t)next * TILE];
                        ne = goffs[(size_t)next * TILE + TILE];
                    }
                }
            }
            if (++s == n_stages) {
                s = 0;
                parity ^= 1u;
            }
        }
    } else {
        // ---- consumer warps ----
        uint64_t kd_lo[G], kd_hi[G]; // register copy of the directory keys, refreshed when the directory grew
        int my_cnt = 0;
#pragma unroll
        for (int q = 0; q < G; q++) kd_lo[q] = kd_hi[q] = 0;
        int s = 0;
        uint32_t parity = 0;
        const uint32_t full0 = smem_u32(&full_bar[0]), empty0 = smem_u32(&empty_bar[0]);
        const uint8_t* stage = smem; // walks the ring with s
        const int nt = (int)n_tiles, tstep = (int)gridDim.x; // a launch covers <= 2^30 rows
        for (int tile = blockIdx.x; tile < nt; tile += tstep) {
            if (S::GMAX > 0 && S::N_KEYS > 0 && my_cnt < G) { // a full directory never changes again
                // my_cnt steers warp votes below, so it must be ONE value per warp even if the lanes poll at different times
                // (the minimum: every lane's own read saw at least that many published entries before it loads them)
                const int c = (int)__reduce_min_sync(0xffffffffu, (unsigned)*(volatile int*)&dir.cnt);
                if (c != my_cnt) {
#pragma unroll
                    for (int q = 0; q < G; q++) {
                        kd_lo[q] = ((volatile uint64_t*)dir.lo)[q];
                        if constexpr (KW == 2) kd_hi[q] = ((volatile uint64_t*)dir.hi)[q];
                    }
                    my_cnt = c;
                }
            }
            mbar_wait_a(full0 + 8u * (uint32_t)s, parity);
            __syncwarp(); // lanes may leave the wait loop on different polls
            bool pass[R];
            if (sa.n_ranges >= 0) {
                eval_ranges_rows<R, THREADS>(sa, stage, tid, pass);
            } else {
                int rows[R];
#pragma unroll
                for (int j = 0; j < R; j++) rows[j] = tid + j * THREADS;
                eval_preds_rows<R>(sa, stage, rows, pass);
            }
            bool any_pass = false;
#pragma unroll
            for (int j = 0; j < R; j++) {
                selected32 += pass[j] ? 1u : 0u;
                any_pass = any_pass || pass[j];
            }
            bool done = false; // warp-uniform: the narrow body handled this tile
            const TileBin tb = tile_bin<S>(p, sa, stage);
            if constexpr (shape_has_narrow<S>() && S::N_KEYS > 0) {
                if (sa.narrow_ok && __any_sync(0xffffffffu, any_pass)) {
                    uint64_t nv[R][S::N_NODES];
                    uint32_t esc = 0;
                    uint32_t ln[R];
                    eval_narrow_all<S, R>(p, sa, nv, ln, stage, tid, esc, NodeSeq{});
                    int gid[R];
                    bool bad = esc != 0;
#pragma unroll
                    for (int j = 0; j < R; j++) {
                        const PackedKey k = pack_key_narrow<S>(nv[j], tb);
                        gid[j] = G;
#pragma unroll
                        for (int q = G - 1; q >= 0; q--)
                            if (kd_lo[q] == k.lo && (KW == 1 || kd_hi[q] == k.hi)) gid[j] = q;
                        bad = bad || (pass[j] && (gid[j] >= my_cnt || k.too_long));
                        if (gid[j] >= my_cnt) gid[j] = 0;
                    }
                    if (!__any_sync(0xffffffffu, bad)) {
                        acc_narrow_rows<S, R>(ra, acc_tid, rows_tid, trash, nv, ln, pass, gid, AggSeq{});
                        done = true;
                    }
                }
            }
            if (!done && __any_sync(0xffffffffu, any_pass)) {
                Val v[R][S::N_NODES];
                eval_static_all<S, R>(p, sa, v, stage, tid, NodeSeq{});
                if constexpr (S::N_KEYS == 0) {
                    int gid[R];
#pragma unroll
                    for (int j = 0; j < R; j++) gid[j] = 0;
                    acc_update_rows<S, R>(ra, acc_tid, rows_tid, trash, v, pass, gid, AggSeq{});
                } else {
                    PackedKey k[R];
                    bool miss[R]; // row j still has to reach its group
#pragma unroll
                    for (int j = 0; j < R; j++) {
                        k[j] = pack_key_static<S>(v[j], tb);
                        if constexpr (shape_has_binary<S>()) {
                            if (k[j].too_long && pass[j]) { // a serialized key beyond 16 bytes: the row is left to the replay, which reports it
                                const int64_t i = (int64_t)tile * TILE + tid + j * THREADS;
                                atomicOr(&defer_out[i >> 5], 1u << (i & 31));
                                deferred++;
                                pass[j] = false;
                            }
                        }
                        miss[j] = pass[j];
                    }
                    if constexpr (S::GMAX > 0) {
                        int gid[R];
                        bool hit[R], need_any = false;
                        // descending: the lowest matching entry wins, and a match at or past my_cnt (a stale register) is discarded
                        auto lookup = [&]() {
#pragma unroll
                            for (int j = 0; j < R; j++) {
                                gid[j] = G;
#pragma unroll
                                for (int q = G - 1; q >= 0; q--)
                                    if (kd_lo[q] == k[j].lo && (KW == 1 || kd_hi[q] == k[j].hi)) gid[j] = q;
                                hit[j] = gid[j] < my_cnt;
                            }
                        };
                        lookup();
#pragma unroll
                        for (int j = 0; j < R; j++) need_any = need_any || (pass[j] && !hit[j]);
                        if (my_cnt < G && __any_sync(0xffffffffu, need_any)) {
                            // some row's group is not in the directory copy: publish it (dir_resolve), refresh the copy, look again
#pragma unroll
                            for (int j = 0; j < R; j++)
                                inserted += dir_resolve<G, static_key_bytes<S>()>(&dir, tv.slots, tv.mask, tv.slot_bytes, tv.key_words, k[j].lo, k[j].hi,
                                                                                  k[j].len, p.phase_seed, pass[j] && !hit[j]);
                            my_cnt = (int)__reduce_min_sync(0xffffffffu, (unsigned)*(volatile int*)&dir.cnt);
#pragma unroll
                            for (int q = 0; q < G; q++) {
                                kd_lo[q] = ((volatile uint64_t*)dir.lo)[q];
                                if constexpr (KW == 2) kd_hi[q] = ((volatile uint64_t*)dir.hi)[q];
                            }
                            lookup();
                        }
                        bool on[R];
#pragma unroll
                        for (int j = 0; j < R; j++) {
                            on[j] = pass[j] && hit[j];
                            miss[j] = pass[j] && !hit[j];
                            if (!hit[j]) gid[j] = 0;
                        }
                        acc_update_rows<S, R>(ra, acc_tid, rows_tid, trash, v, on, gid, AggSeq{});
                    }
                    bool any_miss = false;
#pragma unroll
                    for (int j = 0; j < R; j++) any_miss = any_miss || miss[j];
                    if (S::GMAX > 0 && !__any_sync(0xffffffffu, any_miss)) any_miss = false; // one test per tile on the common path
#pragma unroll
                    for (int j = 0; j < R; j++) {
                        if (any_miss && miss[j]) {
                            // global-table path: the directory is full of other groups (or the shape keeps none)
                            int rc;
                            if constexpr (S::GMAX > 0) { // the exception here: keep its code out of the loop body
                                rc = global_row_cold<S>(&prog, tv.slots, tv.mask, tv.slot_bytes, tv.key_words, k[j].lo, k[j].hi, k[j].len,
                                                        pack_agg_args<S>(v[j], AggSeq{}), AggSeq{});
                            } else {
                                bool fresh;
                                uint8_t* slot =
                                        table_find_or_insert(tv, k[j].lo, k[j].hi, hash_packed_key_static<static_key_bytes<S>()>(k[j], p.phase_seed), &fresh);
                                rc = (fresh ? 1 : 0) | (slot ? 0 : 2);
                                if (slot) row_update_global<S>(p, slot, v[j], AggSeq{});
                            }
                            inserted += rc & 1;
                            if (rc & 2) {
                                const int64_t i = (int64_t)tile * TILE + tid + j * THREADS;
                                atomicOr(&defer_out[i >> 5], 1u << (i & 31));
                                deferred++;
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive_a(empty0 + 8u * (uint32_t)s); // this warp is done with stage s
            stage += stage_bytes;
            if (++s == n_stages) {
                s = 0;
                parity ^= 1u;
                stage = smem;
            }
        }
    }
    __syncthreads(); // the directory is final and the ring is idle: its first bytes become the stash
    if constexpr (S::GMAX > 0) {
        static_assert(Flusher<S>::STASH_BYTES <= 32 * 1024, "stash must fit the first stage");
        uint64_t* stash = (uint64_t*)smem;
        const int cnt = S::N_KEYS == 0 ? 1 : dir.cnt;
        if (tid < THREADS)
            for (int q = 0; q < cnt; q++) Flusher<S>::stash(ra, acc_tid, rows_tid, q, stash, AggSeq{});
        __syncthreads();
        if (tid < cnt) Flusher<S>::fold(p, stash, tid, S::N_KEYS == 0 ? tv.slots : dir.slot[tid], AggSeq{});
    }
    if (tid >= THREADS) return;
    const unsigned long long selected = warp_sum_u64(selected32);
    deferred = warp_sum_u64(deferred);
    inserted = warp_sum_u64(inserted);
    if ((tid & 31) == 0) {
        if (selected) atomicAdd(ctr.rows_selected, selected);
        if (deferred) atomicAdd(ctr.deferred, deferred);
        if (inserted) atomicAdd(ctr.inserted, inserted);
    }
}


// What the host must know about a shape to lay out a stage and launch its kernel. Built-in shapes fill it from their templates; a shape
// compiled at run time (srb_jit.h) exports the same struct as a device constant of its module, so the two can never disagree.
struct ShapeMeta {
    int32_t n_nodes, threads, rows, tile, acc_bytes, has_narrow, gmax, _pad;
    int32_t load_width[MAX_NODES]; // bytes of the column a LOAD node reads, 0 for every other node
    int32_t load_off[MAX_NODES];   // its offset inside a stage
    int32_t load_nullable[MAX_NODES]; // the kernel reads a null byte per row for this LOAD node
    int32_t load_binary[MAX_NODES];   // a BinaryColumn: offsets + byte window (load_width 0)
};
template <class S>
constexpr ShapeMeta shape_meta() {
    ShapeMeta m{};
    m.n_nodes = S::N_NODES;
    m.threads = S::THREADS;
    m.rows = S::ROWS;
    m.tile = tile_rows<S>();
    m.acc_bytes = AccGeom<S>::BYTES;
    m.has_narrow = shape_has_narrow<S>() ? 1 : 0;
    m.gmax = S::GMAX;
    for (int i = 0; i < S::N_NODES; i++) {
        m.load_width[i] = S::node(i).op == DOP_LOAD ? S::node(i).width : 0;
        m.load_off[i] = load_off<S>(i);
        m.load_nullable[i] = (S::node(i).op == DOP_LOAD && S::node(i).nullable != 0) ? 1 : 0;
        m.load_binary[i] = (S::node(i).op == DOP_LOAD && S::node(i).width == 0) ? 1 : 0;
    }
    return m;
}

#if !defined(__CUDACC_RTC__) // host side: shape registry, predicate lowering, launch
// ---- registry ---------------------------------------------------------------------------------------------------------------
inline int match_static_shape(const DProg& p) {
    if (shape_matches<ShapeSumMulFast>(p)) return 0;
    if (shape_matches<ShapeQ1Dict>(p)) return 1;
    if (shape_matches<ShapeHighCard>(p)) return 2;
    if (shape_matches<ShapeQ3Agg>(p)) return 3;
    if (shape_matches<ShapeHighCardMerge>(p)) return 4;
    return -1;
}
inline const char* static_shape_name(int id) {
    switch (id) {
    case 0: return ShapeSumMulFast::name;
    case 1: return ShapeQ1Dict::name;
    case 2: return ShapeHighCard::name;
    case 3: return ShapeQ3Agg::name;
    case 4: return ShapeHighCardMerge::name;
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
    case 10: return 0b0111; // NOT NULL: every non-NULL value
    default: return 0b1000; // IS NULL: bit 3 = the answer for a NULL (spred_eval), no non-NULL value
    }
}

// Lower conjunct predicates over 4- / 8-byte integer columns to per-column ranges. Returns false (caller keeps the
// generic predicate loop) for OR-groups, other widths, or more distinct ranges than ST_MAX_RANGES.
inline bool lower_ranges(const DProg& p, StaticArgs& sa) {
    int n = 0;
    for (int i = 0; i < p.n_preds; i++) {
        const DPred& d = p.preds[i];
        const int w = p.cols[d.col].width;
        if (d.group != 0 || (w != 4 && w != 8) || sa.null_off[d.col] >= 0) return false;
        const i128 tmin = w == 4 ? (i128)INT32_MIN : (i128)INT64_MIN, tmax = w == 4 ? (i128)INT32_MAX : (i128)INT64_MAX;
        const i128 lit = make_i128(d.lit_lo, d.lit_hi);
        i128 lo = tmin, hi = tmax; // inclusive bounds
        bool neg = false;
        switch (d.op) {
        case SRB_PRED_EQ: lo = hi = lit; break;
        case SRB_PRED_NE: lo = hi = lit; neg = true; break;
        case SRB_PRED_GT: lo = lit + 1; break;
        case SRB_PRED_GE: lo = lit; break;
        case SRB_PRED_LT: hi = lit - 1; break;
        case SRB_PRED_LE: hi = lit; break;
        case SRB_PRED_NOT_NULL: continue;  // NULL-free column: always true
        default: neg = true; break;        // IS NULL on a NULL-free column: never true (NOT of the full range)
        }
        if (lo < tmin) lo = tmin;
        if (hi > tmax) hi = tmax;
        if (lo > hi) { // empty range: NOT(full) when plain, full when negated
            lo = tmin;
            hi = tmax;
            neg = !neg;
        }
        // intersect with an earlier plain range on the same column
        bool merged = false;
        if (!neg) {
            for (int k = 0; k < n && !merged; k++) {
                SRange& g = sa.rng[k];
                if (g.col_off != sa.col_off[d.col] || g.neg) continue;
                i128 glo = (i128)g.lo, ghi = (i128)g.lo + (i128)g.span;
                if (lo > glo) glo = lo;
                if (hi < ghi) ghi = hi;
                if (glo > ghi) { // empty intersection
                    g.lo = (int64_t)tmin;
                    g.span = (uint64_t)(tmax - tmin);
                    g.neg = 1;
                } else {
                    g.lo = (int64_t)glo;
                    g.span = (uint64_t)(ghi - glo);
                }
                merged = true;
            }
        }
        if (merged) continue;
        if (n == ST_MAX_RANGES) return false;
        sa.rng[n++] = SRange{sa.col_off[d.col], w, neg ? 1 : 0, 0, (int64_t)lo, (uint64_t)(hi - lo)};
    }
    sa.n_ranges = n;
    return true;
}

// uniform part of the narrow body's assumption, checked once on the host: literals in [0, 2^32), casts that do not rescale (the program's
// nodes are the shape's nodes: shape_matches / the generator guarantee it)
inline bool narrow_literals_ok_rt(const ShapeMeta& m, const DProg& p) {
    if (!m.has_narrow) return false;
    for (int i = 0; i < p.n_nodes; i++) {
        const DNode& rt = p.nodes[i];
        if (rt.op == DOP_LIT && !(rt.lit_hi == 0 && (rt.lit_lo >> 32) == 0)) return false;
        if (rt.op == DOP_CAST && !(rt.lit_hi == 0 && rt.lit_lo == 1)) return false;
    }
    return true;
}

// Stage layout and kernel arguments of one launch of a static kernel over the leading whole tiles of [base, base+n): the number of tiles,
// 0 when the preconditions (no null maps, 16-byte aligned columns, every column bound once, two stages beside the accumulators) do not hold.
inline int64_t prepare_static(const ShapeMeta& m, const DProg& p, int64_t base, int64_t n, bool allow_narrow, StaticArgs& sa, size_t& smem) {
    // the columns of the LOAD nodes first, in node order (compile-time offsets), then predicate-only columns
    bool placed[MAX_COLS] = {};
    int node_of[MAX_COLS];
    int bytes = 0;
    bool any_binary = false;
    for (int c = 0; c < MAX_COLS; c++) {
        node_of[c] = -1;
        sa.null_off[c] = -1;
        sa.bin_win_off[c] = -1;
        sa.bin_hdr_off[c] = -1;
    }
    for (int i = 0; i < MAX_NODES; i++) sa.node_null_off[i] = -1;
    for (int i = 0; i < m.n_nodes; i++) {
        if (!m.load_width[i] && !m.load_binary[i]) continue;
        const int c = p.nodes[i].col;
        if (c < 0 || c >= p.n_cols || placed[c] || p.cols[c].width != m.load_width[i]) return 0; // a column bound twice: generic kernel
        if ((p.cols[c].kind == CK_BINARY) != (m.load_binary[i] != 0)) return 0;
        placed[c] = true;
        node_of[c] = i;
        sa.col_off[c] = m.load_off[i];
        bytes = m.load_off[i] + (m.load_binary[i] ? 4 * m.tile + 16 : m.load_width[i] * m.tile);
    }
    for (int c = 0; c < p.n_cols; c++) {
        if (p.cols[c].kind == CK_BINARY) {
            // offsets and bytes both start 16-byte aligned (bulk copies); the bytes keep the reference's 16 bytes of tail padding
            // (column/bytes.h:27): the last window is rounded up to 16
            if (!placed[c] || !p.cols[c].offsets || (((uintptr_t)p.cols[c].offsets + (size_t)base * 4) & 15) || ((uintptr_t)p.cols[c].data & 15)) return 0;
            any_binary = true;
            sa.bin_win_off[c] = bytes;
            bytes += bin_window_bytes(m.tile);
            sa.bin_hdr_off[c] = bytes;
            bytes += 16;
        } else {
            if (!placed[c]) {
                sa.col_off[c] = bytes;
                bytes += p.cols[c].width * m.tile;
            }
            // TMA bulk copies need 16-byte aligned sources: tile starts are multiples of 16 rows, so the base pointers decide
            if (((uintptr_t)p.cols[c].data + (size_t)base * p.cols[c].width) & 15) return 0;
        }
        // null maps: the kernel was compiled for exactly the LOAD columns that carry one (predicate-only columns may or may not)
        const bool compiled_nullable = node_of[c] >= 0 && m.load_nullable[node_of[c]];
        if (node_of[c] >= 0 && compiled_nullable != (p.cols[c].nulls != nullptr)) return 0;
        if (p.cols[c].nulls) {
            if (((uintptr_t)p.cols[c].nulls + (size_t)base) & 15) return 0;
            sa.null_off[c] = bytes;
            if (node_of[c] >= 0) sa.node_null_off[node_of[c]] = bytes;
            bytes += m.tile;
        }
    }
    int64_t n_tiles = n / m.tile;
    if (any_binary && n_tiles * m.tile + 3 > n) n_tiles--; // a tile copies tile + 4 offsets (n + 1 exist): the last one must leave three rows behind it
    if (n_tiles <= 0 || bytes == 0) return 0;
    sa.stage_bytes = bytes;
    int stages = (ST_SMEM_BUDGET - m.acc_bytes) / sa.stage_bytes;
    if (stages > ST_MAX_STAGES) stages = ST_MAX_STAGES;
    if (stages < 2) return 0;
    sa.n_stages = stages;
    sa.acc_off = stages * sa.stage_bytes;
    sa.n_preds = p.n_preds;
    for (int i = 0; i < p.n_preds; i++) {
        const DPred& d = p.preds[i];
        sa.preds[i] = SPred{sa.col_off[d.col], p.cols[d.col].width, pred_opmask(d.op), d.group, sa.null_off[d.col], 0, d.lit_lo, d.lit_hi};
    }
    if (!lower_ranges(p, sa)) sa.n_ranges = -1;
    sa.narrow_ok = (allow_narrow && narrow_literals_ok_rt(m, p)) ? 1 : 0; // allow_narrow = 0: A/B runs and tests of the general body (SRB_NO_NARROW)
    smem = (size_t)sa.acc_off + m.acc_bytes;
    return n_tiles;
}

template <class S>
inline int64_t launch_shape(const DProg& p, const TableView& tv, int64_t base, int64_t n, uint32_t* defer_out, unsigned long long* rows_selected,
                            unsigned long long* deferred, unsigned long long* inserted, int n_sms, cudaStream_t stream, bool allow_narrow) {
    constexpr ShapeMeta m = shape_meta<S>();
    StaticArgs sa{};
    size_t smem = 0;
    const int64_t n_tiles = prepare_static(m, p, base, n, allow_narrow, sa, smem);
    if (n_tiles == 0) return 0;
    static SmemOptIn opt_in; // per (shape, device)
    opt_in.set(k_scan_agg_static<S>, ST_SMEM_BUDGET);
    const int grid = (int)(n_tiles < n_sms ? n_tiles : n_sms);
    StaticCtr ctr{rows_selected, deferred, inserted};
    k_scan_agg_static<S><<<grid, S::THREADS + 32, smem, stream>>>(p, sa, tv, base, n_tiles, defer_out, ctr);
    return n_tiles * m.tile;
}


// Launches the static kernel of `shape` over the leading whole tiles of [base, base+n); returns the number of rows it
// covered (0 when the preconditions — no null maps, 16-byte aligned columns, at least one tile — do not hold). The caller
// runs the generic kernel over the remaining rows.
inline int64_t launch_static_shape(int shape, const DProg& p, const TableView& tv, int64_t base, int64_t n, uint32_t* defer_out,
                                   unsigned long long* rows_selected, unsigned long long* deferred, unsigned long long* inserted, int n_sms,
                                   cudaStream_t stream, bool allow_narrow = true) {
    switch (shape) {
    case 0: return launch_shape<ShapeSumMulFast>(p, tv, base, n, defer_out, rows_selected, deferred, inserted, n_sms, stream, allow_narrow);
    case 1: return launch_shape<ShapeQ1Dict>(p, tv, base, n, defer_out, rows_selected, deferred, inserted, n_sms, stream, allow_narrow);
    case 2: return launch_shape<ShapeHighCard>(p, tv, base, n, defer_out, rows_selected, deferred, inserted, n_sms, stream, allow_narrow);
    case 3: return launch_shape<ShapeQ3Agg>(p, tv, base, n, defer_out, rows_selected, deferred, inserted, n_sms, stream, allow_narrow);
    case 4: return launch_shape<ShapeHighCardMerge>(p, tv, base, n, defer_out, rows_selected, deferred, inserted, n_sms, stream, allow_narrow);
    }
    return 0;
}
#endif // !__CUDACC_RTC__

} // namespace srb
