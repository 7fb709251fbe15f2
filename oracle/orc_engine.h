// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_arith.h header).
// Chunk-at-a-time column model + predicate / expression evaluation, restating:
//   ColumnPredicateCmpBase::t_evaluate            be/src/storage/column_predicate_cmp.cpp:222-239
//   ConjunctivePredicates::evaluate / evaluate_or  be/src/storage/conjunctive_predicates.cpp:35-63
//   Chunk::filter / ColumnHelper::t_filter_range   be/src/column/chunk.cpp:302, column/column_helper.h:468-577
//   VectorizedArithmeticExpr::evaluate_checked      be/src/exprs/arithmetic_expr.cpp:70-128
//   DecimalBinaryFunction::evaluate                 be/src/exprs/decimal_binary_function.h:47-213
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/srb/srb_c_api.h"
#include "orc_arith.h"
#include "orc_hash.h"

namespace orc {

inline int type_width(int t) {
    switch (t) {
    case SRB_TYPE_TINYINT:
    case SRB_TYPE_BOOLEAN: return 1;
    case SRB_TYPE_SMALLINT: return 2;
    case SRB_TYPE_INT:
    case SRB_TYPE_DATE:
    case SRB_TYPE_DECIMAL32: return 4;
    case SRB_TYPE_BIGINT:
    case SRB_TYPE_DECIMAL64:
    case SRB_TYPE_DOUBLE: return 8;
    case SRB_TYPE_LARGEINT:
    case SRB_TYPE_DECIMAL128: return 16;
    default: return 0; // binary
    }
}
inline bool is_decimal(int t) { return t == SRB_TYPE_DECIMAL32 || t == SRB_TYPE_DECIMAL64 || t == SRB_TYPE_DECIMAL128; }
inline bool is_binary(int t) { return t == SRB_TYPE_VARCHAR || t == SRB_TYPE_VARBINARY; }

// An evaluated column for one chunk. Integer-like values are widened to int128 (exact); doubles kept apart;
// binary values kept as strings.
struct Col {
    int type = 0, precision = 0, scale = 0;
    bool nullable = false; // has a null map
    std::vector<int128_t> v;
    std::vector<double> d;
    std::vector<std::string> s;
    std::vector<uint8_t> nulls; // 1 = null; sized n when nullable
    size_t size() const { return type == SRB_TYPE_DOUBLE ? d.size() : (is_binary(type) ? s.size() : v.size()); }
    bool is_null(size_t i) const { return nullable && nulls[i]; }
};

inline int128_t load_int(const srb_column& c, int64_t row) {
    const uint8_t* p = (const uint8_t*)c.data;
    switch (type_width(c.type)) {
    case 1: return ((const int8_t*)p)[row];
    case 2: return ((const int16_t*)p)[row];
    case 4: return ((const int32_t*)p)[row];
    case 8: return ((const int64_t*)p)[row];
    case 16: {
        int128_t x;
        memcpy(&x, p + 16 * row, 16);
        return x;
    }
    }
    throw std::runtime_error("load_int: bad type");
}

inline const srb_column* find_slot(const srb_chunk& ch, int slot) {
    for (int i = 0; i < ch.n_cols; i++)
        if (ch.slot_ids[i] == slot) return &ch.cols[i];
    throw std::runtime_error("slot " + std::to_string(slot) + " not in chunk");
}

// ---- predicates -> byte selection -----------------------------------------------------------------
inline bool cmp_apply(int op, int128_t a, int128_t b) {
    switch (op) {
    case SRB_PRED_EQ: return a == b;
    case SRB_PRED_NE: return a != b;
    case SRB_PRED_GT: return a > b;
    case SRB_PRED_GE: return a >= b;
    case SRB_PRED_LT: return a < b;
    case SRB_PRED_LE: return a <= b;
    }
    throw std::runtime_error("bad predicate op");
}
inline bool cmp_apply_d(int op, double a, double b) {
    switch (op) {
    case SRB_PRED_EQ: return a == b;
    case SRB_PRED_NE: return a != b;
    case SRB_PRED_GT: return a > b;
    case SRB_PRED_GE: return a >= b;
    case SRB_PRED_LT: return a < b;
    case SRB_PRED_LE: return a <= b;
    }
    throw std::runtime_error("bad predicate op");
}

// one predicate over rows [from,to) of the chunk window starting at `base`; mode 0 = assign, 1 = and, 2 = or
inline void pred_evaluate(const srb_predicate& p, const srb_chunk& ch, int64_t base, uint8_t* sel, size_t from, size_t to,
                          int mode) {
    const srb_column* c = find_slot(ch, p.slot_id);
    int128_t lit = ((int128_t)p.lit_hi << 64) | (uint64_t)p.lit_lo;
    double dlit;
    memcpy(&dlit, &p.lit_lo, 8);
    for (size_t i = from; i < to; i++) {
        int64_t row = base + (int64_t)i;
        bool isnull = c->nulls && c->nulls[row];
        uint8_t r;
        if (p.op == SRB_PRED_IS_NULL)
            r = isnull;
        else if (p.op == SRB_PRED_NOT_NULL)
            r = !isnull;
        else if (is_binary(c->type)) {
            // BinaryColumnPredicateCmpBase (storage/column_predicate_cmp.cpp:559-700): Slice::compare = memcmp over the common length, then lengths
            r = 0;
            if (!isnull) {
                const uint8_t* sp = (const uint8_t*)c->data + c->offsets[row];
                const size_t n = c->offsets[row + 1] - c->offsets[row], m = (size_t)p.lit_len;
                int d = memcmp(sp, p.lit_bytes, n < m ? n : m);
                if (d == 0) d = n == m ? 0 : (n < m ? -1 : 1);
                r = (uint8_t)cmp_apply(p.op, (int128_t)d, (int128_t)0);
            }
        } else if (c->type == SRB_TYPE_DOUBLE)
            r = (uint8_t)((!isnull) & cmp_apply_d(p.op, ((const double*)c->data)[row], dlit));
        else
            r = (uint8_t)((!isnull) & cmp_apply(p.op, load_int(*c, row), lit));
        if (mode == 0)
            sel[i] = r;
        else if (mode == 1)
            sel[i] = sel[i] & r;
        else
            sel[i] = sel[i] | r;
    }
}

// AND of plain conjuncts and OR-groups, in plan order; first predicate assigns (conjunctive_predicates.cpp:35-52)
inline void eval_predicates(const srb_predicate* preds, int n, const srb_chunk& ch, int64_t base, size_t rows,
                            std::vector<uint8_t>& sel) {
    sel.assign(rows, 1);
    std::vector<uint8_t> tmp;
    std::vector<bool> done(n, false);
    bool first = true;
    for (int i = 0; i < n; i++) {
        if (done[i]) continue;
        if (preds[i].or_group == 0) {
            pred_evaluate(preds[i], ch, base, sel.data(), 0, rows, first ? 0 : 1);
        } else {
            tmp.assign(rows, 0);
            for (int j = i; j < n; j++)
                if (preds[j].or_group == preds[i].or_group) {
                    pred_evaluate(preds[j], ch, base, tmp.data(), 0, rows, 2);
                    done[j] = true;
                }
            for (size_t r = 0; r < rows; r++) sel[r] = first ? tmp[r] : (sel[r] & tmp[r]);
        }
        first = false;
    }
}

// ---- compaction: gather one input column for the selected rows (Chunk::filter on that column) --------
inline Col gather_column(const srb_column& c, int64_t base, const std::vector<uint32_t>& idx) {
    Col out;
    out.type = c.type;
    out.precision = c.precision;
    out.scale = c.scale;
    out.nullable = c.nulls != nullptr;
    size_t n = idx.size();
    if (out.nullable) out.nulls.resize(n);
    if (c.type == SRB_TYPE_DOUBLE) {
        out.d.resize(n);
        for (size_t i = 0; i < n; i++) out.d[i] = ((const double*)c.data)[base + idx[i]];
    } else if (is_binary(c.type)) {
        out.s.resize(n);
        for (size_t i = 0; i < n; i++) {
            int64_t r = base + idx[i];
            out.s[i].assign((const char*)c.data + c.offsets[r], c.offsets[r + 1] - c.offsets[r]);
        }
    } else {
        out.v.resize(n);
        for (size_t i = 0; i < n; i++) out.v[i] = load_int(c, base + idx[i]);
    }
    if (out.nullable)
        for (size_t i = 0; i < n; i++) out.nulls[i] = c.nulls[base + idx[i]];
    return out;
}

// ---- expressions -------------------------------------------------------------------------------------
template <typename T>
inline bool checked_binop(int op, int128_t a, int128_t b, int128_t* out) {
    T r;
    bool ov;
    if (op == SRB_EXPR_ADD)
        ov = add_overflow((T)a, (T)b, &r);
    else if (op == SRB_EXPR_SUB)
        ov = sub_overflow((T)a, (T)b, &r);
    else
        ov = mul_overflow((T)a, (T)b, &r);
    *out = (int128_t)r;
    return ov;
}

inline int128_t wrap_to(int type, int128_t x) {
    switch (type_width(type)) {
    case 1: return (int8_t)x;
    case 2: return (int16_t)x;
    case 4: return (int32_t)x;
    case 8: return (int64_t)x;
    }
    return x;
}

struct ExprEval {
    const srb_expr_node* nodes;
    int n_nodes;
    const srb_chunk* chunk;
    int64_t base;
    const std::vector<uint32_t>* idx;
    std::vector<std::unique_ptr<Col>> memo;

    ExprEval(const srb_expr_node* nd, int n, const srb_chunk* ch, int64_t b, const std::vector<uint32_t>* ix)
            : nodes(nd), n_nodes(n), chunk(ch), base(b), idx(ix), memo(n) {}

    // a CAST whose child is decimal with the same scale: eliminate_trivial_cast_for_decimal_mul (arithmetic_expr.cpp:52-64)
    int trivial_cast_child_type(int ni) const {
        const srb_expr_node& e = nodes[ni];
        if (e.op != SRB_EXPR_CAST) return 0;
        const srb_expr_node& c = nodes[e.lhs];
        if (is_decimal(e.type) && is_decimal(c.type) && e.scale == c.scale) return c.type;
        return 0;
    }

    const Col& eval(int ni) {
        if (memo[ni]) return *memo[ni];
        const srb_expr_node& e = nodes[ni];
        auto out = std::make_unique<Col>();
        size_t n = idx->size();
        out->type = e.type;
        out->precision = e.precision;
        out->scale = e.scale;
        switch (e.op) {
        case SRB_EXPR_SLOT_REF: {
            *out = gather_column(*find_slot(*chunk, e.slot_id), base, *idx);
            break;
        }
        case SRB_EXPR_LITERAL: {
            int128_t lit = ((int128_t)e.lit_hi << 64) | (uint64_t)e.lit_lo;
            if (e.type == SRB_TYPE_DOUBLE) {
                double dl;
                memcpy(&dl, &e.lit_lo, 8);
                out->d.assign(n, dl);
            } else
                out->v.assign(n, lit);
            break;
        }
        case SRB_EXPR_CAST: {
            const Col& c = eval(e.lhs);
            out->nullable = c.nullable;
            out->nulls = c.nulls;
            int up = e.scale - c.scale;
            if (up < 0) throw std::runtime_error("cast: scale down not supported");
            if (c.type == SRB_TYPE_DOUBLE || e.type == SRB_TYPE_DOUBLE) throw std::runtime_error("cast: double not supported");
            out->v.resize(n);
            int128_t f = exp10_128(up);
            bool any_null = false;
            std::vector<uint8_t> nl(n, 0);
            for (size_t i = 0; i < n; i++) {
                int128_t r;
                bool ov = false;
                // DecimalV3Cast::to_decimal<From,To,To,true,check> -> scale_up in the target type (decimalv3.h:240-256)
                if (type_width(e.type) == 16)
                    ov = up ? checked_binop<int128_t>(SRB_EXPR_MUL, c.v[i], f, &r) : (r = c.v[i], false);
                else if (type_width(e.type) == 8) {
                    ov = up ? checked_binop<int64_t>(SRB_EXPR_MUL, c.v[i], f, &r) : (r = c.v[i], false);
                    ov = ov || c.v[i] > INT64_MAX || c.v[i] < INT64_MIN;
                } else {
                    ov = up ? checked_binop<int32_t>(SRB_EXPR_MUL, c.v[i], f, &r) : (r = c.v[i], false);
                    ov = ov || c.v[i] > INT32_MAX || c.v[i] < INT32_MIN;
                }
                out->v[i] = r;
                if (ov && is_decimal(e.type)) {
                    nl[i] = 1;
                    any_null = true;
                }
            }
            if (any_null) {
                if (!out->nullable) {
                    out->nullable = true;
                    out->nulls.assign(n, 0);
                }
                for (size_t i = 0; i < n; i++) out->nulls[i] |= nl[i];
            }
            break;
        }
        case SRB_EXPR_ADD:
        case SRB_EXPR_SUB:
        case SRB_EXPR_MUL: {
            bool fast_mul = false;
            int li = e.lhs, ri = e.rhs;
            if (e.op == SRB_EXPR_MUL && e.type == SRB_TYPE_DECIMAL128) {
                int lt = trivial_cast_child_type(li), rt = trivial_cast_child_type(ri);
                if ((lt == SRB_TYPE_DECIMAL64 || lt == SRB_TYPE_DECIMAL32) &&
                    (rt == SRB_TYPE_DECIMAL64 || rt == SRB_TYPE_DECIMAL32)) {
                    fast_mul = true; // children evaluated below the cast (arithmetic_expr.cpp:76-78)
                    li = nodes[li].lhs;
                    ri = nodes[ri].lhs;
                }
            }
            const Col& a = eval(li);
            const Col& b = eval(ri);
            if (a.type == SRB_TYPE_DOUBLE || b.type == SRB_TYPE_DOUBLE) throw std::runtime_error("double arithmetic not supported");
            out->nullable = a.nullable || b.nullable;
            out->v.resize(n);
            if (out->nullable) {
                out->nulls.assign(n, 0);
                for (size_t i = 0; i < n; i++) out->nulls[i] = (uint8_t)(a.is_null(i) | b.is_null(i));
            }
            if (fast_mul) {
                for (size_t i = 0; i < n; i++) out->v[i] = i64_x_i64_produce_i128((int64_t)a.v[i], (int64_t)b.v[i]);
            } else if (is_decimal(e.type)) {
                // scale adjustment for add/sub (decimal_binary_function.h:142-170): the lower-scale side is scaled up, checked
                int adj_a = 0, adj_b = 0;
                if (e.op != SRB_EXPR_MUL) {
                    int s = a.scale > b.scale ? a.scale : b.scale;
                    adj_a = s - a.scale;
                    adj_b = s - b.scale;
                }
                int128_t fa = exp10_128(adj_a), fb = exp10_128(adj_b);
                bool any_null = false;
                std::vector<uint8_t> nl(n, 0);
                int w = type_width(e.type);
                for (size_t i = 0; i < n; i++) {
                    int128_t x = a.v[i], y = b.v[i], r = 0;
                    bool ov = false;
                    auto bin = [&](int op, int128_t p, int128_t q, int128_t* o) {
                        return w == 16 ? checked_binop<int128_t>(op, p, q, o)
                                       : (w == 8 ? checked_binop<int64_t>(op, p, q, o) : checked_binop<int32_t>(op, p, q, o));
                    };
                    if (adj_a) ov |= bin(SRB_EXPR_MUL, x, fa, &x);
                    if (adj_b) ov |= bin(SRB_EXPR_MUL, y, fb, &y);
                    if (!ov) ov = bin(e.op, x, y, &r);
                    out->v[i] = r;
                    if (ov) {
                        nl[i] = 1;
                        any_null = true;
                    }
                }
                if (any_null) {
                    if (!out->nullable) {
                        out->nullable = true;
                        out->nulls.assign(n, 0);
                    }
                    for (size_t i = 0; i < n; i++) out->nulls[i] |= nl[i];
                }
            } else {
                // plain integer arithmetic wraps in the result type (arithmetic_operation.h:127-141)
                for (size_t i = 0; i < n; i++) {
                    uint128_t x = (uint128_t)a.v[i], y = (uint128_t)b.v[i], r;
                    r = e.op == SRB_EXPR_ADD ? x + y : (e.op == SRB_EXPR_SUB ? x - y : x * y);
                    out->v[i] = wrap_to(e.type, (int128_t)r);
                }
            }
            break;
        }
        default: throw std::runtime_error("bad expr op");
        }
        out->type = e.type;
        out->precision = e.precision;
        out->scale = e.scale;
        memo[ni] = std::move(out);
        return *memo[ni];
    }
};

} // namespace orc
