// srb_lower.h — host-side lowering of an srb_scan_agg_plan into the device program (srb_prog.h).
//
// Typing rules restated from the reference:
//   fast-mul detection (both children same-scale casts of DECIMAL32/64 into DECIMAL128)  be/src/exprs/arithmetic_expr.cpp:52-101
//   add/sub scale = max, mul scale = sum; the lower-scale side is scaled up (checked)     be/src/exprs/arithmetic_operation.h:391-427,
//                                                                                         be/src/exprs/decimal_binary_function.h:142-170
//   SUM result: any decimal -> DECIMAL128; ints -> BIGINT; LARGEINT; DOUBLE              be/src/exprs/agg/sum.h:30-31,176-181
//   AVG result: decimal -> DECIMAL128 scale per decimal_div_integer; DATE; else DOUBLE   be/src/exprs/agg/avg.h:36-53,218-245
//   MIN/MAX keep the argument type                                                        be/src/exprs/agg/maxmin.h:30-122
#pragma once
#include <cstring>
#include <map>
#include <vector>

#include "srb_host.h"
#include "srb_prog.h"

namespace srb {

struct ColMeta {
    int slot, type, precision, scale;
    bool nullable;
};

struct AggOutMeta {
    int fn;
    int arg_type, arg_scale;  // original argument type (through the blob for merges)
    int out_type, out_precision, out_scale;
    int out_width;   // bytes per value of the device result column (blob: width of the sum part)
    bool nullable;   // result column carries a null map
    bool blob;       // AVG in a serialize phase: VARBINARY {sum,count}
    bool blob_double;
};

struct Lowered {
    DProg prog;
    std::vector<ColMeta> cols;            // bound column order
    std::vector<AggOutMeta> agg_out;
    std::vector<ColMeta> key_out;         // one per key
    bool finalize = true;
    bool merge = false;
};

inline void agg_result_type(int fn, int arg_type, int arg_scale, int* type, int* precision, int* scale) {
    *precision = 0;
    *scale = 0;
    switch (fn) {
    case SRB_AGG_COUNT: *type = SRB_TYPE_BIGINT; return;
    case SRB_AGG_SUM:
        if (is_decimal(arg_type)) {
            *type = SRB_TYPE_DECIMAL128;
            *precision = 38;
            *scale = arg_scale;
        } else if (arg_type == SRB_TYPE_DOUBLE)
            *type = SRB_TYPE_DOUBLE;
        else if (arg_type == SRB_TYPE_LARGEINT)
            *type = SRB_TYPE_LARGEINT;
        else
            *type = SRB_TYPE_BIGINT;
        return;
    case SRB_AGG_AVG:
        if (is_decimal(arg_type)) {
            *type = SRB_TYPE_DECIMAL128;
            *precision = 38;
            *scale = div_result_scale(arg_scale);
        } else if (arg_type == SRB_TYPE_DATE)
            *type = SRB_TYPE_DATE;
        else
            *type = SRB_TYPE_DOUBLE;
        return;
    default:
        *type = arg_type;
        *scale = arg_scale;
        return;
    }
}
inline bool avg_in_double(int t) { return !(is_decimal(t) || t == SRB_TYPE_DATE || t == SRB_TYPE_LARGEINT); }

struct PlanCopy {
    std::vector<srb_predicate> preds;
    std::vector<srb_expr_node> nodes;
    std::vector<int32_t> key_nodes;
    std::vector<srb_agg_fn> aggs;
    std::vector<int32_t> out_slots;
    int phase = 0;
    int64_t group_hint = 0;
    explicit PlanCopy(const srb_scan_agg_plan* p) {
        if (p->n_preds < 0 || p->n_nodes < 0 || p->n_keys < 0 || p->n_aggs < 0 || p->n_out_slots < 0)
            throw Error(SRB_INVALID_ARGUMENT, "negative count in plan");
        preds.assign(p->preds, p->preds + p->n_preds);
        nodes.assign(p->nodes, p->nodes + p->n_nodes);
        key_nodes.assign(p->key_nodes, p->key_nodes + p->n_keys);
        aggs.assign(p->aggs, p->aggs + p->n_aggs);
        out_slots.assign(p->out_slots, p->out_slots + p->n_out_slots);
        phase = p->phase;
        group_hint = p->group_hint;
        if ((int)preds.size() > MAX_PREDS) throw Error(SRB_NOT_IMPLEMENTED, fmt("more than %d predicates", MAX_PREDS));
        if ((int)nodes.size() > MAX_NODES) throw Error(SRB_NOT_IMPLEMENTED, fmt("more than %d expression nodes", MAX_NODES));
        if ((int)key_nodes.size() > MAX_KEYS) throw Error(SRB_NOT_IMPLEMENTED, fmt("more than %d group keys", MAX_KEYS));
        if ((int)aggs.size() > MAX_AGGS) throw Error(SRB_NOT_IMPLEMENTED, fmt("more than %d aggregates", MAX_AGGS));
        for (size_t i = 0; i < nodes.size(); i++) {
            const srb_expr_node& e = nodes[i];
            auto bad = [&](int c) { return c < 0 || c >= (int)i; };
            if (e.op == SRB_EXPR_CAST && bad(e.lhs)) throw Error(SRB_INVALID_ARGUMENT, "expr node references a later node");
            if (e.op >= SRB_EXPR_ADD && e.op <= SRB_EXPR_MUL && (bad(e.lhs) || bad(e.rhs)))
                throw Error(SRB_INVALID_ARGUMENT, "expr node references a later node");
            if (e.op < SRB_EXPR_SLOT_REF || e.op > SRB_EXPR_MUL) throw Error(SRB_INVALID_ARGUMENT, "bad expr op");
        }
        for (auto k : key_nodes)
            if (k < 0 || k >= (int)nodes.size() || nodes[k].op != SRB_EXPR_SLOT_REF)
                throw Error(SRB_INVALID_ARGUMENT, "group-by keys must be SLOT_REF nodes");
        for (auto& a : aggs) {
            if (a.fn < SRB_AGG_SUM || a.fn > SRB_AGG_MAX) throw Error(SRB_INVALID_ARGUMENT, "bad aggregate function");
            if (a.arg >= (int)nodes.size()) throw Error(SRB_INVALID_ARGUMENT, "aggregate argument out of range");
            if (a.arg < 0 && a.fn != SRB_AGG_COUNT) throw Error(SRB_INVALID_ARGUMENT, "only count(*) takes no argument");
        }
        for (auto& p2 : preds) {
            bool ok = (p2.op >= SRB_PRED_EQ && p2.op <= SRB_PRED_LE) || p2.op == SRB_PRED_IS_NULL || p2.op == SRB_PRED_NOT_NULL;
            if (!ok) throw Error(SRB_NOT_IMPLEMENTED, fmt("predicate op %d", p2.op));
        }
        if (phase < 0 || phase > 3) throw Error(SRB_INVALID_ARGUMENT, "bad aggregation phase");
    }
};

inline const srb_column* find_slot(const srb_chunk* ch, int slot) {
    for (int i = 0; i < ch->n_cols; i++)
        if (ch->slot_ids[i] == slot) return &ch->cols[i];
    return nullptr;
}

// Lower `plan` against the column shapes of `first` (the first pushed chunk). Data pointers are NOT bound here.
inline Lowered lower_plan(const PlanCopy& plan, const srb_chunk* first) {
    Lowered L;
    memset(&L.prog, 0, sizeof(L.prog));
    DProg& P = L.prog;
    L.finalize = plan.phase == SRB_AGG_UPDATE_FINALIZE || plan.phase == SRB_AGG_MERGE_FINALIZE;
    L.merge = plan.phase >= SRB_AGG_MERGE_FINALIZE;
    P.phase_seed = L.merge ? 2 : 1;
    std::map<int, int> slot2col;
    auto bind = [&](int slot) -> int {
        auto it = slot2col.find(slot);
        if (it != slot2col.end()) return it->second;
        const srb_column* c = find_slot(first, slot);
        if (!c) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d not in chunk", slot));
        if ((int)L.cols.size() >= MAX_COLS) throw Error(SRB_NOT_IMPLEMENTED, fmt("more than %d referenced columns", MAX_COLS));
        int w = type_width(c->type);
        if (w == 0 && !is_binary(c->type)) throw Error(SRB_NOT_IMPLEMENTED, fmt("column type %d", c->type));
        int idx = (int)L.cols.size();
        L.cols.push_back({slot, c->type, c->precision, c->scale, c->nulls != nullptr});
        P.cols[idx].width = w;
        P.cols[idx].kind = is_binary(c->type) ? CK_BINARY : (c->type == SRB_TYPE_DOUBLE ? CK_DOUBLE : CK_INT);
        slot2col[slot] = idx;
        return idx;
    };
    // predicates: conjuncts first, OR groups made contiguous
    std::vector<bool> done(plan.preds.size(), false);
    int np = 0, next_group = 1;
    for (size_t i = 0; i < plan.preds.size(); i++) {
        if (done[i]) continue;
        auto put = [&](const srb_predicate& s, int group) {
            DPred& d = P.preds[np++];
            d.col = bind(s.slot_id);
            d.op = s.op;
            d.group = group;
            d.lit_lo = (uint64_t)s.lit_lo;
            d.lit_hi = s.lit_hi;
            d.lit_len = 0;
            if (P.cols[d.col].kind == CK_BINARY && s.op != SRB_PRED_IS_NULL && s.op != SRB_PRED_NOT_NULL) {
                // BinaryColumnPredicateCmpBase (column_predicate_cmp.cpp:559-700): the literal's bytes ride inside the program
                if (s.lit_len < 0 || s.lit_len > 16 || (s.lit_len && !s.lit_bytes))
                    throw Error(SRB_NOT_IMPLEMENTED, "string predicate literal longer than 16 bytes");
                uint8_t raw[16] = {0};
                if (s.lit_len) memcpy(raw, s.lit_bytes, (size_t)s.lit_len);
                memcpy(&d.lit_lo, raw, 8);
                memcpy(&d.lit_hi, raw + 8, 8);
                d.lit_len = (int32_t)s.lit_len;
            }
        };
        if (plan.preds[i].or_group == 0) {
            put(plan.preds[i], 0);
        } else {
            int g = next_group++;
            for (size_t j = i; j < plan.preds.size(); j++)
                if (plan.preds[j].or_group == plan.preds[i].or_group) {
                    put(plan.preds[j], g);
                    done[j] = true;
                }
        }
    }
    P.n_preds = np;
    // expression nodes
    auto trivial_cast_child_type = [&](int ni) -> int {
        const srb_expr_node& e = plan.nodes[ni];
        if (e.op != SRB_EXPR_CAST) return 0;
        const srb_expr_node& c = plan.nodes[e.lhs];
        if (is_decimal(e.type) && is_decimal(c.type) && e.scale == c.scale) return c.type;
        return 0;
    };
    P.n_nodes = (int)plan.nodes.size();
    for (int i = 0; i < P.n_nodes; i++) {
        const srb_expr_node& e = plan.nodes[i];
        DNode& d = P.nodes[i];
        d.lhs = d.rhs = -1;
        d.width = type_width(e.type);
        d.kind = is_binary(e.type) ? CK_BINARY : (e.type == SRB_TYPE_DOUBLE ? CK_DOUBLE : CK_INT);
        switch (e.op) {
        case SRB_EXPR_SLOT_REF: {
            d.op = DOP_LOAD;
            d.col = bind(e.slot_id);
            if (type_width(L.cols[d.col].type) != d.width)
                throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d: node type %d does not match column type %d", e.slot_id, e.type,
                                                      L.cols[d.col].type));
            break;
        }
        case SRB_EXPR_LITERAL:
            d.op = DOP_LIT;
            d.lit_lo = (uint64_t)e.lit_lo;
            d.lit_hi = e.lit_hi;
            break;
        case SRB_EXPR_CAST: {
            const srb_expr_node& c = plan.nodes[e.lhs];
            if (c.type == SRB_TYPE_DOUBLE || e.type == SRB_TYPE_DOUBLE || is_binary(c.type) || is_binary(e.type))
                throw Error(SRB_NOT_IMPLEMENTED, "cast: only integer/decimal casts are supported");
            int up = e.scale - c.scale;
            if (up < 0) throw Error(SRB_NOT_IMPLEMENTED, "cast: scale down not supported");
            if (!is_decimal(e.type) && (up != 0 || d.width < type_width(c.type)))
                throw Error(SRB_NOT_IMPLEMENTED, "cast: narrowing integer cast not supported");
            if (up > 38) throw Error(SRB_INVALID_ARGUMENT, "cast: scale difference too large");
            d.op = DOP_CAST;
            d.lhs = e.lhs;
            d.checked = is_decimal(e.type) ? 1 : 0;
            i128 f = exp10_i128(up);
            d.lit_lo = lo64(f);
            d.lit_hi = hi64(f);
            break;
        }
        default: {
            const srb_expr_node& a = plan.nodes[e.lhs];
            const srb_expr_node& b = plan.nodes[e.rhs];
            if (a.type == SRB_TYPE_DOUBLE || b.type == SRB_TYPE_DOUBLE || e.type == SRB_TYPE_DOUBLE)
                throw Error(SRB_NOT_IMPLEMENTED, "double arithmetic not supported");
            if (is_binary(a.type) || is_binary(b.type) || d.width == 0) throw Error(SRB_INVALID_ARGUMENT, "arithmetic on binary");
            d.lhs = e.lhs;
            d.rhs = e.rhs;
            d.op = e.op == SRB_EXPR_ADD ? DOP_ADD : (e.op == SRB_EXPR_SUB ? DOP_SUB : DOP_MUL);
            bool fast = false;
            if (e.op == SRB_EXPR_MUL && e.type == SRB_TYPE_DECIMAL128) {
                int lt = trivial_cast_child_type(e.lhs), rt = trivial_cast_child_type(e.rhs);
                if ((lt == SRB_TYPE_DECIMAL64 || lt == SRB_TYPE_DECIMAL32) && (rt == SRB_TYPE_DECIMAL64 || rt == SRB_TYPE_DECIMAL32)) fast = true;
            }
            if (fast) {
                d.op = DOP_MULFAST;
                d.lhs = plan.nodes[e.lhs].lhs;
                d.rhs = plan.nodes[e.rhs].lhs;
            } else if (is_decimal(e.type)) {
                if (d.width < 4) throw Error(SRB_INVALID_ARGUMENT, "decimal narrower than 4 bytes");
                d.checked = 1;
                if (e.op != SRB_EXPR_MUL) {
                    int s = a.scale > b.scale ? a.scale : b.scale;
                    if (s - a.scale > 38 || s - b.scale > 38) throw Error(SRB_INVALID_ARGUMENT, "scale difference too large");
                    if (s != a.scale) {
                        d.adj_l = 1;
                        i128 f = exp10_i128(s - a.scale);
                        d.fl_lo = lo64(f);
                        d.fl_hi = hi64(f);
                    }
                    if (s != b.scale) {
                        d.adj_r = 1;
                        i128 f = exp10_i128(s - b.scale);
                        d.fr_lo = lo64(f);
                        d.fr_hi = hi64(f);
                    }
                }
            }
        }
        }
    }
    // keys
    P.n_keys = (int)plan.key_nodes.size();
    int kb = 0;
    bool var = false;
    for (int k = 0; k < P.n_keys; k++) {
        DKey& d = P.keys[k];
        d.node = plan.key_nodes[k];
        const DNode& nd = P.nodes[d.node];
        const ColMeta& cm = L.cols[nd.col];
        d.nullable = cm.nullable ? 1 : 0;
        d.width = nd.kind == CK_BINARY ? 0 : nd.width;
        // a CHAR / VARCHAR key column may carry its declared length (TypeDescriptor::len) in `precision`: the key area is sized for it.
        // Without one (0) the key area stays at 16 bytes and a row whose serialized key is longer is reported when it is met.
        kb += d.nullable + (d.width ? d.width : 4 + std::max(0, cm.precision));
        if (d.width == 0) var = true;
        L.key_out.push_back({k, cm.type, cm.precision, cm.scale, cm.nullable});
    }
    if (kb > 8 * MAX_KEY_WORDS) throw Error(SRB_NOT_IMPLEMENTED, fmt("packed group key of %d bytes (limit %d)", kb, 8 * MAX_KEY_WORDS));
    P.key_bytes = var ? 0 : kb;
    P.key_words = kb > 16 ? (kb + 7) / 8 : ((var || kb > 8) ? 2 : 1); // beyond two words: wide keys (the interpreter's, srb_rowpipe.cuh)
    // aggregates + slot layout
    P.n_aggs = (int)plan.aggs.size();
    auto can_be_null = [&](int ni) {
        // conservative: a column with a null map, or any checked operation anywhere below
        std::vector<int> st{ni};
        while (!st.empty()) {
            int x = st.back();
            st.pop_back();
            const DNode& nd = P.nodes[x];
            if (nd.op == DOP_LOAD && L.cols[nd.col].nullable) return true;
            if (nd.checked) return true;
            if (nd.lhs >= 0) st.push_back(nd.lhs);
            if (nd.rhs >= 0) st.push_back(nd.rhs);
        }
        return false;
    };
    // slot header: pair A {meta, key word 0}, pair B {key word 1, ready flag} for 16-byte keys (srb_prog.h)
    int off = (P.n_keys && P.key_words >= 2) ? 32 + (((P.key_words - 2) * 8 + 15) & ~15) : 16;
    P.off_state = off;
    // pairs need 16-byte alignment
    std::vector<int> need_pair(P.n_aggs, 0), need_n(P.n_aggs, 0);
    for (int a = 0; a < P.n_aggs; a++) {
        DAgg& g = P.aggs[a];
        const srb_agg_fn& s = plan.aggs[a];
        g.fn = s.fn;
        g.arg = s.arg;
        g.merge = L.merge ? 1 : 0;
        g.blob_col = -1;
        g.off_lo = g.off_hi = g.off_n = -1;
        AggOutMeta m{};
        m.fn = s.fn;
        int at = 0, as = 0;
        if (s.arg >= 0) {
            const srb_expr_node& nd = plan.nodes[s.arg];
            at = nd.type;
            as = nd.scale;
            if (L.merge && s.fn == SRB_AGG_AVG) {
                if (!is_binary(nd.type) || P.nodes[s.arg].op != DOP_LOAD)
                    throw Error(SRB_INVALID_ARGUMENT, "merge of AVG expects a VARBINARY {sum,count} slot");
                at = nd.precision; // blob column: precision carries the original argument type
                g.blob_col = P.nodes[s.arg].col;
            } else if (is_binary(nd.type))
                throw Error(SRB_NOT_IMPLEMENTED, "aggregate over a binary column");
            if (L.merge && s.fn == SRB_AGG_COUNT && nd.type != SRB_TYPE_BIGINT)
                throw Error(SRB_INVALID_ARGUMENT, "merge of COUNT expects a BIGINT slot");
        } else if (L.merge)
            throw Error(SRB_INVALID_ARGUMENT, "merge phases need an intermediate column per aggregate");
        m.arg_type = at;
        m.arg_scale = as;
        agg_result_type(s.fn, at, as, &m.out_type, &m.out_precision, &m.out_scale);
        bool dbl_arg = at == SRB_TYPE_DOUBLE;
        g.is_double = (s.fn != SRB_AGG_COUNT && dbl_arg) ? 1 : 0;
        g.avg_as_double = (s.fn == SRB_AGG_AVG && !dbl_arg && avg_in_double(at)) ? 1 : 0;
        m.blob = !L.finalize && s.fn == SRB_AGG_AVG;
        m.blob_double = m.blob && (g.is_double || g.avg_as_double);
        if (m.blob) {
            m.out_type = SRB_TYPE_VARBINARY;
            m.out_precision = at;
            m.out_scale = as;
            m.out_width = m.blob_double ? 8 : 16;
        } else
            m.out_width = type_width(m.out_type);
        m.nullable = s.fn != SRB_AGG_COUNT;
        bool state_double = g.is_double || g.avg_as_double;
        if (s.fn == SRB_AGG_COUNT) {
            need_n[a] = 1;
        } else if (s.fn == SRB_AGG_SUM) {
            need_pair[a] = (!state_double && m.out_width == 16) ? 1 : 0;
            need_n[a] = (L.merge || P.n_keys == 0 || can_be_null(s.arg)) ? 1 : 0;
        } else if (s.fn == SRB_AGG_AVG) {
            need_pair[a] = state_double ? 0 : 1;
            need_n[a] = 1;
        } else {
            g.wide = (!state_double && type_width(at) == 16) ? 1 : 0;
            need_pair[a] = g.wide;
            need_n[a] = (L.merge || P.n_keys == 0 || can_be_null(s.arg)) ? 1 : 0;
        }
        L.agg_out.push_back(m);
    }
    off = (off + 15) & ~15;
    for (int a = 0; a < P.n_aggs; a++)
        if (need_pair[a]) {
            P.aggs[a].off_lo = off;
            P.aggs[a].off_hi = off + 8;
            off += 16;
        }
    for (int a = 0; a < P.n_aggs; a++) {
        if (!need_pair[a] && P.aggs[a].fn != FN_COUNT) {
            P.aggs[a].off_lo = off;
            off += 8;
        }
        if (need_n[a]) {
            P.aggs[a].off_n = off;
            off += 8;
        }
    }
    P.slot_bytes = (off + 15) & ~15;
    P.n_cols = (int)L.cols.size();
    return L;
}

} // namespace srb
