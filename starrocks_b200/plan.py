"""Builders for the plan PODs of include/srb/srb_c_api.h.

An `ExprBuilder` flattens an expression tree into the SSA node list the C ABI takes; typing follows the FE's
rules for DecimalV3 (be/src/exprs/arithmetic_operation.h:391-427: add/sub scale = max, mul scale = sum).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

from . import capi as K


@dataclass
class Node:
    builder: "ExprBuilder"
    index: int
    type: int
    precision: int
    scale: int


class ExprBuilder:
    def __init__(self):
        self.nodes: List[K.srb_expr_node] = []

    def _add(self, op, type_, precision, scale, lhs=-1, rhs=-1, slot=-1, lit=0) -> Node:
        n = K.srb_expr_node()
        n.op, n.type, n.precision, n.scale, n.lhs, n.rhs, n.slot_id = op, type_, precision, scale, lhs, rhs, slot
        n.lit_lo, n.lit_hi = K.split128(int(lit))
        self.nodes.append(n)
        return Node(self, len(self.nodes) - 1, type_, precision, scale)

    def slot(self, slot_id: int, type_: int, precision: int = 0, scale: int = 0) -> Node:
        return self._add(K.EXPR_SLOT_REF, type_, precision, scale, slot=slot_id)

    def lit(self, value: int, type_: int, precision: int = 0, scale: int = 0) -> Node:
        """`value` is already in the node's representation (a DECIMAL64(16,2) literal 1 is 100)."""
        return self._add(K.EXPR_LITERAL, type_, precision, scale, lit=value)

    def cast(self, child: Node, type_: int, precision: int = 0, scale: int = 0) -> Node:
        return self._add(K.EXPR_CAST, type_, precision, scale, lhs=child.index)

    def _bin(self, op, a: Node, b: Node, type_: Optional[int], precision: Optional[int]) -> Node:
        if type_ is None:
            type_ = a.type if K.type_width(a.type) >= K.type_width(b.type) else b.type
        if K.is_decimal(type_):
            scale = a.scale + b.scale if op == K.EXPR_MUL else max(a.scale, b.scale)
            if precision is None:
                precision = {4: 9, 8: 18, 16: 38}[K.type_width(type_)]
        else:
            scale, precision = 0, 0
        return self._add(op, type_, precision, scale, lhs=a.index, rhs=b.index)

    def add(self, a, b, type_=None, precision=None):
        return self._bin(K.EXPR_ADD, a, b, type_, precision)

    def sub(self, a, b, type_=None, precision=None):
        return self._bin(K.EXPR_SUB, a, b, type_, precision)

    def mul(self, a, b, type_=None, precision=None):
        return self._bin(K.EXPR_MUL, a, b, type_, precision)


def pred(slot_id: int, op: int, type_: int, literal=0, or_group: int = 0) -> K.srb_predicate:
    p = K.srb_predicate()
    p.slot_id, p.op, p.type, p.or_group = slot_id, op, type_, or_group
    if K.is_binary(type_):  # a string literal: its bytes travel with the predicate (kept alive on the Python object)
        raw = bytes(literal) if not isinstance(literal, str) else literal.encode()
        p._keep = C.create_string_buffer(raw, max(1, len(raw)))
        p.lit_bytes = C.cast(p._keep, C.c_void_p)
        p.lit_len = len(raw)
    elif type_ == K.TYPE_DOUBLE:
        import struct
        p.lit_lo = struct.unpack("<q", struct.pack("<d", float(literal)))[0]
        p.lit_hi = 0
    else:
        p.lit_lo, p.lit_hi = K.split128(int(literal))
    return p


def pred_in(slot_id: int, type_: int, values, or_group: int, negate: bool = False):
    """`col IN (v1, ..., vk)` / `col NOT IN (...)` as the predicates the plan takes (ColumnInPredicate / ColumnNotInPredicate,
    be/src/storage/column_predicate.h:249-277, column_in_predicate.cpp): IN = the OR of k EQ predicates (one or_group), NOT IN = k NE
    conjuncts. A NULL row satisfies neither, as in the reference (a NULL never passes a comparison). At most 16 predicates per plan."""
    if negate:
        return [pred(slot_id, K.PRED_NE, type_, v) for v in values]
    return [pred(slot_id, K.PRED_EQ, type_, v, or_group=or_group) for v in values]


class ScanAggPlan:
    """Owns the ctypes arrays behind one srb_scan_agg_plan."""

    def __init__(self, preds: Sequence[K.srb_predicate] = (), exprs: Optional[ExprBuilder] = None,
                 keys: Sequence[Node] = (), aggs: Sequence[tuple] = (), phase: int = K.AGG_UPDATE_FINALIZE,
                 out_slots: Sequence[int] = (), group_hint: int = 0):
        """aggs: sequence of (fn, Node | None)."""
        self.preds = (K.srb_predicate * max(1, len(preds)))(*preds)
        self._pred_keep = [getattr(q, "_keep", None) for q in preds]  # string literals stay alive with the plan
        nodes = exprs.nodes if exprs else []
        self.nodes = (K.srb_expr_node * max(1, len(nodes)))(*nodes)
        self.key_nodes = (C.c_int32 * max(1, len(keys)))(*[k.index for k in keys])
        self.aggs = (K.srb_agg_fn * max(1, len(aggs)))()
        for i, (fn, arg) in enumerate(aggs):
            self.aggs[i].fn = fn
            self.aggs[i].arg = -1 if arg is None else arg.index
        self.out_slots = (C.c_int32 * max(1, len(out_slots)))(*out_slots)
        self.c = K.srb_scan_agg_plan(len(preds), self.preds, len(nodes), self.nodes, len(keys), self.key_nodes,
                                     len(aggs), self.aggs, phase, len(out_slots), self.out_slots, group_hint)
        self.n_keys, self.n_aggs = len(keys), len(aggs)


class JoinPlan:
    def __init__(self, build_keys: Sequence[int], probe_keys: Sequence[int], build_out: Sequence[int] = (),
                 probe_out: Sequence[int] = (), join_type: int = K.JOIN_INNER, emit_indexes: bool = False,
                 conjuncts: Sequence[tuple] = ()):
        """conjuncts: other-join-conjuncts as (left_slot, K.PRED_*, right_slot) over key / output columns of either side (INNER joins)"""
        assert len(build_keys) == len(probe_keys)
        self.bk = (C.c_int32 * max(1, len(build_keys)))(*build_keys)
        self.pk = (C.c_int32 * max(1, len(probe_keys)))(*probe_keys)
        self.bo = (C.c_int32 * max(1, len(build_out)))(*build_out)
        self.po = (C.c_int32 * max(1, len(probe_out)))(*probe_out)
        self.c = K.srb_join_plan(join_type, len(build_keys), self.bk, self.pk, len(build_out), self.bo,
                                 len(probe_out), self.po, 1 if emit_indexes else 0, len(conjuncts), None)
        if conjuncts:
            self.cj = (K.srb_join_conjunct * len(conjuncts))(*[K.srb_join_conjunct(l, op, r, 0) for l, op, r in conjuncts])
            self.c.conjuncts = C.cast(self.cj, C.c_void_p)
