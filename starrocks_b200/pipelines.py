"""Multi-operator pipelines of the benchmark configs, wired the way the BE's pipeline builder wires the fragments: every operator
output is handed to the next operator as a device chunk (no host round trip between operators).

SSB Q4.1: the four dimensions are filtered and become hash-join build sides (replicated on every GPU in
the multi-GPU plan: broadcast join); lineorder probes supplier, customer, part and dates in turn with late materialisation
(each probe carries only the columns later operators read), then GROUP BY d_year, c_nation.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

from . import capi as K, ops, queries as Q

SSB_DIM_COLS = {"ssb_customer": ["sc_custkey", "sc_nation_code", "sc_region_code"], "ssb_supplier": ["ss_suppkey", "ss_region_code"],
                "ssb_part": ["sp_partkey", "sp_mfgr_code"], "ssb_dates": ["sd_datekey", "sd_year"]}


def _filter(ctx, plan, chunk):
    op = ops.ScanAggOperator(ctx, plan)
    op.push_chunk(chunk)
    op.set_finishing()
    return op, op.pull_device()


class SsbQ41:
    """build(): filter the dimensions and build the four join tables (once per query); probe(lineorder): one pass of the fact
    table through the probes and the aggregation. `replicate` (multi-GPU): callable applied to every filtered dimension before
    its build — the broadcast exchange (ops.Communicator.all_gather)."""

    def __init__(self, ctx, replicate: Optional[Callable[[K.Chunk], K.Chunk]] = None):
        self.ctx, self.replicate = ctx, replicate
        self.joins, self.keep = [], []
        self.launches = 0

    def build(self, dims: Dict[str, K.Chunk]):
        plans = [("ssb_supplier", Q.ssb_supplier_filter(), Q.ssb_join_supplier()), ("ssb_customer", Q.ssb_customer_filter(), Q.ssb_join_customer()),
                 ("ssb_part", Q.ssb_part_filter(), Q.ssb_join_part()), ("ssb_dates", None, Q.ssb_join_dates())]
        for name, fplan, jplan in plans:
            side = dims[name]
            if fplan is not None:
                fop, side = _filter(self.ctx, fplan, side)
                self.keep.append(fop)
            if self.replicate is not None:
                side = self.replicate(side)
            j = ops.HashJoinOperator(self.ctx, jplan)
            j.push_chunk(side)
            j.set_finishing()
            self.joins.append(j)
        return self

    def probe(self, lineorder: K.Chunk, phase: int = K.AGG_UPDATE_FINALIZE, stage_rows=None):
        """-> the aggregation operator after set_finishing (pull from it); stage_rows (list) collects the rows after each probe"""
        cur = lineorder
        for j in self.joins:
            j.probe.push_chunk(cur)
            cur = j.probe.pull_device()
            if stage_rows is not None:
                stage_rows.append(cur.n_rows if cur is not None else 0)
            if cur is None:
                break
        agg = ops.ScanAggOperator(self.ctx, Q.ssb_q41_agg_plan(phase=phase))
        if cur is not None and cur.n_rows:
            agg.push_chunk(cur)
        agg.set_finishing()
        return agg

    def close(self):
        for o in self.joins + self.keep:
            o.close()
        self.joins, self.keep = [], []


def ssb_q41_oracle(oracle, dims: Dict[str, K.Chunk], lineorder: K.Chunk) -> K.Chunk:
    """the same chain through the CPU oracle (tests and the CPU baseline only)"""
    s = oracle.scan_agg(Q.ssb_supplier_filter(), dims["ssb_supplier"])
    c = oracle.scan_agg(Q.ssb_customer_filter(), dims["ssb_customer"])
    p = oracle.scan_agg(Q.ssb_part_filter(), dims["ssb_part"])
    cur = oracle.join(Q.ssb_join_supplier(), s, lineorder)
    cur = oracle.join(Q.ssb_join_customer(), c, cur)
    cur = oracle.join(Q.ssb_join_part(), p, cur)
    cur = oracle.join(Q.ssb_join_dates(), dims["ssb_dates"], cur)
    return oracle.scan_agg(Q.ssb_q41_agg_plan(), cur)
