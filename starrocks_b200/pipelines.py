"""Multi-operator pipelines of the benchmark configs, wired the way the BE's pipeline builder wires the fragments: every operator
output is handed to the next operator as a device chunk (no host round trip between operators).

TPC-H Q3: customer (filtered) is the build side of the orders join; at N > 1 it is replicated on every rank
(all-gather: the FE plans this join as BROADCAST, fe/fe-core/src/test/resources/sql/external/hive/tpch/q3.sql:100-115); the joined
orders and the filtered lineitem rows are both hash-partitioned on the order key and exchanged all-to-all (the "INNER JOIN
(PARTITIONED)" of q3.sql:57-79), so every rank builds and probes its own key range and its groups (the key contains l_orderkey) are
disjoint from every other rank's.

SSB Q4.1: the four dimensions are filtered and become hash-join build sides (replicated on every GPU in
the multi-GPU plan: broadcast join); lineorder probes supplier, customer, part and dates in turn with late materialisation
(each probe carries only the columns later operators read), then GROUP BY d_year, c_nation.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

from . import capi as K, ops, pages as PG, queries as Q

SSB_DIM_COLS = {"ssb_customer": ["sc_custkey", "sc_nation_code", "sc_region_code"], "ssb_supplier": ["ss_suppkey", "ss_region_code"],
                "ssb_part": ["sp_partkey", "sp_mfgr_code"], "ssb_dates": ["sd_datekey", "sd_year"]}


def _filter(ctx, plan, chunk):
    """chunk: a Chunk, or a pages.PageTable (encoded segment pages, decoded on the device in front of the scan)"""
    op = ops.ScanAggOperator(ctx, plan)
    PG.push_source(op, chunk)
    op.set_finishing()
    return op, op.pull_device()


def having(ctx, agg_op, preds, out_slots):
    """HAVING: the conjuncts AggregateBlockingSourceOperator::pull_chunk applies to the chunks it emits (eval_conjuncts_and_in_filters,
    be/src/exec/pipeline/aggregate/aggregate_blocking_source_operator.cpp:62-68) = a filter operator over the aggregate's device result,
    no host hop in between. `preds` name the result's slots (group keys 0..k-1, then the aggregates). -> the filter operator after
    set_finishing (pull from it; close it and the aggregate afterwards)."""
    f = ops.ScanAggOperator(ctx, Q.filter_plan(preds, out_slots))
    groups = agg_op.pull_device()
    if groups is not None and groups.n_rows:
        f.push_chunk(groups)
    f.set_finishing()
    return f


class TpchQ3:
    """One pass of Q3 over this rank's row ranges of customer / orders / lineitem. `comm` (ops.Communicator) and `part`
    (ops.Partitioner over world_size destinations) switch on the multi-GPU plan; without them everything is local."""

    STAGES = ["customer filter (+ broadcast) + build J1", "orders filter", "probe J1 (+ shuffle) + build J2", "lineitem filter (+ shuffle)",
              "probe J2 + gather", "group by"]

    def __init__(self, ctx, comm=None, part=None, fused: Optional[bool] = None, runtime_filter: Optional[bool] = None):
        """fused (default: unless SRB_NO_P2P is set): shuffle with Communicator.shuffle (peer-memory scatter) instead of
        Partitioner.run + Communicator.all_to_all (NCCL send/recv).
        runtime_filter (default: unless SRB_NO_RF is set): the orders join publishes a runtime filter on its build keys and the
        lineitem scan applies it to l_orderkey before compacting (and before the shuffle); at N > 1 the ranks' partial filters are
        merged into the global one first (hash_join_build_operator.cpp:97-208, the FE's plan: q3.sql "build runtime filters")."""
        import os
        self.ctx, self.comm, self.part = ctx, comm, part
        self.fused = (not os.environ.get("SRB_NO_P2P")) if fused is None else fused
        self.use_rf = (not os.environ.get("SRB_NO_RF")) if runtime_filter is None else runtime_filter
        self.open = []

    def _shuffle(self, chunk, key_slot):
        if self.comm is None:
            return chunk
        if chunk is None:
            raise ValueError("a rank with no chunk to shuffle cannot describe the schema of the exchange")
        if self.fused:  # partitioning fused with the exchange: rows are scattered straight into the peers' receive buffers
            recv, _ = self.comm.shuffle(chunk, [key_slot])
            return recv
        grouped, counts = self.part.run(chunk, [key_slot])
        recv, _ = self.comm.all_to_all(grouped, counts)
        return recv

    def run(self, cust: K.Chunk, orders: K.Chunk, line: K.Chunk, mark: Optional[Callable[[], None]] = None, rows: Optional[dict] = None):
        """-> the aggregation operator after set_finishing (pull from it; close() afterwards). `mark()` is called after every stage
        (STAGES), `rows` collects the row counts between the operators."""
        mark = mark or (lambda: None)
        rows = rows if rows is not None else {}
        fc, dc = _filter(self.ctx, Q.q3_customer_filter(), cust)
        self.open.append(fc)
        rows["customer_kept"] = dc.n_rows
        if self.comm is not None:
            dc = self.comm.all_gather(dc)
        j1 = ops.HashJoinOperator(self.ctx, Q.q3_join_customer_orders())
        self.open.append(j1)
        j1.push_chunk(dc)
        j1.set_finishing()
        mark()
        fo, do = _filter(self.ctx, Q.q3_orders_filter(), orders)
        self.open.append(fo)
        rows["orders_kept"] = do.n_rows
        mark()
        j1.probe.push_chunk(do)
        d1 = j1.probe.pull_device()
        rows["j1_out"] = d1.n_rows if d1 is not None else 0
        d1 = self._shuffle(d1, Q.O_ORDERKEY)
        j2 = ops.HashJoinOperator(self.ctx, Q.q3_join_orders_lineitem())
        self.open.append(j2)
        if d1 is not None and d1.n_rows:
            j2.push_chunk(d1)
        j2.set_finishing()
        rf = None
        if self.use_rf:
            n_build = d1.n_rows if d1 is not None else 0
            if self.comm is not None:  # the partial filters are merged, so every rank sizes its own for the global build side
                import numpy as np
                mine = K.Chunk(1, {0: K.Column.from_ints(K.TYPE_BIGINT, np.array([n_build], dtype=np.int64))})
                n_build = int(sum(r[0] for r in K.chunk_rows(self.comm.all_gather(mine), sort=False)))
            # order keys use 8 of every 32 values: under the reference tree's identity hash three quarters of the buckets stay empty
            # and most lines pass by accident, so the filter is built with the mixing hash (ops.HashJoinOperator.build_runtime_filter)
            # one rank: this builder holds the whole build side, so the filter may be a RuntimeBitsetFilter (create_join_runtime_filter)
            rf = j2.build_runtime_filter(0, max(1, n_build), mix=True) if self.comm is not None else j2.build_runtime_filter(0, 0, mix=True, bitset=True)
            if self.comm is not None:
                self.comm.merge_runtime_filter(rf)
            self.open.append(rf)
        mark()
        fl = ops.ScanAggOperator(self.ctx, Q.q3_lineitem_filter())
        self.open.insert(0, fl)  # closed before the filter it reads
        if rf is not None:
            fl.add_runtime_filter(rf, Q.L_ORDERKEY)
        PG.push_source(fl, line)
        fl.set_finishing()
        dl = fl.pull_device()
        if dl is None:
            raise ValueError("the lineitem scan kept no row on this rank")
        rows["lineitem_kept"] = dl.n_rows
        dl = self._shuffle(dl, Q.L_ORDERKEY)
        mark()
        d2 = None
        if dl is not None and dl.n_rows:
            j2.probe.push_chunk(dl)
            d2 = j2.probe.pull_device()
        rows["j2_out"] = d2.n_rows if d2 is not None else 0
        mark()
        agg = ops.ScanAggOperator(self.ctx, Q.q3_agg_plan(group_hint=max(1024, rows["j2_out"] // 2)))
        self.open.append(agg)
        if d2 is not None and d2.n_rows:
            agg.push_chunk(d2)
        agg.set_finishing()
        return agg

    # q3.sql:36-38: ORDER BY revenue DESC, o_orderdate LIMIT 10 (DESC puts NULLs last, ASC first: StarRocks' defaults)
    TOP_KEYS, TOP_ASC, TOP_NULLS_FIRST = [3, 1], [0, 1], [0, 1]

    def top(self, agg, limit: int = 10, groups: Optional[K.Chunk] = None) -> K.Chunk:
        """the plan's last operator: the aggregation's device result (`groups`, pulled here when not given) through the device TopN
        (ops.TopNOperator). At N > 1 every rank's groups are disjoint: the ranks' top rows are gathered (<= limit rows each) and
        pass through a second TopN on every rank (the gather exchange + final ORDER BY ... LIMIT of q3.sql)."""
        top = ops.TopNOperator(self.ctx, self.TOP_KEYS, self.TOP_ASC, self.TOP_NULLS_FIRST, limit)
        self.open.append(top)
        if groups is None:
            groups = agg.pull_device()
        if groups is not None and groups.n_rows:
            top.push_chunk(groups)
        top.set_finishing()
        mine = top.pull_all()
        if self.comm is None or self.comm.ctx.world_size == 1:
            return mine
        if not mine.columns:
            raise ValueError("a rank without groups cannot describe the schema of the gather")
        final = ops.TopNOperator(self.ctx, self.TOP_KEYS, self.TOP_ASC, self.TOP_NULLS_FIRST, limit)
        self.open.append(final)
        final.push_chunk(self.comm.all_gather(mine))
        final.set_finishing()
        return final.pull_all()

    def launches(self) -> int:
        return sum(o.counters()["kernel_launches"] for o in self.open if hasattr(o, "counters"))

    def close(self):
        for o in self.open:
            o.close()
        self.open = []


class SsbQ41:
    """build(): filter the dimensions and build the four join tables (once per query); probe(lineorder): one pass of the fact
    table through the probes and the aggregation. `replicate` (multi-GPU): callable applied to every filtered dimension before
    its build — the broadcast exchange (ops.Communicator.all_gather)."""

    def __init__(self, ctx, replicate: Optional[Callable[[K.Chunk], K.Chunk]] = None, runtime_filter: Optional[bool] = None):
        """runtime_filter (default: unless SRB_NO_RF is set): the three filtered dimension joins publish runtime filters and the
        lineorder scan applies them (supplier, customer, part: 1/5 x 1/5 x 2/5 of the rows survive) before anything is copied or
        probed — the local runtime filters of a broadcast join (hash_join_build_operator.cpp:97-208). Dimension keys are dense, so
        the filters use the reference tree's own hash (hash_mix = 0)."""
        import os
        self.ctx, self.replicate = ctx, replicate
        self.use_rf = (not os.environ.get("SRB_NO_RF")) if runtime_filter is None else runtime_filter
        self.joins, self.keep, self.rfs = [], [], []
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
            if self.use_rf and fplan is not None:
                probe_slot = {"ssb_supplier": Q.LO_SUPPKEY, "ssb_customer": Q.LO_CUSTKEY, "ssb_part": Q.LO_PARTKEY}[name]
                self.rfs.append((j.build_runtime_filter(0, bitset=True), probe_slot))  # a broadcast join: every rank builds from the whole dimension
        return self

    def probe(self, lineorder: K.Chunk, phase: int = K.AGG_UPDATE_FINALIZE, stage_rows=None):
        """-> the aggregation operator after set_finishing (pull from it); stage_rows (list) collects the rows after each probe"""
        cur = lineorder
        scan = None
        if self.rfs or isinstance(lineorder, PG.PageTable):  # the lineorder scan: no predicate of its own, three runtime filters (or just the page decode)
            scan = ops.ScanAggOperator(self.ctx, Q.filter_plan([], [Q.LO_CUSTKEY, Q.LO_PARTKEY, Q.LO_SUPPKEY, Q.LO_ORDERDATE, Q.LO_REVENUE, Q.LO_SUPPLYCOST]))
            for rf, slot in self.rfs:
                scan.add_runtime_filter(rf, slot)
            PG.push_source(scan, lineorder)
            scan.set_finishing()
            cur = scan.pull_device()
            self.scan_launches = scan.counters()["kernel_launches"]
            if stage_rows is not None:
                stage_rows.append(cur.n_rows if cur is not None else 0)
        for j in (self.joins if cur is not None else []):
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
        if scan is not None:
            scan.close()  # its rows have been consumed by the first probe
        return agg

    def close(self):
        for o in self.joins + self.keep + [rf for rf, _ in self.rfs]:
            o.close()
        self.joins, self.keep, self.rfs = [], [], []
