"""The benchmark pipelines through the CPU oracle, operator by operator (TEST INFRASTRUCTURE: tests, __graft_entry__.smoke() and
bench.py's CPU legs only). Same plans as starrocks_b200/pipelines.py; n_threads = the pipelines' degree of parallelism (scan/filter
and aggregation workers with private tables + merge, probe drivers over one shared join table)."""
from __future__ import annotations

from typing import Dict

from starrocks_b200 import capi as K, queries as Q
from starrocks_b200.pipelines import TpchQ3


def q3_top_oracle(oracle, groups: K.Chunk, limit: int = 10) -> K.Chunk:
    return oracle.sort(groups, TpchQ3.TOP_KEYS, TpchQ3.TOP_ASC, TpchQ3.TOP_NULLS_FIRST, 0, limit)


def q3_oracle(oracle, cust: K.Chunk, orders: K.Chunk, line: K.Chunk, n_threads: int = 1) -> K.Chunk:
    """the same chain through the CPU oracle (tests and the CPU baseline only)"""
    c = oracle.scan_agg(Q.q3_customer_filter(), cust, n_threads=n_threads)
    o = oracle.scan_agg(Q.q3_orders_filter(), orders, n_threads=n_threads)
    l = oracle.scan_agg(Q.q3_lineitem_filter(), line, n_threads=n_threads)
    d1 = oracle.join(Q.q3_join_customer_orders(), c, o, n_threads=n_threads)
    d2 = oracle.join(Q.q3_join_orders_lineitem(), d1, l, n_threads=n_threads)
    return oracle.scan_agg(Q.q3_agg_plan(), d2, n_threads=n_threads)


def ssb_q41_oracle(oracle, dims: Dict[str, K.Chunk], lineorder: K.Chunk, n_threads: int = 1) -> K.Chunk:
    """the same chain through the CPU oracle (tests and the CPU baseline only)"""
    s = oracle.scan_agg(Q.ssb_supplier_filter(), dims["ssb_supplier"], n_threads=n_threads)
    c = oracle.scan_agg(Q.ssb_customer_filter(), dims["ssb_customer"], n_threads=n_threads)
    p = oracle.scan_agg(Q.ssb_part_filter(), dims["ssb_part"], n_threads=n_threads)
    cur = oracle.join(Q.ssb_join_supplier(), s, lineorder, n_threads=n_threads)
    cur = oracle.join(Q.ssb_join_customer(), c, cur, n_threads=n_threads)
    cur = oracle.join(Q.ssb_join_part(), p, cur, n_threads=n_threads)
    cur = oracle.join(Q.ssb_join_dates(), dims["ssb_dates"], cur, n_threads=n_threads)
    return oracle.scan_agg(Q.ssb_q41_agg_plan(), cur, n_threads=n_threads)
