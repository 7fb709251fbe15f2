"""Typed operator pipelines of the benchmark queries, as the FE plans them for the BE.

Expression typing is read from the reference's FE plan goldens (text only):
  Q1  fe/fe-core/src/test/resources/sql/external/hive/tpch/q1.sql:71-102
  Q6  fe/fe-core/src/test/resources/sql/external/hive/tpch/q6.sql:8,24
  Q3  fe/fe-core/src/test/resources/sql/external/hive/tpch/q3.sql:34-36,57-59,100-102
Predicate constants: docs/en/benchmarking/TPC-H_Benchmarking.md:229-251,303-311.
Slot ids are this package's own numbering of the synthetic tables (tpch_synth.py).
"""
from __future__ import annotations

import functools

from . import capi as K
from .plan import ExprBuilder, JoinPlan, ScanAggPlan, pred


def from_date(year: int, month: int, day: int) -> int:
    """Julian day number as date::from_date computes it (be/src/runtime/time_types.cpp:474-491)."""
    if month > 2:
        month += 1
        year += 4800
    else:
        month += 13
        year += 4799
    century = year // 100
    julian = year * 365 - 32167
    julian += year // 4 - century + century // 4
    julian += 7834 * month // 256 + day
    return julian


# lineitem slots
L_ORDERKEY, L_QUANTITY, L_EXTENDEDPRICE, L_DISCOUNT, L_TAX, L_SHIPDATE = 0, 1, 2, 3, 4, 5
L_RETURNFLAG, L_LINESTATUS, L_RETURNFLAG_CODE, L_LINESTATUS_CODE = 6, 7, 8, 9
# orders / customer slots
O_ORDERKEY, O_CUSTKEY, O_ORDERDATE, O_SHIPPRIORITY = 20, 21, 22, 23
C_CUSTKEY, C_MKTSEGMENT_CODE = 30, 31
# high-cardinality table
HC_K, HC_V = 40, 41
# SSB star schema (test/common/sql/ssb/create.sql: every key and measure is INT; the VARCHAR attributes the query touches
# are carried as dictionary codes, as the low-cardinality dictionary rewrite does)
LO_CUSTKEY, LO_PARTKEY, LO_SUPPKEY, LO_ORDERDATE, LO_REVENUE, LO_SUPPLYCOST = 50, 51, 52, 53, 54, 55
SC_CUSTKEY, SC_NATION_CODE, SC_REGION_CODE = 60, 61, 62
SS_SUPPKEY, SS_REGION_CODE = 65, 66
SP_PARTKEY, SP_MFGR_CODE = 70, 71
SD_DATEKEY, SD_YEAR = 75, 76

D64 = K.TYPE_DECIMAL64
D128 = K.TYPE_DECIMAL128


def q6_plan(phase: int = K.AGG_UPDATE_FINALIZE) -> ScanAggPlan:
    """select sum(l_extendedprice * l_discount) from lineitem
       where l_shipdate >= '1994-01-01' and l_shipdate < '1995-01-01' and l_discount between 0.05 and 0.07 and l_quantity < 24"""
    e = ExprBuilder()
    ext = e.slot(L_EXTENDEDPRICE, D64, 15, 2)
    disc = e.slot(L_DISCOUNT, D64, 15, 2)
    prod = e.mul(e.cast(ext, D128, 15, 2), e.cast(disc, D128, 15, 2), D128, 38)  # fast-mul: i64 x i64 -> i128, scale 4
    preds = [pred(L_SHIPDATE, K.PRED_GE, K.TYPE_DATE, from_date(1994, 1, 1)),
             pred(L_SHIPDATE, K.PRED_LT, K.TYPE_DATE, from_date(1995, 1, 1)),
             pred(L_DISCOUNT, K.PRED_GE, D64, 5), pred(L_DISCOUNT, K.PRED_LE, D64, 7),
             pred(L_QUANTITY, K.PRED_LT, D64, 2400)]
    return ScanAggPlan(preds=preds, exprs=e, aggs=[(K.AGG_SUM, prod)], phase=phase)


def q1_plan(dict_keys: bool = True, phase: int = K.AGG_UPDATE_FINALIZE, group_hint: int = 4, with_min_max: bool = False,
            more_aggs: bool = False) -> ScanAggPlan:
    """select l_returnflag, l_linestatus, sum(qty), sum(ext), sum(ext*(1-disc)), sum(ext*(1-disc)*(1+tax)),
              avg(qty), avg(ext), avg(disc), count(*) from lineitem where l_shipdate <= '1998-09-02' group by 1, 2"""
    e = ExprBuilder()
    qty = e.slot(L_QUANTITY, D64, 15, 2)
    ext = e.slot(L_EXTENDEDPRICE, D64, 15, 2)
    disc = e.slot(L_DISCOUNT, D64, 15, 2)
    tax = e.slot(L_TAX, D64, 15, 2)
    if dict_keys:
        k1 = e.slot(L_RETURNFLAG_CODE, K.TYPE_INT)
        k2 = e.slot(L_LINESTATUS_CODE, K.TYPE_INT)
    else:
        k1 = e.slot(L_RETURNFLAG, K.TYPE_VARCHAR)
        k2 = e.slot(L_LINESTATUS, K.TYPE_VARCHAR)
    one = e.lit(100, D64, 16, 2)
    one_minus_disc = e.sub(one, disc, D64, 16)                                                        # DECIMAL64(16,2)
    disc_price = e.mul(e.cast(ext, D128, 15, 2), e.cast(one_minus_disc, D128, 16, 2), D128, 31)       # DECIMAL128(31,4), fast-mul
    one_plus_tax = e.add(one, tax, D64, 16)
    charge = e.mul(disc_price, e.cast(one_plus_tax, D128, 16, 2), D128, 38)                           # DECIMAL128(38,6), checked
    aggs = [(K.AGG_SUM, qty), (K.AGG_SUM, ext), (K.AGG_SUM, disc_price), (K.AGG_SUM, charge),
            (K.AGG_AVG, qty), (K.AGG_AVG, ext), (K.AGG_AVG, disc), (K.AGG_COUNT, None)]
    if with_min_max:  # a ninth and tenth aggregate: the plan no longer matches a static shape
        aggs += [(K.AGG_MIN, ext), (K.AGG_MAX, ext)]
    if more_aggs:  # Q1 as someone else would have written it: two more aggregates, one of them COUNT(x)
        aggs += [(K.AGG_SUM, tax), (K.AGG_COUNT, tax)]
    preds = [pred(L_SHIPDATE, K.PRED_LE, K.TYPE_DATE, from_date(1998, 9, 2))]
    return ScanAggPlan(preds=preds, exprs=e, keys=[k1, k2], aggs=aggs, phase=phase, group_hint=group_hint)


def q1_merge_plan(inter: K.Chunk) -> ScanAggPlan:
    """phase 2 of Q1 over the intermediate rows of q1_plan(phase=AGG_UPDATE_SERIALIZE): slots 0,1 = keys, 2..9 = the eight
    states in aggregate order (SUM values, AVG {sum,count} blobs, COUNT) (aggregator.cpp:871-875, avg.h:161-181)"""
    e = ExprBuilder()
    c = inter.columns
    keys = [e.slot(0, c[0].type), e.slot(1, c[1].type)]
    fns = [K.AGG_SUM] * 4 + [K.AGG_AVG] * 3 + [K.AGG_COUNT]
    args = [e.slot(2 + a, c[2 + a].type, c[2 + a].precision, c[2 + a].scale) for a in range(8)]
    return ScanAggPlan(exprs=e, keys=keys, aggs=list(zip(fns, args)), phase=K.AGG_MERGE_FINALIZE)


def q6_merge_plan(inter: K.Chunk) -> ScanAggPlan:
    """phase 2 of Q6 over the intermediate rows of q6_plan(phase=AGG_UPDATE_SERIALIZE): one DECIMAL128 partial sum per rank
    (NULL when the rank selected no row), merged by SUM (aggregator.cpp:871-875, sum.h:113-117)"""
    e = ExprBuilder()
    c = inter.columns[0]
    return ScanAggPlan(exprs=e, aggs=[(K.AGG_SUM, e.slot(0, c.type, c.precision, c.scale))], phase=K.AGG_MERGE_FINALIZE)


def highcard_plan(phase: int = K.AGG_UPDATE_FINALIZE, group_hint: int = 0) -> ScanAggPlan:
    """select k, sum(v), count(*) from t group by k   (k, v BIGINT)"""
    e = ExprBuilder()
    k = e.slot(HC_K, K.TYPE_BIGINT)
    v = e.slot(HC_V, K.TYPE_BIGINT)
    return ScanAggPlan(exprs=e, keys=[k], aggs=[(K.AGG_SUM, v), (K.AGG_COUNT, None)], phase=phase, group_hint=group_hint)


def highcard_merge_plan(group_hint: int = 0) -> ScanAggPlan:
    """phase 2 of the two-phase aggregation: rows are (k, partial sum, partial count) tuples (aggregator.cpp:871-875)"""
    e = ExprBuilder()
    k = e.slot(0, K.TYPE_BIGINT)
    s = e.slot(1, K.TYPE_BIGINT)
    c = e.slot(2, K.TYPE_BIGINT)
    return ScanAggPlan(exprs=e, keys=[k], aggs=[(K.AGG_SUM, s), (K.AGG_COUNT, c)], phase=K.AGG_MERGE_FINALIZE, group_hint=group_hint)


def filter_plan(preds, out_slots) -> ScanAggPlan:
    return ScanAggPlan(preds=preds, out_slots=out_slots)


# ---- Q3: customer (filtered) |><| orders (filtered) |><| lineitem (filtered) -> group by -----------------------------
Q3_DATE = from_date(1995, 3, 15)
SEG_BUILDING = 1  # code of 'BUILDING' in tpch_synth.MKTSEGMENTS


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def q3_customer_filter() -> ScanAggPlan:
    return filter_plan([pred(C_MKTSEGMENT_CODE, K.PRED_EQ, K.TYPE_TINYINT, SEG_BUILDING)], [C_CUSTKEY])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def q3_orders_filter() -> ScanAggPlan:
    return filter_plan([pred(O_ORDERDATE, K.PRED_LT, K.TYPE_DATE, Q3_DATE)], [O_ORDERKEY, O_CUSTKEY, O_ORDERDATE, O_SHIPPRIORITY])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def q3_lineitem_filter() -> ScanAggPlan:
    return filter_plan([pred(L_SHIPDATE, K.PRED_GT, K.TYPE_DATE, Q3_DATE)], [L_ORDERKEY, L_EXTENDEDPRICE, L_DISCOUNT])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def q3_join_customer_orders() -> JoinPlan:
    # build: filtered customer; probe: filtered orders; output: the orders columns the next join needs
    return JoinPlan([C_CUSTKEY], [O_CUSTKEY], build_out=[], probe_out=[O_ORDERKEY, O_ORDERDATE, O_SHIPPRIORITY])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def q3_join_orders_lineitem() -> JoinPlan:
    return JoinPlan([O_ORDERKEY], [L_ORDERKEY], build_out=[O_ORDERDATE, O_SHIPPRIORITY],
                    probe_out=[L_ORDERKEY, L_EXTENDEDPRICE, L_DISCOUNT])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def q3_agg_plan(group_hint: int = 0) -> ScanAggPlan:
    """group by l_orderkey, o_orderdate, o_shippriority; sum(l_extendedprice * (1 - l_discount))"""
    e = ExprBuilder()
    ok = e.slot(L_ORDERKEY, K.TYPE_INT)
    od = e.slot(O_ORDERDATE, K.TYPE_DATE)
    sp = e.slot(O_SHIPPRIORITY, K.TYPE_INT)
    ext = e.slot(L_EXTENDEDPRICE, D64, 15, 2)
    disc = e.slot(L_DISCOUNT, D64, 15, 2)
    one = e.lit(100, D64, 16, 2)
    omd = e.sub(one, disc, D64, 16)
    rev = e.mul(e.cast(ext, D128, 15, 2), e.cast(omd, D128, 16, 2), D128, 31)
    return ScanAggPlan(exprs=e, keys=[ok, od, sp], aggs=[(K.AGG_SUM, rev)], group_hint=group_hint)


# ---- SSB Q4.1: lineorder |><| supplier |><| customer |><| part |><| dates -> group by d_year, c_nation ----------------------
#   select d_year, c_nation, sum(lo_revenue - lo_supplycost) as profit from dates, customer, supplier, part, lineorder
#   where lo_custkey = c_custkey and lo_suppkey = s_suppkey and lo_partkey = p_partkey and lo_orderdate = d_datekey
#     and c_region = 'AMERICA' and s_region = 'AMERICA' and (p_mfgr = 'MFGR#1' or p_mfgr = 'MFGR#2') group by d_year, c_nation
# The dimensions are filtered, then replicated on every GPU (broadcast join); lineorder is range-partitioned and probes them in turn.
REGION_AMERICA = 1  # code of 'AMERICA' in tpch_synth.REGIONS
MFGR_1, MFGR_2 = 0, 1


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_customer_filter() -> ScanAggPlan:
    return filter_plan([pred(SC_REGION_CODE, K.PRED_EQ, K.TYPE_INT, REGION_AMERICA)], [SC_CUSTKEY, SC_NATION_CODE])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_supplier_filter() -> ScanAggPlan:
    return filter_plan([pred(SS_REGION_CODE, K.PRED_EQ, K.TYPE_INT, REGION_AMERICA)], [SS_SUPPKEY])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_part_filter() -> ScanAggPlan:
    return filter_plan([pred(SP_MFGR_CODE, K.PRED_EQ, K.TYPE_INT, MFGR_1, or_group=1), pred(SP_MFGR_CODE, K.PRED_EQ, K.TYPE_INT, MFGR_2, or_group=1)],
                       [SP_PARTKEY])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_join_supplier() -> JoinPlan:
    return JoinPlan([SS_SUPPKEY], [LO_SUPPKEY], build_out=[], probe_out=[LO_CUSTKEY, LO_PARTKEY, LO_ORDERDATE, LO_REVENUE, LO_SUPPLYCOST])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_join_customer() -> JoinPlan:
    return JoinPlan([SC_CUSTKEY], [LO_CUSTKEY], build_out=[SC_NATION_CODE], probe_out=[LO_PARTKEY, LO_ORDERDATE, LO_REVENUE, LO_SUPPLYCOST])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_join_part() -> JoinPlan:
    return JoinPlan([SP_PARTKEY], [LO_PARTKEY], build_out=[], probe_out=[SC_NATION_CODE, LO_ORDERDATE, LO_REVENUE, LO_SUPPLYCOST])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_join_dates() -> JoinPlan:
    return JoinPlan([SD_DATEKEY], [LO_ORDERDATE], build_out=[SD_YEAR], probe_out=[SC_NATION_CODE, LO_REVENUE, LO_SUPPLYCOST])


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_q41_agg_plan(phase: int = K.AGG_UPDATE_FINALIZE, group_hint: int = 64) -> ScanAggPlan:
    """group by d_year, c_nation; sum(lo_revenue - lo_supplycost): INT - INT is computed in BIGINT, the sum is BIGINT"""
    e = ExprBuilder()
    year = e.slot(SD_YEAR, K.TYPE_INT)
    nation = e.slot(SC_NATION_CODE, K.TYPE_INT)
    rev = e.cast(e.slot(LO_REVENUE, K.TYPE_INT), K.TYPE_BIGINT)
    cost = e.cast(e.slot(LO_SUPPLYCOST, K.TYPE_INT), K.TYPE_BIGINT)
    return ScanAggPlan(exprs=e, keys=[year, nation], aggs=[(K.AGG_SUM, e.sub(rev, cost, K.TYPE_BIGINT))], phase=phase, group_hint=group_hint)


@functools.lru_cache(maxsize=None)  # a plan is immutable once built: the FE hands it over once per query
def ssb_q41_merge_plan(group_hint: int = 64) -> ScanAggPlan:
    e = ExprBuilder()
    year = e.slot(0, K.TYPE_INT)
    nation = e.slot(1, K.TYPE_INT)
    return ScanAggPlan(exprs=e, keys=[year, nation], aggs=[(K.AGG_SUM, e.slot(2, K.TYPE_BIGINT))], phase=K.AGG_MERGE_FINALIZE, group_hint=group_hint)
