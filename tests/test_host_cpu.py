"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol include/srb/srb_c_api.h declares
(no compute calls without a GPU), the plan builders produce the typed plans the kernels' static shapes expect, and the
synthetic-table generator is identical on its numpy and torch back ends."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from starrocks_b200 import capi as K
from starrocks_b200 import queries as Q
from starrocks_b200 import tpch_synth as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def libsrb():
    from starrocks_b200 import build, ops
    build.build()
    return ops.lib()


def test_library_exports_every_declared_symbol(libsrb):
    from starrocks_b200 import ops
    hdr = open(os.path.join(ROOT, "include", "srb", "srb_c_api.h")).read()
    declared = set(re.findall(r"^(?:srb_status|void\*?|int32_t|const char\*)\s+(srb_[a-z0-9_]+)\(", hdr, flags=re.M))
    assert len(declared) >= 40
    for name in sorted(declared):
        assert hasattr(libsrb, name), f"{name} declared in the header but not exported"
    assert declared == set(ops.EXPORTS), declared ^ set(ops.EXPORTS)


def test_no_cpu_fallback_without_device(libsrb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from starrocks_b200 import ops
    with pytest.raises(ops.SrbError) as ei:
        ops.Context(device=0)
    assert "no CUDA device" in str(ei.value)


def test_struct_sizes_match_the_header():
    # byte-for-byte mirrors of the C PODs
    assert C.sizeof(K.srb_column) == 48
    assert C.sizeof(K.srb_chunk) == 32
    assert C.sizeof(K.srb_predicate) == 48  # + the string literal (pointer, length)
    assert C.sizeof(K.srb_expr_node) == 48
    assert C.sizeof(K.srb_agg_fn) == 8
    assert C.sizeof(K.srb_config) == 32


def test_ctypes_mirrors_have_the_sizes_the_c_compiler_gives_the_header(tmp_path):
    # sizeof of every POD of include/srb/srb_c_api.h as gcc lays it out, against the ctypes mirrors the Python host side passes
    import subprocess
    from starrocks_b200 import pages as PG
    mirrors = {"srb_column": K.srb_column, "srb_chunk": K.srb_chunk, "srb_predicate": K.srb_predicate, "srb_expr_node": K.srb_expr_node,
               "srb_agg_fn": K.srb_agg_fn, "srb_config": K.srb_config, "srb_scan_agg_plan": K.srb_scan_agg_plan, "srb_join_plan": K.srb_join_plan,
               "srb_join_conjunct": K.srb_join_conjunct, "srb_topn_plan": K.srb_topn_plan, "srb_page": PG.srb_page, "srb_page_column": PG.srb_page_column}
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "srb/srb_c_api.h"\nint main(void) {\n' +
                   "".join(f'    printf("{n} %zu\\n", sizeof({n}));\n' for n in mirrors) + "    return 0;\n}\n")
    exe = tmp_path / "sizes"
    subprocess.check_call(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = dict(line.split() for line in subprocess.check_output([str(exe)], text=True).splitlines())
    for n, cls in mirrors.items():
        assert int(got[n]) == C.sizeof(cls), (n, got[n], C.sizeof(cls))


def test_from_date_matches_oracle(oracle):
    for y, m, d in [(1970, 1, 1), (1992, 1, 2), (1994, 1, 1), (1995, 3, 15), (1998, 9, 2), (2000, 2, 29), (1000, 1, 1)]:
        assert Q.from_date(y, m, d) == oracle.lib.orc_from_date(y, m, d)
    assert Q.from_date(1970, 1, 1) == 2440588


def test_q1_plan_is_the_typed_fe_plan():
    p = Q.q1_plan()
    ops_ = [p.nodes[i].op for i in range(p.c.n_nodes)]
    assert ops_ == [K.EXPR_SLOT_REF] * 6 + [K.EXPR_LITERAL, K.EXPR_SUB, K.EXPR_CAST, K.EXPR_CAST, K.EXPR_MUL, K.EXPR_ADD, K.EXPR_CAST, K.EXPR_MUL]
    assert (p.nodes[10].type, p.nodes[10].scale) == (K.TYPE_DECIMAL128, 4)   # ext * (1 - disc): DECIMAL128(31,4)
    assert (p.nodes[13].type, p.nodes[13].scale) == (K.TYPE_DECIMAL128, 6)   # * (1 + tax): DECIMAL128(38,6)
    assert p.c.n_aggs == 8 and p.c.n_keys == 2


def test_synth_numpy_and_torch_agree():
    import torch
    for table, cols, kw in (("lineitem", T.LINEITEM_Q1 + ["l_orderkey"], {}), ("orders", ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"], {"n_customers": 15000}),
                            ("customer", ["c_custkey", "c_mktsegment_code"], {}), ("highcard", ["k", "v"], {"n_keys": 100_000_000})):
        a = T._gen(table, T._Np, 123456789, 5000, cols, **kw)
        b = T._gen(table, T._Torch("cpu"), 123456789, 5000, cols, **kw)
        for c in cols:
            assert np.array_equal(np.asarray(a[c]).astype(np.int64), b[c].to(torch.int64).numpy()), (table, c)


def test_synth_q1_has_four_groups_and_q6_selectivity():
    ch = T.host_chunk("lineitem", 0, 200000, T.LINEITEM_Q1)
    rf, ls = ch.columns[Q.L_RETURNFLAG_CODE].data, ch.columns[Q.L_LINESTATUS_CODE].data
    assert sorted(set(zip(rf.tolist(), ls.tolist()))) == [(0, 0), (1, 0), (1, 1), (2, 0)]  # (A,F) (N,F) (N,O) (R,F)
    ship, disc, qty = ch.columns[Q.L_SHIPDATE].data, ch.columns[Q.L_DISCOUNT].data, ch.columns[Q.L_QUANTITY].data
    sel = ((ship >= Q.from_date(1994, 1, 1)) & (ship < Q.from_date(1995, 1, 1)) & (disc >= 5) & (disc <= 7) & (qty < 2400)).mean()
    assert 0.015 < sel < 0.025
    assert (ship <= Q.from_date(1998, 9, 2)).mean() > 0.95


def test_oracle_runs_the_benchmark_pipelines(oracle):
    ch = T.host_chunk("lineitem", 0, 50000, T.LINEITEM_Q1)
    r1 = oracle.scan_agg(Q.q1_plan(), ch, n_threads=2)
    assert r1.n_rows == 4
    rows = K.chunk_rows(r1)
    assert sum(r[-1] for r in rows) == int((ch.columns[Q.L_SHIPDATE].data <= Q.from_date(1998, 9, 2)).sum())
    # sum(qty) / count == avg(qty) under decimal_div_integer
    for r in rows:
        assert r[6] == oracle.decimal_div_integer(r[2], r[9], 2)
    r6 = oracle.scan_agg(Q.q6_plan(), T.host_chunk("lineitem", 0, 50000, T.LINEITEM_Q6))
    ship, disc, qty, ext = (ch.columns[s].data for s in (Q.L_SHIPDATE, Q.L_DISCOUNT, Q.L_QUANTITY, Q.L_EXTENDEDPRICE))
    m = (ship >= Q.from_date(1994, 1, 1)) & (ship < Q.from_date(1995, 1, 1)) & (disc >= 5) & (disc <= 7) & (qty < 2400)
    assert K.column_pyvalues(r6.columns[0])[0] == int((ext[m].astype(object) * disc[m].astype(object)).sum())


def test_boundary_header_is_plain_c():
    """the drop-in boundary is a C ABI: include/srb/srb_c_api.h compiles as C99 on its own (no C++ types, no missing includes)"""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", os.path.join(root, "include", "srb", "srb_c_api.h")])


def test_oracle_join_with_probe_drivers_returns_the_same_rows_in_the_same_order(oracle):
    """the oracle's probe side runs with DOP drivers over one shared table (hash_join_probe_operator.cpp:104): any thread count, same rows"""
    import numpy as np
    from starrocks_b200 import capi as K
    from starrocks_b200.plan import JoinPlan
    rng = np.random.default_rng(8)
    nb, npb = 20_000, 300_000
    build = K.Chunk(nb, {0: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 15_000, nb).astype(np.int32)), 1: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(0, 1 << 40, nb))})
    probe = K.Chunk(npb, {2: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 30_000, npb).astype(np.int32), (rng.random(npb) < 0.05).astype(np.uint8)),
                          3: K.Column.from_ints(K.TYPE_DECIMAL128, [int(x) for x in rng.integers(-2**40, 2**40, npb)], None, 27, 2)})
    for jt in (K.JOIN_INNER, K.JOIN_LEFT_SEMI, K.JOIN_LEFT_ANTI):
        plan = JoinPlan([0], [2], build_out=[1], probe_out=[2, 3], join_type=jt)
        one = K.chunk_rows(oracle.join(plan, build, probe), sort=False)
        for t in (2, 7):
            assert K.chunk_rows(oracle.join(plan, build, probe, n_threads=t), sort=False) == one
