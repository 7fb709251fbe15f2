"""CPU checks of the page builder used for synthetic tables (starrocks_b200/pages.py) against the oracle's restatement of the reference's
ForEncoder (oracle/orc_page.cpp, pinned to the reference's page tests in tests/test_oracle_golden.py): byte-identical pages."""
import numpy as np
import pytest
import torch

from starrocks_b200 import pages as PG


def oracle_pages(oracle, a, page_rows):
    out, pages, off = [], [], 0
    for s in range(0, max(len(a), 1), page_rows):
        p = oracle.for_encode(a[s:s + page_rows])
        out.append(p)
        pages.append((off, len(p), len(a[s:s + page_rows])))
        off += len(p)
    return b"".join(out), pages


# (ranges stay within half the type's range: an overflow frame (format 2) that is not the page's last one is laid out by the reference's
# encoder with 8x the bytes its own decoder skips — frame_of_reference_coding.cpp:176-180 vs :262-268 — so such pages do not round-trip there)
@pytest.mark.parametrize("dt,lo,hi", [(np.int8, -30, 30), (np.int16, -8000, 8000), (np.int32, -2**29, 2**29), (np.int64, -2**61, 2**61),
                                       (np.int32, 2448000, 2451500), (np.int64, 0, 11), (np.int64, 90100, 10494951)])
@pytest.mark.parametrize("n", [1, 127, 128, 129, 5000, 70_001])
def test_builder_pages_are_the_reference_encoders_pages(oracle, dt, lo, hi, n):
    rng = np.random.default_rng(n + lo % 97)
    for kind in ("random", "sorted", "constant", "runs"):
        a = rng.integers(lo, hi, n).astype(dt)
        if kind == "sorted":
            a = np.sort(a)
        elif kind == "constant":
            a[:] = a[0]
        elif kind == "runs":  # sorted stretches (delta frames) between random ones
            for s in range(0, n, 384):
                a[s:s + 128] = np.sort(a[s:s + 128])
        want, want_pages = oracle_pages(oracle, a, 1 << 12)
        got, pages = PG.for_encode_column(torch.from_numpy(a), 1 << 12)
        assert pages == want_pages, kind
        assert bytes(got.numpy().tobytes()) == want, kind
        # and every page decodes to the values
        for off, nb, nr in pages:
            assert len(oracle.for_decode(want[off:off + nb], dt)) == nr
        dec = np.concatenate([oracle.for_decode(want[off:off + nb], dt) for off, nb, nr in pages])
        assert (dec == a).all()


def test_builder_overflow_frame_and_empty_column(oracle):
    a = np.array([5, 7, -2**31, 2**31 - 1, 0], np.int32)  # the range does not fit int32: the values themselves are stored (format 2)
    got, pages = PG.for_encode_column(torch.from_numpy(a))
    assert bytes(got.numpy().tobytes()) == oracle.for_encode(a)
    got, pages = PG.for_encode_column(torch.zeros(0, dtype=torch.int32))
    assert bytes(got.numpy().tobytes()) == oracle.for_encode(np.zeros(0, np.int32)) and pages == [(0, 5, 0)]


def test_plain_pages():
    a = np.arange(10, dtype=np.int64) * 3
    got, pages = PG.plain_encode_column(torch.from_numpy(a), 4)
    raw = got.numpy().tobytes()
    assert pages == [(0, 36, 4), (36, 36, 4), (72, 20, 2)]
    assert raw[:4] == (4).to_bytes(4, "little") and np.frombuffer(raw[4:36], np.int64).tolist() == [0, 3, 6, 9]
