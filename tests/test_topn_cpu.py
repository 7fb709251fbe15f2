"""Design check of the device TopN's composite sort key (starrocks_b200/csrc/srb_topn.cu, k_topn_encode), CPU only: a byte-for-byte
Python model of the encoding — one NULL-rank byte + the value's big-endian, sign-flipped (DESC: complemented) bytes per ORDER BY
column, arrival number as the tie-break — must order rows exactly as the oracle's restatement of the reference's multi-column
compare (oracle/orc_sort.cpp, pinned to be/test/exec/chunks_sorter_test.cpp). The CUDA kernel itself is covered by
tests/test_gpu_topn.py."""
import struct

import numpy as np
import pytest

from starrocks_b200 import capi as K


def encode_value(type_, v, desc):
    w = K.type_width(type_)
    if type_ == K.TYPE_DOUBLE:
        bits = struct.unpack("<Q", struct.pack("<d", float(v)))[0]
        if bits == 1 << 63:
            bits = 0  # -0.0 compares equal to +0.0
        u = (~bits) & (2**64 - 1) if bits >> 63 else bits | (1 << 63)
    else:
        u = (int(v) + (1 << (8 * w - 1))) & ((1 << (8 * w)) - 1)  # two's complement with the sign bit flipped
    if desc:
        u = (~u) & ((1 << (8 * w)) - 1)
    return u.to_bytes(w, "big")


def composite_key(chunk, row, keys, asc, nulls_first):
    out = b""
    for slot, a, nf in zip(keys, asc, nulls_first):
        c = chunk.columns[slot]
        isnull = c.nulls is not None and c.nulls[row]
        out += bytes([(0 if nf else 1) if isnull else (1 if nf else 0)])
        w = K.type_width(c.type)
        if isnull:
            out += b"\0" * w
        else:
            v = c.data[row]
            if w == 16:
                v = K.join128(int(v["lo"]), int(v["hi"]))
                if v >= 1 << 127:
                    v -= 1 << 128
            out += encode_value(c.type, v, not a)
    assert len(out) <= 32
    return out.ljust(32, b"\0")


def random_chunk(rng, n):
    def nulls(frac):
        return (rng.random(n) < frac).astype(np.uint8)
    d128 = K.to_i128_array([int(x) * (10**20) + int(y) for x, y in zip(rng.integers(-50, 50, n), rng.integers(0, 10**6, n))])
    dbl = rng.normal(0, 100, n)
    dbl[::17] = 0.0
    dbl[1::17] = -0.0  # equal to +0.0 in the reference's compare: ties fall back to the next key / arrival order
    dbl[2::17] = -1e-300
    return K.Chunk(n, {0: K.Column.from_ints(K.TYPE_INT, rng.integers(-5, 5, n).astype(np.int32), nulls(0.2)),
                       1: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(-2**62, 2**62, n)),
                       2: K.Column.from_ints(K.TYPE_DECIMAL128, d128, nulls(0.1), 38, 4),
                       3: K.Column.from_ints(K.TYPE_DOUBLE, dbl),
                       4: K.Column.from_ints(K.TYPE_SMALLINT, rng.integers(-300, 300, n).astype(np.int16)),
                       5: K.Column.from_ints(K.TYPE_INT, np.arange(n, dtype=np.int32))})


TOPN_CASES = [([0, 1], [0, 1], [0, 0]), ([2, 0], [0, 1], [1, 1]), ([3, 4], [1, 0], [1, 1]), ([0], [1], [1]), ([4, 0, 3], [0, 0, 1], [0, 1, 0]),
              ([2], [1], [0])]


@pytest.mark.parametrize("keys,asc,nf", TOPN_CASES)
def test_composite_key_order_is_the_reference_order(oracle, keys, asc, nf):
    rng = np.random.default_rng(31)
    n = 3000
    ch = random_chunk(rng, n)
    want = K.column_pyvalues(oracle.sort(ch, keys, asc, nf).columns[5])  # row ids in the reference's order (stable)
    order = sorted(range(n), key=lambda r: (composite_key(ch, r, keys, asc, nf), r))
    assert order == want


def test_bitonic_network_of_the_reduce_kernel_sorts():
    """the compare-exchange schedule of k_topn_reduce (one pair per thread: i = 2j(t / j) + t % j, partner i + j, direction by (i & k)),
    replayed with numpy on the kernel's tile size: every step touches each element once and the result is sorted"""
    tile, threads = 2048, 1024
    rng = np.random.default_rng(9)
    for trial in range(3):
        a = rng.integers(0, 50 if trial == 0 else 2**62, tile)  # trial 0: many ties
        a = a * tile + rng.permutation(tile)                    # the arrival number makes the order total, as in the kernel
        t = np.arange(threads)
        k = 2
        while k <= tile:
            j = k >> 1
            while j > 0:
                i = 2 * j * (t // j) + (t % j)
                p = i + j
                assert len(np.unique(np.concatenate([i, p]))) == tile
                asc = (i & k) == 0
                swap = np.where(asc, a[p] < a[i], a[i] < a[p])
                ai, ap = a[i].copy(), a[p].copy()
                a[i] = np.where(swap, ap, ai)
                a[p] = np.where(swap, ai, ap)
                j >>= 1
            k <<= 1
        assert (np.diff(a) > 0).all()
