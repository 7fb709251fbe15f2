"""ctypes wrapper of oracle/build/liboracle.so — the CPU restatement of the reference (TEST INFRASTRUCTURE).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

from starrocks_b200 import capi as K

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "build", "liboracle.so")


def build_oracle(force: bool = False) -> str:
    srcs = [os.path.join(ROOT, "oracle", f) for f in os.listdir(os.path.join(ROOT, "oracle")) if f.endswith((".cpp", ".h"))]
    srcs.append(os.path.join(ROOT, "include", "srb", "srb_c_api.h"))
    stale = force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    return LIB


class Oracle:
    def __init__(self):
        self.lib = C.CDLL(build_oracle())
        L = self.lib
        L.orc_last_error.restype = C.c_char_p
        L.orc_scan_agg.restype = C.c_void_p
        L.orc_scan_agg.argtypes = [C.POINTER(K.srb_scan_agg_plan), C.POINTER(K.srb_chunk), C.c_int, C.c_int]
        L.orc_agg_lookup_sequence.restype = C.c_int
        L.orc_agg_lookup_sequence.argtypes = [C.POINTER(K.srb_scan_agg_plan), C.c_int, C.POINTER(C.POINTER(K.srb_chunk)), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)]
        L.orc_sort.restype = C.c_void_p
        L.orc_sort.argtypes = [C.POINTER(K.srb_chunk), C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.POINTER(C.c_uint8), C.c_int64, C.c_int64]
        L.orc_scan_agg_streaming.restype = C.c_void_p
        L.orc_scan_agg_streaming.argtypes = [C.POINTER(K.srb_scan_agg_plan), C.POINTER(K.srb_chunk), C.c_int, C.c_int, C.POINTER(C.c_int64)]
        L.orc_join.restype = C.c_void_p
        L.orc_join.argtypes = [C.POINTER(K.srb_join_plan), C.POINTER(K.srb_chunk), C.POINTER(K.srb_chunk)]
        L.orc_join_mt.restype = C.c_void_p
        L.orc_join_mt.argtypes = [C.POINTER(K.srb_join_plan), C.POINTER(K.srb_chunk), C.POINTER(K.srb_chunk), C.c_int]
        L.orc_result_chunk.restype = C.POINTER(K.srb_chunk)
        L.orc_result_chunk.argtypes = [C.c_void_p]
        L.orc_result_counters.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.orc_result_free.argtypes = [C.c_void_p]
        L.orc_from_date.restype = C.c_int32
        L.orc_checked_binop.argtypes = [C.c_int, C.c_int] + [C.c_int64] * 4 + [C.POINTER(C.c_int64)] * 2
        L.orc_decimal_div_integer.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.orc_div_round128.argtypes = [C.c_int64] * 4 + [C.POINTER(C.c_int64)] * 2
        L.orc_div_round64.restype = C.c_int64
        L.orc_div_round64.argtypes = [C.c_int64, C.c_int64]
        L.orc_join_key_hash32.restype = C.c_uint32
        L.orc_join_key_hash32.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_join_key_hash64.restype = C.c_uint32
        L.orc_join_key_hash64.argtypes = [C.c_uint64, C.c_uint32]
        L.orc_join_key_hash_bytes.restype = C.c_uint32
        L.orc_join_key_hash_bytes.argtypes = [C.c_char_p, C.c_int32, C.c_uint32]
        L.orc_calc_bucket_size.restype = C.c_uint32
        L.orc_calc_bucket_size.argtypes = [C.c_uint32]
        L.orc_join_build_raw.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.POINTER(C.c_uint32), C.c_void_p, C.c_void_p]
        L.orc_crc_hash_32.restype = C.c_uint32
        L.orc_crc_hash_32.argtypes = [C.c_char_p, C.c_int32, C.c_uint32]
        L.orc_crc_hash_64.restype = C.c_uint64
        L.orc_crc_hash_64.argtypes = [C.c_char_p, C.c_int32, C.c_uint64]
        L.orc_phmap_mix8.restype = C.c_uint64
        L.orc_phmap_mix8.argtypes = [C.c_uint64, C.c_int]
        L.orc_hash128_with_seed.restype = C.c_uint64
        L.orc_hash128_with_seed.argtypes = [C.c_int64, C.c_int64, C.c_int]
        L.orc_fnv_hash.restype = C.c_uint32
        L.orc_fnv_hash.argtypes = [C.c_char_p, C.c_int32, C.c_uint32]
        L.orc_xorshift32.restype = C.c_uint32
        L.orc_xorshift32.argtypes = [C.c_uint32]
        L.orc_shuffle_channel.restype = C.c_uint32
        L.orc_shuffle_channel.argtypes = [C.c_char_p, C.c_int32, C.c_uint32]
        L.orc_rf_create.restype = C.c_void_p
        L.orc_rf_create.argtypes = [C.c_uint64, C.c_int]
        L.orc_rf_free.argtypes = [C.c_void_p]
        L.orc_rf_insert_hash.argtypes = [C.c_void_p, C.c_uint64]
        L.orc_rf_test_hash.argtypes = [C.c_void_p, C.c_uint64]
        L.orc_rf_insert_column.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_void_p]
        L.orc_rf_evaluate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_rf_test_data.argtypes = [C.c_void_p, C.c_int64]
        L.orc_rf_info.restype = C.c_int64
        L.orc_rf_info.argtypes = [C.c_void_p, C.c_int]
        L.orc_rf_directory.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_rf_choose_bitset.argtypes = [C.c_int64, C.c_int64, C.c_uint64, C.c_uint64, C.c_uint64]
        L.orc_bitset_create.restype = C.c_void_p
        L.orc_bitset_create.argtypes = [C.c_int64, C.c_int64]
        L.orc_bitset_free.argtypes = [C.c_void_p]
        L.orc_bitset_insert_column.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_void_p]
        L.orc_bitset_evaluate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_bitset_bytes.restype = C.c_int64
        L.orc_bitset_bytes.argtypes = [C.c_void_p]
        L.orc_bitset_data.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_for_encode.restype = C.c_int64
        L.orc_for_encode.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int64]
        L.orc_for_decode.restype = C.c_int64
        L.orc_for_decode.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int64]
        L.orc_for_decode_pages.restype = C.c_int64
        L.orc_for_decode_pages.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_int]
        L.orc_bits.restype = C.c_uint8
        L.orc_bits.argtypes = [C.c_int64, C.c_int64, C.c_int]
        if L.orc_has_hw_crc():
            L.orc_hw_crc32c_u64.restype = C.c_uint32
            L.orc_hw_crc32c_u64.argtypes = [C.c_uint32, C.c_uint64]

    def _err(self):
        return self.lib.orc_last_error().decode()

    def scan_agg(self, plan, chunk: K.Chunk, n_threads: int = 1, chunk_size: int = 4096, with_counters=False):
        cch, keep = chunk.to_c()
        r = self.lib.orc_scan_agg(C.byref(plan.c), C.byref(cch), n_threads, chunk_size)
        if not r:
            raise RuntimeError("oracle: " + self._err())
        try:
            out = K.decode_chunk(self.lib.orc_result_chunk(r).contents)
            a, b = C.c_int64(), C.c_int64()
            self.lib.orc_result_counters(r, C.byref(a), C.byref(b))
        finally:
            self.lib.orc_result_free(r)
        return (out, a.value, b.value) if with_counters else out

    def scan_agg_streaming(self, plan, chunk: K.Chunk, mode: int, chunk_size: int = 4096):
        """streaming pre-aggregation sink (TStreamingPreaggregationMode: 0 AUTO, 1 FORCE_STREAMING, 2 FORCE_PREAGGREGATION);
        returns (intermediate chunk, stats dict)"""
        cch, keep = chunk.to_c()
        stats = (C.c_int64 * 9)()
        r = self.lib.orc_scan_agg_streaming(C.byref(plan.c), C.byref(cch), chunk_size, mode, stats)
        if not r:
            raise RuntimeError("oracle: " + self._err())
        try:
            out = K.decode_chunk(self.lib.orc_result_chunk(r).contents)
        finally:
            self.lib.orc_result_free(r)
        names = ["init_preagg", "adjust", "pass_through", "force_preagg", "preagg", "selective_preagg"]
        return out, {"rows_streamed": stats[0], "groups": stats[1], "chunks": dict(zip(names, list(stats[2:8]))), "final_state": names[stats[8]]}

    def agg_lookup_sequence(self, plan, chunks, ops):
        """ops[i]: 0 build_hash_map, 1 build_hash_map_with_selection, 2 build_hash_map_with_selection_and_allocation on chunks[i];
        returns the not-found flags of every chunk run with op 1 or 2"""
        cs = [c.to_c() for c in chunks]
        arr = (C.POINTER(K.srb_chunk) * len(chunks))(*[C.pointer(c[0]) for c in cs])
        n_out = sum(c.n_rows for c, o in zip(chunks, ops) if o)
        out = (C.c_uint8 * max(1, n_out))()
        if self.lib.orc_agg_lookup_sequence(C.byref(plan.c), len(chunks), arr, (C.c_int32 * len(ops))(*ops), out):
            raise RuntimeError("oracle: " + self._err())
        flat, res, pos = list(out)[:n_out], [], 0
        for c, o in zip(chunks, ops):
            if o:
                res.append(flat[pos:pos + c.n_rows])
                pos += c.n_rows
        return res

    def sort(self, chunk: K.Chunk, key_slots, is_asc, nulls_first, offset: int = 0, limit: int = -1) -> K.Chunk:
        """ORDER BY ... OFFSET ... LIMIT (ChunksSorterTopn / ChunksSorterFullSort): every column, rows in sorted order"""
        cch, keep = chunk.to_c()
        n = len(key_slots)
        r = self.lib.orc_sort(C.byref(cch), n, (C.c_int32 * n)(*key_slots), (C.c_uint8 * n)(*[int(bool(x)) for x in is_asc]),
                              (C.c_uint8 * n)(*[int(bool(x)) for x in nulls_first]), offset, limit)
        if not r:
            raise RuntimeError("oracle: " + self._err())
        try:
            return K.decode_chunk(self.lib.orc_result_chunk(r).contents)
        finally:
            self.lib.orc_result_free(r)

    def join(self, plan, build: K.Chunk, probe: K.Chunk, n_threads: int = 1):
        """n_threads = the probe side's degree of parallelism (drivers over one shared table); same rows, same order"""
        b, k1 = build.to_c()
        p, k2 = probe.to_c()
        r = self.lib.orc_join_mt(C.byref(plan.c), C.byref(b), C.byref(p), n_threads)
        if not r:
            raise RuntimeError("oracle: " + self._err())
        try:
            return K.decode_chunk(self.lib.orc_result_chunk(r).contents)
        finally:
            self.lib.orc_result_free(r)

    # ---- segment data pages (oracle/orc_page.cpp) -------------------------------------------------------------------------------
    def for_encode(self, values) -> bytes:
        """one frame-of-reference page of a numpy array of a signed integer type (int8..int64, or capi.I128)"""
        import numpy as np
        a = np.ascontiguousarray(values)
        cap = a.nbytes * 9 + 4096 + 2 * (len(a) // 128 + 1) * (a.dtype.itemsize + 2)
        out = np.empty(cap, np.uint8)
        n = self.lib.orc_for_encode(a.ctypes.data, a.dtype.itemsize, len(a), out.ctypes.data, cap)
        if n <= 0:
            raise RuntimeError("oracle: for_encode")
        return out[:n].tobytes()

    def for_decode(self, page: bytes, dtype):
        import numpy as np
        buf = np.frombuffer(page, np.uint8)
        n = int.from_bytes(page[-4:], "little") if len(page) >= 5 else 0
        out = np.zeros(max(n, 1), dtype)
        got = self.lib.orc_for_decode(buf.ctypes.data, len(page), np.dtype(dtype).itemsize, out.ctypes.data, len(out))
        if got < 0:
            raise RuntimeError("oracle: corrupt frame-of-reference page")
        return out[:got]

    def for_decode_pages(self, data, offs, rows, dtype, n_threads: int = 1):
        """pages [offs[p], offs[p+1]) of `data` (uint8 array) whose values start at rows[p] -> the decoded column"""
        import numpy as np
        offs = np.ascontiguousarray(offs, np.int64)
        rows = np.ascontiguousarray(rows, np.int64)
        out = np.empty(int(rows[-1]), dtype)
        bad = self.lib.orc_for_decode_pages(data.ctypes.data, offs.ctypes.data, rows.ctypes.data, len(offs) - 1, np.dtype(dtype).itemsize, out.ctypes.data, n_threads)
        if bad:
            raise RuntimeError("oracle: corrupt frame-of-reference page")
        return out

    def bits(self, v: int, width: int) -> int:
        lo, hi = K.split128(v)
        return self.lib.orc_bits(lo, hi, width)

    def checked_binop(self, width, op, a, b):
        alo, ahi = K.split128(a)
        blo, bhi = K.split128(b)
        rlo, rhi = C.c_int64(), C.c_int64()
        ov = self.lib.orc_checked_binop(width, op, alo, ahi, blo, bhi, C.byref(rlo), C.byref(rhi))
        v = K.join128(rlo.value, rhi.value)
        if v >= 1 << 127:
            v -= 1 << 128
        return bool(ov), v

    def decimal_div_integer(self, s, count, scale):
        lo, hi = K.split128(s)
        rlo, rhi = C.c_int64(), C.c_int64()
        self.lib.orc_decimal_div_integer(lo, hi, count, scale, C.byref(rlo), C.byref(rhi))
        v = K.join128(rlo.value, rhi.value)
        return v - (1 << 128) if v >= 1 << 127 else v


class OracleBitsetFilter:
    """oracle/orc_rf.h RuntimeBitsetFilter: one bit per value of [min, max] behind the min / max test; NULL rows pass iff a NULL was inserted"""

    def __init__(self, oracle: Oracle, mn: int, mx: int):
        self.L = oracle.lib
        self.h = self.L.orc_bitset_create(mn, mx)

    @staticmethod
    def chosen(oracle: Oracle, mn: int, mx: int, row_count: int, l2_cache: int, l3_cache: int = 0) -> bool:
        """try_create_runtime_bitset_filter's rule"""
        return bool(oracle.lib.orc_rf_choose_bitset(mn, mx, row_count, l2_cache, l3_cache))

    def insert_column(self, col: K.Column):
        import numpy as np
        data = np.ascontiguousarray(col.data)
        nulls = None if col.nulls is None else np.ascontiguousarray(col.nulls)
        self.L.orc_bitset_insert_column(self.h, data.ctypes.data, data.dtype.itemsize, len(data), None if nulls is None else nulls.ctypes.data)
        return self

    def evaluate(self, col: K.Column, selection=None):
        import numpy as np
        data = np.ascontiguousarray(col.data)
        nulls = None if col.nulls is None else np.ascontiguousarray(col.nulls)
        sel = np.ones(len(data), np.uint8) if selection is None else np.ascontiguousarray(selection, dtype=np.uint8).copy()
        self.L.orc_bitset_evaluate(self.h, data.ctypes.data, data.dtype.itemsize, len(data), None if nulls is None else nulls.ctypes.data, sel.ctypes.data)
        return sel

    def data(self):
        import numpy as np
        out = np.empty(self.L.orc_bitset_bytes(self.h), dtype=np.uint8)
        self.L.orc_bitset_data(self.h, out.ctypes.data)
        return out

    def close(self):
        if self.h:
            self.L.orc_bitset_free(self.h)
            self.h = None


class OracleRuntimeFilter:
    """oracle/orc_rf.h: SimdBlockFilter + min / max / has_null over integer keys (the CPU restatement the GPU filter is checked against)"""

    def __init__(self, oracle: Oracle, nums: int, mix: bool = False):
        self.L = oracle.lib
        self.h = self.L.orc_rf_create(nums, 1 if mix else 0)

    def insert_column(self, col: K.Column):
        import numpy as np
        data = np.ascontiguousarray(col.data)
        nulls = None if col.nulls is None else np.ascontiguousarray(col.nulls)
        self.L.orc_rf_insert_column(self.h, data.ctypes.data, data.dtype.itemsize, len(data), None if nulls is None else nulls.ctypes.data)
        return self

    def evaluate(self, col: K.Column, selection=None):
        import numpy as np
        data = np.ascontiguousarray(col.data)
        nulls = None if col.nulls is None else np.ascontiguousarray(col.nulls)
        sel = np.ones(len(data), np.uint8) if selection is None else np.ascontiguousarray(selection, dtype=np.uint8).copy()
        self.L.orc_rf_evaluate(self.h, data.ctypes.data, data.dtype.itemsize, len(data), None if nulls is None else nulls.ctypes.data, sel.ctypes.data)
        return sel

    def info(self):
        g = lambda w: self.L.orc_rf_info(self.h, w)
        return {"n_buckets": g(0), "min": g(1), "max": g(2), "has_null": bool(g(3)), "inserted": g(4)}

    def directory(self):
        import numpy as np
        out = np.empty((self.info()["n_buckets"], 8), dtype=np.uint32)
        self.L.orc_rf_directory(self.h, out.ctypes.data)
        return out

    def close(self):
        if self.h:
            self.L.orc_rf_free(self.h)
            self.h = None
