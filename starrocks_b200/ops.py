"""Host-side mirror of the reference's pipeline operator interface over libsrb.so (ctypes, no torch types).

Method names and meaning follow starrocks::pipeline::Operator (be/src/exec/pipeline/operator.h:43-344):
push_chunk / pull_chunk / set_finishing / has_output / need_input / is_finished / set_cancelled / close.
`ScanAggOperator` stands where OlapScan predicate evaluation + Project + AggregateBlockingSink/Source sit;
`HashJoinOperator` carries the build side (HashJoinBuildOperator) and exposes the probe side as `.probe`
(HashJoinProbeOperator). Errors raise `SrbError` carrying the TStatusCode (Status in the reference).

There is no CPU fallback: importing this module works anywhere, but creating a Context needs libsrb.so and a B200.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

from . import capi as K

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libsrb.so")
_lib = None

EXPORTS = [
    "srb_ctx_create", "srb_ctx_destroy", "srb_last_error", "srb_ctx_device_info", "srb_ctx_trim", "srb_ctx_alloc_stats", "srb_op_stream", "srb_jit_wait", "srb_jit_stats", "srb_jit_check_plan", "srb_runtime_filter_kind", "srb_runtime_filter_bitset",
    "srb_scan_agg_create", "srb_join_create", "srb_op_push_chunk", "srb_op_push_pages", "srb_op_set_finishing_merged", "srb_op_set_finishing", "srb_op_has_output",
    "srb_op_need_input", "srb_op_is_finished", "srb_op_pull_chunk", "srb_op_pull_all", "srb_op_pull_device",
    "srb_op_cancel", "srb_op_close", "srb_op_sync", "srb_op_counters", "srb_op_profile",
    "srb_join_probe_push_chunk", "srb_join_probe_set_finishing", "srb_join_probe_has_output",
    "srb_join_probe_need_input", "srb_join_probe_is_finished", "srb_join_probe_pull_chunk",
    "srb_join_probe_pull_all", "srb_join_probe_pull_device",
    "srb_device_alloc", "srb_device_free", "srb_pinned_alloc", "srb_pinned_free", "srb_memcpy_h2d", "srb_memcpy_d2h",
    "srb_partition_create", "srb_partition_run", "srb_partition_destroy", "srb_partition_run_noalloc",
    "srb_comm_unique_id", "srb_comm_create", "srb_comm_destroy", "srb_comm_abort", "srb_comm_check", "srb_comm_stream", "srb_comm_all_to_all", "srb_comm_all_gather", "srb_comm_shuffle", "srb_comm_counters", "srb_comm_merge_runtime_filter",
    "srb_join_build_runtime_filter", "srb_op_add_runtime_filter", "srb_op_set_streaming_mode", "srb_op_set_streaming_memory_limit", "srb_op_streaming_stats", "srb_topn_create", "srb_runtime_filter_info", "srb_runtime_filter_directory", "srb_runtime_filter_destroy",
]


class SrbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[status {code}] {msg}")
        self.code = code


def lib() -> C.CDLL:
    """Loads libsrb.so (built by `python -m starrocks_b200.build`). Fails loudly when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise ImportError(f"{_LIB_PATH} not built: run `python -m starrocks_b200.build` (there is no CPU fallback)")
    L = C.CDLL(_LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    L.srb_last_error.restype = C.c_char_p
    L.srb_ctx_create.argtypes = [C.POINTER(K.srb_config), C.POINTER(vp)]
    L.srb_ctx_destroy.argtypes = [vp]
    L.srb_ctx_destroy.restype = None
    L.srb_ctx_device_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i64), C.POINTER(i32)]
    L.srb_ctx_trim.argtypes = [vp]
    L.srb_ctx_alloc_stats.argtypes = [vp] + [C.POINTER(i64)] * 5
    L.srb_op_stream.argtypes = [vp]
    L.srb_op_stream.restype = vp
    L.srb_jit_wait.argtypes = []
    L.srb_jit_stats.argtypes = [C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), C.c_char_p, i64]
    L.srb_jit_check_plan.argtypes = [C.POINTER(K.srb_scan_agg_plan), C.POINTER(K.srb_chunk), C.c_char_p, i64]
    L.srb_scan_agg_create.argtypes = [vp, C.POINTER(K.srb_scan_agg_plan), C.POINTER(vp)]
    L.srb_join_create.argtypes = [vp, C.POINTER(K.srb_join_plan), C.POINTER(vp)]
    L.srb_topn_create.argtypes = [vp, C.POINTER(K.srb_topn_plan), C.POINTER(vp)]
    for f in ("srb_op_push_chunk", "srb_join_probe_push_chunk"):
        getattr(L, f).argtypes = [vp, C.POINTER(K.srb_chunk)]
    L.srb_op_set_finishing_merged.argtypes = [vp, vp]
    from . import pages as PG
    L.srb_op_push_pages.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(PG.srb_page_column), i64]
    for f in ("srb_op_set_finishing", "srb_join_probe_set_finishing", "srb_op_cancel", "srb_op_sync", "srb_op_has_output",
              "srb_op_need_input", "srb_op_is_finished", "srb_join_probe_has_output", "srb_join_probe_need_input",
              "srb_join_probe_is_finished"):
        getattr(L, f).argtypes = [vp]
    for f in ("srb_op_pull_chunk", "srb_op_pull_all", "srb_op_pull_device", "srb_join_probe_pull_chunk",
              "srb_join_probe_pull_all", "srb_join_probe_pull_device"):
        getattr(L, f).argtypes = [vp, C.POINTER(K.srb_chunk)]
    L.srb_op_close.argtypes = [vp]
    L.srb_op_close.restype = None
    L.srb_op_counters.argtypes = [vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]
    L.srb_op_profile.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(i64), C.POINTER(i64), C.POINTER(i32)]
    L.srb_device_alloc.argtypes = [vp, i64, C.POINTER(vp)]
    L.srb_device_free.argtypes = [vp, vp]
    L.srb_device_free.restype = None
    L.srb_pinned_alloc.argtypes = [vp, i64, C.POINTER(vp)]
    L.srb_pinned_free.argtypes = [vp, vp]
    L.srb_pinned_free.restype = None
    L.srb_memcpy_h2d.argtypes = [vp, vp, vp, i64]
    L.srb_memcpy_d2h.argtypes = [vp, vp, vp, i64]
    L.srb_partition_create.argtypes = [vp, i32, i32, C.POINTER(vp)]
    L.srb_partition_run.argtypes = [vp, C.POINTER(K.srb_chunk), i32, vp, C.POINTER(K.srb_chunk), C.POINTER(i64)]
    L.srb_partition_destroy.argtypes = [vp]
    L.srb_partition_run_noalloc.argtypes = [i32, vp, i32, C.POINTER(K.srb_chunk), i32, vp, C.POINTER(K.srb_chunk), vp]
    L.srb_partition_destroy.restype = None
    L.srb_comm_unique_id.argtypes = [vp]
    L.srb_comm_create.argtypes = [vp, vp, C.POINTER(vp)]
    L.srb_comm_destroy.argtypes = [vp]
    L.srb_comm_destroy.restype = None
    L.srb_comm_abort.argtypes = [vp]
    L.srb_comm_check.argtypes = [vp]
    L.srb_comm_stream.argtypes = [vp]
    L.srb_comm_stream.restype = vp
    L.srb_comm_all_to_all.argtypes = [vp, C.POINTER(K.srb_chunk), C.POINTER(i64), C.POINTER(K.srb_chunk), C.POINTER(i64)]
    L.srb_comm_all_gather.argtypes = [vp, C.POINTER(K.srb_chunk), C.POINTER(K.srb_chunk)]
    L.srb_comm_shuffle.argtypes = [vp, C.POINTER(K.srb_chunk), C.c_int32, C.POINTER(C.c_int32), C.POINTER(K.srb_chunk), C.POINTER(i64)]
    L.srb_comm_counters.argtypes = [vp, C.POINTER(i64)]
    L.srb_comm_merge_runtime_filter.argtypes = [vp, vp]
    L.srb_join_build_runtime_filter.argtypes = [vp, C.c_int32, i64, C.c_int32, C.POINTER(vp)]
    L.srb_op_add_runtime_filter.argtypes = [vp, vp, C.c_int32]
    L.srb_op_set_streaming_mode.argtypes = [vp, C.c_int32]
    L.srb_op_set_streaming_memory_limit.argtypes = [vp, C.c_int64]
    L.srb_op_streaming_stats.argtypes = [vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int32)]
    L.srb_runtime_filter_info.argtypes = [vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), C.POINTER(C.c_int32), C.POINTER(i64)]
    L.srb_runtime_filter_directory.argtypes = [vp, vp, i64]
    L.srb_runtime_filter_kind.argtypes = [vp, C.POINTER(i32), C.POINTER(i64)]
    L.srb_runtime_filter_bitset.argtypes = [vp, vp, i64]
    L.srb_runtime_filter_destroy.argtypes = [vp]
    L.srb_runtime_filter_destroy.restype = None
    _lib = L
    return L


def _check(st: int):
    if st not in (K.OK, K.END_OF_FILE):
        raise SrbError(st, lib().srb_last_error().decode())
    return st


def jit_wait():
    """blocks until every run-time kernel compilation started so far has finished (srb_jit_wait)"""
    _check(lib().srb_jit_wait())


def jit_stats() -> dict:
    """process-wide counts of run-time compiled kernel shapes (srb_jit_stats)"""
    v = [C.c_int64() for _ in range(4)]
    err = C.create_string_buffer(4096)
    _check(lib().srb_jit_stats(*[C.byref(x) for x in v], err, len(err)))
    return {"compiled": v[0].value, "from_disk_cache": v[1].value, "failed": v[2].value, "pending": v[3].value, "last_error": err.value.decode(errors="replace")}


def jit_check_plan(plan, first_chunk: K.Chunk) -> str:
    """compiles the kernel shape of `plan` for chunks typed like `first_chunk` on the calling thread (no GPU needed); returns its description,
    raises SrbError NOT_IMPLEMENTED when the plan is outside what the static kernel covers (srb_jit_check_plan)"""
    cch, keep = first_chunk.to_c()
    out = C.create_string_buffer(512)
    _check(lib().srb_jit_check_plan(C.byref(plan.c), C.byref(cch), out, len(out)))
    return out.value.decode()


class Context:
    """srb_ctx: one per process/GPU (ExecEnv + RuntimeState stand-in: device ordinal, chunk_size, rank/world)."""

    def __init__(self, device: int = 0, chunk_size: int = 4096, staging_rows: int = 0, rank: int = 0, world_size: int = 1,
                 mem_limit: int = 0):
        self.L = lib()
        cfg = K.srb_config(device, chunk_size, staging_rows, rank, world_size, mem_limit)
        h = C.c_void_p()
        _check(self.L.srb_ctx_create(C.byref(cfg), C.byref(h)))
        self.h = h
        self.chunk_size = chunk_size
        self.rank, self.world_size = rank, world_size

    def device_info(self):
        a, b, c = C.c_int32(), C.c_int64(), C.c_int32()
        _check(self.L.srb_ctx_device_info(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return {"n_sms": a.value, "hbm_bytes": b.value, "sm_arch": c.value}

    def alloc_stats(self) -> dict:
        """cudaMalloc / cudaFree calls and microseconds of this context, and the bytes parked in its pool (srb_ctx_alloc_stats)"""
        v = [C.c_int64() for _ in range(5)]
        _check(self.L.srb_ctx_alloc_stats(self.h, *[C.byref(x) for x in v]))
        return dict(zip(("n_malloc", "n_free", "malloc_us", "free_us", "pooled_bytes"), [x.value for x in v]))

    def trim(self):
        """give the idle pooled device buffers back to the driver"""
        _check(self.L.srb_ctx_trim(self.h))

    def close(self):
        if self.h:
            self.L.srb_ctx_destroy(self.h)
            self.h = None

    # plain cudaMalloc / cudaMemcpy helpers so a harness needs no other CUDA binding
    def device_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        _check(self.L.srb_device_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def device_free(self, p: int):
        self.L.srb_device_free(self.h, p)

    def to_device(self, arr) -> int:
        import numpy as np
        a = np.ascontiguousarray(arr)
        p = self.device_alloc(max(a.nbytes, 16))
        if a.nbytes:
            _check(self.L.srb_memcpy_h2d(self.h, p, a.ctypes.data, a.nbytes))
        return p

    def device_chunk(self, chunk: K.Chunk):
        """Copies a host chunk to the device; returns (device Chunk, list of device pointers to free)."""
        out = K.Chunk(chunk.n_rows)
        ptrs = []
        for slot, c in chunk.columns.items():
            d = self.to_device(c.data)
            ptrs.append(d)
            n = o = None
            if c.nulls is not None:
                n = self.to_device(c.nulls)
                ptrs.append(n)
            if c.offsets is not None:
                o = self.to_device(c.offsets)
                ptrs.append(o)
            out.columns[slot] = K.Column(c.type, d, n, o, c.precision, c.scale, K.MEM_DEVICE, c.n_bytes)
        return out, ptrs

    def fetch_chunk(self, chunk: K.Chunk) -> K.Chunk:
        """Copies a device-resident chunk (fixed-width columns) back into numpy columns."""
        import numpy as np
        out = K.Chunk(chunk.n_rows)
        n = chunk.n_rows
        for slot, c in chunk.columns.items():
            w = K.type_width(c.type)
            raw = np.zeros(n * w, np.uint8)
            if n:
                _check(self.L.srb_memcpy_d2h(self.h, raw.ctypes.data, c.data, n * w))
            data = raw.view(K.I128) if w == 16 else (raw.view(np.float64) if c.type == K.TYPE_DOUBLE else raw.view(K._NP[w]))
            nulls = None
            if c.nulls:
                nulls = np.zeros(n, np.uint8)
                if n:
                    _check(self.L.srb_memcpy_d2h(self.h, nulls.ctypes.data, c.nulls, n))
            out.columns[slot] = K.Column(c.type, data, nulls, None, c.precision, c.scale)
        return out


def _raw_chunk(ch: K.srb_chunk) -> K.Chunk:
    """A C chunk whose columns are device pointers -> Chunk of MEM_DEVICE columns (no copy)."""
    out = K.Chunk(int(ch.n_rows))
    for i in range(ch.n_cols):
        c = ch.cols[i]
        out.columns[int(ch.slot_ids[i])] = K.Column(c.type, c.data, c.nulls, c.offsets, c.precision, c.scale, K.MEM_DEVICE, c.n_bytes)
    return out


class _Operator:
    _push = "srb_op_push_chunk"
    _finish = "srb_op_set_finishing"
    _has_output = "srb_op_has_output"
    _need_input = "srb_op_need_input"
    _is_finished = "srb_op_is_finished"
    _pull = "srb_op_pull_chunk"
    _pull_all = "srb_op_pull_all"
    _pull_device = "srb_op_pull_device"

    def __init__(self, ctx: Context, handle):
        self.ctx, self.L, self.h = ctx, ctx.L, handle

    def push_chunk(self, chunk: K.Chunk):
        cch, keep = chunk.to_c()
        _check(getattr(self.L, self._push)(self.h, C.byref(cch)))

    def set_finishing(self):
        _check(getattr(self.L, self._finish)(self.h))

    def has_output(self) -> bool:
        return bool(getattr(self.L, self._has_output)(self.h))

    def need_input(self) -> bool:
        return bool(getattr(self.L, self._need_input)(self.h))

    def is_finished(self) -> bool:
        return bool(getattr(self.L, self._is_finished)(self.h))

    def pull_chunk(self) -> Optional[K.Chunk]:
        """At most chunk_size rows; None at end of stream (Status::EndOfFile)."""
        out = K.srb_chunk()
        st = _check(getattr(self.L, self._pull)(self.h, C.byref(out)))
        return None if st == K.END_OF_FILE else K.decode_chunk(out)

    def pull_all(self) -> K.Chunk:
        out = K.srb_chunk()
        st = _check(getattr(self.L, self._pull_all)(self.h, C.byref(out)))
        return K.Chunk(0) if st == K.END_OF_FILE else K.decode_chunk(out)

    def pull_device(self) -> Optional[K.Chunk]:
        out = K.srb_chunk()
        st = _check(getattr(self.L, self._pull_device)(self.h, C.byref(out)))
        return None if st == K.END_OF_FILE else _raw_chunk(out)

    def set_cancelled(self):
        _check(self.L.srb_op_cancel(self.h))

    def sync(self):
        _check(self.L.srb_op_sync(self.h))

    def counters(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        _check(self.L.srb_op_counters(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return {"rows_in": a.value, "rows_selected": b.value, "kernel_launches": c.value}

    def close(self):
        """Operator::close (operator.h:92): idempotent"""
        if self.h:
            self.L.srb_op_close(self.h)
            self.h = None


class ScanAggOperator(_Operator):
    """scan-filter -> project -> aggregate for one driver (see include/srb/srb_c_api.h: srb_scan_agg_create)."""

    def __init__(self, ctx: Context, plan):
        h = C.c_void_p()
        _check(ctx.L.srb_scan_agg_create(ctx.h, C.byref(plan.c), C.byref(h)))
        super().__init__(ctx, h)
        self._plan = plan

    def set_finishing_merged(self, comm: "Communicator"):
        """collective over the communicator's ranks: set_finishing + gather of the ranks' aggregate states on the device + merge; every
        rank then pulls the global result (srb_op_set_finishing_merged)"""
        _check(self.L.srb_op_set_finishing_merged(self.h, comm.h))

    def push_pages_prepared(self, prepared):
        """prepared = pages.prepare(columns): the C structures built once (a hot loop pushes the same page sets again and again)"""
        n, slots, cols, keep, rows = prepared
        _check(self.L.srb_op_push_pages(self.h, n, slots, cols, rows))

    def push_pages(self, columns):
        """columns: {slot: pages.PageColumn}, every column's encoded data pages over the same row range (srb_op_push_pages): decoded on
        the device in front of the scan (SegmentIterator -> page decoders on the CPU in the reference)"""
        from . import pages as PG
        n, slots, cols, keep = PG.to_c(columns)
        rows = next(iter(columns.values())).n_rows() if columns else 0
        _check(self.L.srb_op_push_pages(self.h, n, slots, cols, rows))

    def add_runtime_filter(self, rf: "RuntimeFilter", slot_id: int):
        """scan-filter plans only, before the first push_chunk (Operator::eval_runtime_bloom_filters)"""
        _check(self.L.srb_op_add_runtime_filter(self.h, rf.h, slot_id))
        self._rfs = getattr(self, "_rfs", []) + [rf]

    def set_streaming_mode(self, mode: int):
        """first phase of a two-phase GROUP BY as a streaming sink (AggregateStreamingSinkOperator): capi.STREAMING_AUTO /
        STREAMING_FORCE_STREAMING / STREAMING_FORCE_PREAGGREGATION (TStreamingPreaggregationMode); before the first push_chunk"""
        _check(self.L.srb_op_set_streaming_mode(self.h, mode))

    def set_streaming_memory_limit(self, nbytes: int):
        """STREAMING_LIMITED_MEM: bytes of group table from which on every row passes through (LimitedMemAggState)"""
        _check(self.L.srb_op_set_streaming_memory_limit(self.h, nbytes))

    def streaming_stats(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int32()
        _check(self.L.srb_op_streaming_stats(self.h, C.byref(a), C.byref(b), C.byref(c)))
        names = ["init_preagg", "adjust", "pass_through", "force_preagg", "preagg", "selective_preagg"]
        return {"rows_streamed": a.value, "groups": b.value, "state": names[c.value]}

    def stream(self) -> int:
        return self.L.srb_op_stream(self.h)

    def profile(self):
        ms, b, sl, sh = C.c_double(), C.c_int64(), C.c_int64(), C.c_int32()
        _check(self.L.srb_op_profile(self.h, C.byref(ms), C.byref(b), C.byref(sl), C.byref(sh)))
        return {"kernel_ms": ms.value, "batches": b.value, "static_launches": sl.value, "shape": sh.value}


class RuntimeFilter:
    """A join's runtime filter (bloom directory on the device + min / max / has_null), see srb_join_build_runtime_filter."""

    def __init__(self, ctx: Context, handle):
        self.ctx, self.L, self.h = ctx, ctx.L, handle

    def info(self):
        nb, mn, mx, ins, hn = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64(), C.c_int32()
        _check(self.L.srb_runtime_filter_info(self.h, C.byref(nb), C.byref(mn), C.byref(mx), C.byref(hn), C.byref(ins)))
        return {"n_buckets": nb.value, "min": mn.value, "max": mx.value, "has_null": bool(hn.value), "inserted": ins.value}

    def kind(self) -> int:
        """K.RF_KIND_BLOOM or K.RF_KIND_BITSET: what the probe side tests"""
        k, nb = C.c_int32(), C.c_int64()
        _check(self.L.srb_runtime_filter_kind(self.h, C.byref(k), C.byref(nb)))
        return k.value

    def bitset(self):
        """the bitset as a numpy uint32 array (bit (x - min) of a little-endian bit string)"""
        import numpy as np
        k, nb = C.c_int32(), C.c_int64()
        _check(self.L.srb_runtime_filter_kind(self.h, C.byref(k), C.byref(nb)))
        out = np.empty(nb.value // 4, dtype=np.uint32)
        _check(self.L.srb_runtime_filter_bitset(self.h, out.ctypes.data, out.nbytes))
        return out

    def directory(self):
        """the directory as a numpy array [n_buckets, 8] of uint32"""
        import numpy as np
        nb = self.info()["n_buckets"]
        out = np.empty((nb, 8), dtype=np.uint32)
        _check(self.L.srb_runtime_filter_directory(self.h, out.ctypes.data, out.nbytes))
        return out

    def close(self):
        if self.h:
            self.L.srb_runtime_filter_destroy(self.h)
            self.h = None


class _ProbeSide(_Operator):
    _push = "srb_join_probe_push_chunk"
    _finish = "srb_join_probe_set_finishing"
    _has_output = "srb_join_probe_has_output"
    _need_input = "srb_join_probe_need_input"
    _is_finished = "srb_join_probe_is_finished"
    _pull = "srb_join_probe_pull_chunk"
    _pull_all = "srb_join_probe_pull_all"
    _pull_device = "srb_join_probe_pull_device"


class HashJoinOperator(_Operator):
    """Build side is this object (HashJoinBuildOperator); `.probe` is the HashJoinProbeOperator sharing its table."""

    def __init__(self, ctx: Context, plan):
        h = C.c_void_p()
        _check(ctx.L.srb_join_create(ctx.h, C.byref(plan.c), C.byref(h)))
        super().__init__(ctx, h)
        self._plan = plan
        self.probe = _ProbeSide(ctx, h)

    def build_runtime_filter(self, key_index: int = 0, expected_rows: int = 0, mix: bool = False, bitset: bool = False) -> "RuntimeFilter":
        """after set_finishing: the partial runtime filter over build key `key_index` (HashJoinBuildOperator -> RuntimeBloomFilter).
        mix=False hashes as the reference tree builds (identity); mix=True applies phmap_mix<8> (ENABLE_PHMAP_MIX); bitset=True: this
        builder holds the whole build side (broadcast join), so a RuntimeBitsetFilter may be chosen (SRB_RF_MAYBE_BITSET)"""
        h = C.c_void_p()
        _check(self.L.srb_join_build_runtime_filter(self.h, key_index, expected_rows, (1 if mix else 0) | (2 if bitset else 0), C.byref(h)))
        return RuntimeFilter(self.ctx, h)

    def close(self):
        if self.h:
            self.L.srb_op_close(self.h)
            self.h = None
            self.probe.h = None


class TopNOperator(_Operator):
    """ORDER BY key_slots (is_asc / nulls_first per key) OFFSET offset LIMIT limit on the device (ChunksSorterTopn): push chunks,
    set_finishing, pull_all -> every column of the input, rows [offset, offset + limit) of the sorted order"""

    def __init__(self, ctx: Context, key_slots, is_asc, nulls_first, limit: int, offset: int = 0):
        n = len(key_slots)
        self._keep = ((C.c_int32 * n)(*key_slots), (C.c_uint8 * n)(*[int(bool(x)) for x in is_asc]), (C.c_uint8 * n)(*[int(bool(x)) for x in nulls_first]))
        plan = K.srb_topn_plan(n, 0, self._keep[0], self._keep[1], self._keep[2], offset, limit)
        h = C.c_void_p()
        _check(ctx.L.srb_topn_create(ctx.h, C.byref(plan), C.byref(h)))
        super().__init__(ctx, h)


def run_scan_agg(ctx: Context, plan, chunk: K.Chunk, piece_rows: int = 0) -> K.Chunk:
    """Convenience driver loop: push (optionally in pieces, as the pipeline driver would), finish, pull everything."""
    op = ScanAggOperator(ctx, plan)
    try:
        if piece_rows and chunk.n_rows > piece_rows and all(c.mem_kind == K.MEM_HOST for c in chunk.columns.values()):
            for a in range(0, chunk.n_rows, piece_rows):
                op.push_chunk(slice_chunk(chunk, a, min(chunk.n_rows, a + piece_rows)))
        else:
            op.push_chunk(chunk)
        op.set_finishing()
        return op.pull_all()
    finally:
        op.close()


def slice_chunk(chunk: K.Chunk, a: int, b: int) -> K.Chunk:
    """Rows [a,b) of a host chunk (numpy views; binary columns keep their shared byte buffer)."""
    out = K.Chunk(b - a)
    for slot, c in chunk.columns.items():
        nulls = None if c.nulls is None else c.nulls[a:b]
        if c.offsets is not None:
            out.columns[slot] = K.Column(c.type, c.data, nulls, c.offsets[a:b + 1], c.precision, c.scale, c.mem_kind, c.n_bytes)
        else:
            out.columns[slot] = K.Column(c.type, c.data[a:b], nulls, None, c.precision, c.scale, c.mem_kind)
    return out


class Partitioner:
    """Hash partitioning of a device chunk by the reference's shuffle hash (ExchangeSinkOperator HASH_PARTITIONED)."""

    def __init__(self, ctx: Context, n_partitions: int):
        self.ctx, self.L, self.n = ctx, ctx.L, n_partitions
        h = C.c_void_p()
        _check(self.L.srb_partition_create(ctx.h, n_partitions, 0, C.byref(h)))
        self.h = h

    def run(self, chunk: K.Chunk, key_slots):
        """-> (device Chunk grouped by destination, [count per destination])"""
        cch, keep = chunk.to_c()
        ks = (C.c_int32 * len(key_slots))(*key_slots)
        out = K.srb_chunk()
        counts = (C.c_int64 * self.n)()
        _check(self.L.srb_partition_run(self.h, C.byref(cch), len(key_slots), ks, C.byref(out), counts))
        return _raw_chunk(out), list(counts)

    def close(self):
        if self.h:
            self.L.srb_partition_destroy(self.h)
            self.h = None


class Communicator:
    """NCCL communicator of libsrb (all-to-all shuffle, all-gather/broadcast). `store_set` / `store_get` move the 128-byte
    ncclUniqueId between ranks (torch.distributed's store in the harness)."""

    _nccl_preloaded = False

    @classmethod
    def _preload_nccl(cls):
        """libsrb binds NCCL with dlopen("libnccl.so.2"). When PyTorch is installed its wheel carries a newer NCCL under the
        same soname; whichever copy is mapped first wins for the whole process, and libtorch_cuda needs symbols of its own
        copy. So map the wheel's copy first (no-op when torch is already imported or no wheel copy exists)."""
        if cls._nccl_preloaded:
            return
        cls._nccl_preloaded = True
        try:
            import importlib.util
            spec = importlib.util.find_spec("nvidia.nccl")
            for d in (spec.submodule_search_locations if spec else []):
                p = os.path.join(d, "lib", "libnccl.so.2")
                if os.path.exists(p):
                    C.CDLL(p, mode=C.RTLD_GLOBAL)
                    return
        except Exception:
            pass

    def __init__(self, ctx: Context, store_set, store_get):
        self._preload_nccl()
        self.ctx, self.L = ctx, ctx.L
        buf = (C.c_char * 128)()
        if ctx.rank == 0:
            _check(self.L.srb_comm_unique_id(buf))
            store_set(bytes(buf.raw))
        else:
            raw = store_get()
            C.memmove(buf, raw, 128)
        h = C.c_void_p()
        _check(self.L.srb_comm_create(ctx.h, buf, C.byref(h)))
        self.h = h

    def all_to_all(self, chunk: K.Chunk, send_counts):
        cch, keep = chunk.to_c()
        w = self.ctx.world_size
        sc = (C.c_int64 * w)(*send_counts)
        rc = (C.c_int64 * w)()
        out = K.srb_chunk()
        _check(self.L.srb_comm_all_to_all(self.h, C.byref(cch), sc, C.byref(out), rc))
        return _raw_chunk(out), list(rc)

    def shuffle(self, chunk: K.Chunk, key_slots):
        """Hash partitioning fused with the exchange: rows are scattered straight into the destination ranks' receive buffers over
        peer memory (srb_comm_shuffle). -> (device Chunk of the rows this rank received, [rows per source rank])"""
        cch, keep = chunk.to_c()
        ks = (C.c_int32 * len(key_slots))(*key_slots)
        rc = (C.c_int64 * self.ctx.world_size)()
        out = K.srb_chunk()
        _check(self.L.srb_comm_shuffle(self.h, C.byref(cch), len(key_slots), ks, C.byref(out), rc))
        return _raw_chunk(out), list(rc)

    def abort(self):
        """ncclCommAbort: in-flight and later collectives on this rank return; every later call fails with CANCELLED"""
        _check(self.L.srb_comm_abort(self.h))

    def check(self):
        """raises when the communicator was aborted or NCCL reports an asynchronous error"""
        _check(self.L.srb_comm_check(self.h))

    def launches(self) -> int:
        n = C.c_int64()
        _check(self.L.srb_comm_counters(self.h, C.byref(n)))
        return n.value

    def merge_runtime_filter(self, rf: "RuntimeFilter") -> "RuntimeFilter":
        """global runtime filter: the OR of every rank's partial filter, in place on every rank"""
        _check(self.L.srb_comm_merge_runtime_filter(self.h, rf.h))
        return rf

    def all_gather(self, chunk: K.Chunk) -> K.Chunk:
        """Every rank's rows on every rank, rank 0 first. Device chunks (fixed-width columns) stay on the device; host chunks
        (any column kind: the few intermediate rows of a low-cardinality plan) come back as host chunks."""
        cch, keep = chunk.to_c()
        out = K.srb_chunk()
        _check(self.L.srb_comm_all_gather(self.h, C.byref(cch), C.byref(out)))
        on_host = chunk.columns and all(c.mem_kind == K.MEM_HOST for c in chunk.columns.values())
        return K.decode_chunk(out) if on_host else _raw_chunk(out)

    def close(self):
        if self.h:
            self.L.srb_comm_destroy(self.h)
            self.h = None
