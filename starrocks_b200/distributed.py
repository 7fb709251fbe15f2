"""Host-side logic of the multi-GPU path: which rows a rank owns, and the two-phase aggregation around the exchange.

The reference distributes an aggregation as phase 1 (update + serialize, per fragment instance) -> shuffle ->
phase 2 (merge + finalize) (be/src/exec/pipeline/aggregate/aggregate_streaming_sink_operator.cpp:105,
be/src/exec/aggregator.cpp:858-880; ExchangeSinkOperator HASH_PARTITIONED, exchange_sink_operator.cpp:569-633).
One process per GPU plays the role of one fragment instance:

  * low cardinality (Q1, Q6): every rank aggregates its row range; the <= a few intermediate rows per rank are
    all-gathered and merged everywhere (no hash partitioning: the table is tiny);
  * high cardinality: every rank pre-aggregates its rows, hash-partitions the intermediate (key, state) tuples by the
    reference's shuffle hash, exchanges them all-to-all and merges the tuples it receives; the result stays partitioned.

The functions take the local engine and the exchange as arguments, so the same code runs on the GPU (ops.ScanAggOperator +
ops.Partitioner + ops.Communicator over NCCL) and in the CPU tests (oracle + torch.distributed/gloo).
"""
from __future__ import annotations

from typing import Callable, List, Sequence, Tuple

import numpy as np

from . import capi as K


def shard_range(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    """Rows [begin, end) of rank `rank`: contiguous ranges, the remainder spread over the first ranks."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    base, rem = divmod(n_rows, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def concat_chunks(chunks: Sequence[K.Chunk]) -> K.Chunk:
    """Row-wise concatenation of host chunks with identical schemas (a rank's received pieces, in source-rank order)."""
    chunks = [c for c in chunks if c is not None]
    if not chunks:
        raise ValueError("nothing to concatenate")
    first = chunks[0]
    out = K.Chunk(sum(c.n_rows for c in chunks))
    for slot, c0 in first.columns.items():
        cols = [c.columns[slot] for c in chunks]
        any_nulls = any(c.nulls is not None for c in cols)
        nulls = None
        if any_nulls:
            nulls = np.concatenate([c.nulls if c.nulls is not None else np.zeros(ch.n_rows, np.uint8) for c, ch in zip(cols, chunks)])
        if K.is_binary(c0.type):
            datas, offs, base = [], [np.zeros(1, np.uint32)], 0
            for c in cols:
                o = np.asarray(c.offsets, dtype=np.int64)
                datas.append(np.asarray(c.data[int(o[0]):int(o[-1])], dtype=np.uint8))
                offs.append((o[1:] - o[0] + base).astype(np.uint32))
                base += int(o[-1] - o[0])
            data = np.concatenate(datas + [np.zeros(16, np.uint8)])
            out.columns[slot] = K.Column(c0.type, data, nulls, np.concatenate(offs), c0.precision, c0.scale, n_bytes=base)
        else:
            out.columns[slot] = K.Column(c0.type, np.concatenate([c.data for c in cols]), nulls, None, c0.precision, c0.scale)
    return out


def take_rows(chunk: K.Chunk, idx: np.ndarray) -> K.Chunk:
    """Rows `idx` of a host chunk, in that order (fixed-width and binary columns)."""
    out = K.Chunk(int(len(idx)))
    for slot, c in chunk.columns.items():
        nulls = None if c.nulls is None else c.nulls[idx]
        if K.is_binary(c.type):
            o = np.asarray(c.offsets, dtype=np.int64)
            vals = [bytes(c.data[int(o[i]):int(o[i + 1])]) for i in idx]
            col = K.Column.from_strings(vals, nulls, c.type)
            out.columns[slot] = col
        else:
            out.columns[slot] = K.Column(c.type, np.ascontiguousarray(c.data[idx]), nulls, None, c.precision, c.scale)
    return out


def group_by_destination(chunk: K.Chunk, dest: np.ndarray, world: int) -> Tuple[K.Chunk, List[int]]:
    """Stable counting sort of the rows by destination (what Shuffler + append_selective do per chunk,
    exchange/shuffler.h:70-85): returns the reordered chunk and the row count per destination."""
    dest = np.asarray(dest, dtype=np.int64)
    if len(dest) != chunk.n_rows:
        raise ValueError("one destination per row")
    if len(dest) and (dest.min() < 0 or dest.max() >= world):
        raise ValueError("destination out of range")
    order = np.argsort(dest, kind="stable")
    counts = np.bincount(dest, minlength=world).astype(np.int64).tolist()
    return take_rows(chunk, order), counts


class ObjectExchange:
    """Exchange over torch.distributed object collectives (any backend; host chunks). The GPU path uses
    ops.Communicator (NCCL, device buffers) with the same two methods."""

    def __init__(self, dist, rank: int, world: int, group=None):
        self.dist, self.rank, self.world, self.group = dist, rank, world, group

    SMALL = 16 << 10  # bytes per rank of the one-collective path (a handful of intermediate rows)

    def all_gather(self, chunk: K.Chunk) -> K.Chunk:
        # Small payloads (the <= 4 intermediate rows of a low-cardinality plan) travel in ONE fixed-size collective:
        # [u64 length][pickle] padded to SMALL. all_gather_object costs two collectives plus tensor round trips.
        import pickle
        import torch
        blob = pickle.dumps(chunk, protocol=pickle.HIGHEST_PROTOCOL)
        fits = len(blob) + 8 <= self.SMALL
        mine = torch.zeros(self.SMALL, dtype=torch.uint8)
        if fits:
            mine[:8] = torch.frombuffer(bytearray(len(blob).to_bytes(8, "little")), dtype=torch.uint8)
            mine[8:8 + len(blob)] = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
        dev = None
        if self.dist.get_backend(self.group) == "nccl":
            dev = torch.device("cuda", torch.cuda.current_device())
            mine = mine.to(dev)
        allb = torch.empty(self.SMALL * self.world, dtype=torch.uint8, device=dev)
        self.dist.all_gather_into_tensor(allb, mine, group=self.group)
        allb = allb.cpu().numpy().reshape(self.world, self.SMALL)
        lens = [int.from_bytes(allb[r, :8].tobytes(), "little") for r in range(self.world)]
        if all(lens):  # every rank's payload fitted (length 0 marks "too large")
            return concat_chunks([pickle.loads(allb[r, 8:8 + lens[r]].tobytes()) for r in range(self.world)])
        got = [None] * self.world
        self.dist.all_gather_object(got, chunk, group=self.group)
        return concat_chunks(got)

    def all_to_all(self, grouped: K.Chunk, send_counts: Sequence[int]) -> Tuple[K.Chunk, List[int]]:
        if len(send_counts) != self.world or sum(send_counts) != grouped.n_rows:
            raise ValueError("send counts do not cover the chunk")
        from .ops import slice_chunk
        pieces, pos = [], 0
        for c in send_counts:
            pieces.append(slice_chunk(grouped, pos, pos + c))
            pos += c
        # binary columns of a slice still carry the whole byte buffer: trim before shipping
        pieces = [concat_chunks([p]) for p in pieces]
        got = [None] * self.world
        self.dist.all_gather_object(got, pieces, group=self.group)  # got[src][dst]
        mine = [got[src][self.rank] for src in range(self.world)]
        return concat_chunks(mine), [m.n_rows for m in mine]


def two_phase_gather(local_update: Callable[[K.Chunk], K.Chunk], merge: Callable[[K.Chunk], K.Chunk], exchange, shard: K.Chunk) -> K.Chunk:
    """Low-cardinality plan: phase 1 on the rank's rows, all-gather the intermediate rows, phase 2 on every rank."""
    partial = local_update(shard)
    return merge(exchange.all_gather(partial))


def two_phase_shuffle(local_update: Callable[[K.Chunk], K.Chunk], partition: Callable[[K.Chunk], Tuple[K.Chunk, List[int]]],
                      merge: Callable[[K.Chunk], K.Chunk], exchange, shard: K.Chunk) -> K.Chunk:
    """High-cardinality plan: phase 1 locally, hash-partition the intermediate tuples, all-to-all, phase 2 on the received
    tuples. Every group ends on exactly one rank (the one its key hashes to)."""
    partial = local_update(shard)
    grouped, counts = partition(partial)
    received, _ = exchange.all_to_all(grouped, counts)
    return merge(received)
