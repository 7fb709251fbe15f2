"""Segment data pages for the page-fed scan (srb_op_push_pages): the page-set description the C ABI takes, and a frame-of-reference page
BUILDER for synthetic tables.

The builder restates FrameOfReferencePageBuilder / ForEncoder (be/src/storage/rowset/frame_of_reference_page.h:46-104,
be/src/util/frame_of_reference_coding.cpp:44-241): frames of 128 values, [min LE][bit-packed deltas, MSB first], footer of
(storage_format, bit_width) pairs + u8 128 + u32 count; format 0 = value - min, 1 = value - previous (non-decreasing frame), 2 = the value itself
(the subtraction would overflow). It is vectorised over all frames of a column with torch integer ops, so it runs on the CPU (tests) and on the
GPU (bench.py fills its host-side "segment" from resident synthetic columns in a fraction of a second). Building pages is what the BE's segment
writer does, not part of the measured path; decoding is (csrc/srb_pages.cuh). Types up to 64 bits; the oracle's builder covers int128.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Sequence, Tuple

from . import capi as K

ENC_PLAIN, ENC_FOR = 2, 7  # EncodingTypePB (gensrc/proto/segment.proto:49-58)
FRAME = 128


class srb_page(C.Structure):
    _fields_ = [("offset", C.c_int64), ("bytes", C.c_int64), ("n_rows", C.c_int64)]


class srb_page_column(C.Structure):
    _fields_ = [("type", C.c_int32), ("precision", C.c_int32), ("scale", C.c_int32), ("encoding", C.c_int32), ("data", C.c_void_p),
                ("bytes", C.c_int64), ("mem_kind", C.c_int32), ("n_pages", C.c_int32), ("pages", C.POINTER(srb_page))]


class PageColumn:
    """the encoded pages of one column over a row range: `data` = int address or numpy uint8 array; pages = [(offset, bytes, n_rows)]"""

    def __init__(self, type_: int, data, nbytes: int, pages: Sequence[Tuple[int, int, int]], encoding: int = ENC_FOR, precision: int = 0,
                 scale: int = 0, mem_kind: int = K.MEM_HOST, keep=None):
        self.type, self.precision, self.scale, self.encoding, self.mem_kind = type_, precision, scale, encoding, mem_kind
        self.data, self.nbytes, self.pages, self.keep = data, nbytes, list(pages), keep

    def n_rows(self) -> int:
        return sum(p[2] for p in self.pages)

    def address(self) -> int:
        return self.data if isinstance(self.data, int) else self.data.ctypes.data


def to_c(columns: Dict[int, PageColumn]):
    """-> (n_cols, slot id array, srb_page_column array, keepalive)"""
    n = len(columns)
    slots = (C.c_int32 * n)(*columns.keys())
    cols = (srb_page_column * n)()
    keep = [slots, cols, columns]
    for i, c in enumerate(columns.values()):
        pg = (srb_page * max(1, len(c.pages)))()
        for j, (off, nb, nr) in enumerate(c.pages):
            pg[j].offset, pg[j].bytes, pg[j].n_rows = off, nb, nr
        keep.append(pg)
        cols[i].type, cols[i].precision, cols[i].scale, cols[i].encoding = c.type, c.precision, c.scale, c.encoding
        cols[i].data, cols[i].bytes, cols[i].mem_kind, cols[i].n_pages, cols[i].pages = c.address(), c.nbytes, c.mem_kind, len(c.pages), pg
    return n, slots, cols, keep


def prepare(columns: Dict[int, PageColumn]):
    """-> the argument tuple of ops.ScanAggOperator.push_pages_prepared"""
    n, slots, cols, keep = to_c(columns)
    rows = next(iter(columns.values())).n_rows() if columns else 0
    return n, slots, cols, keep, rows


def slice_pages(col: PageColumn, first: int, last: int) -> PageColumn:
    """pages [first, last) of a column as a page set of their own (same byte buffer)"""
    return PageColumn(col.type, col.data, col.nbytes, col.pages[first:last], col.encoding, col.precision, col.scale, col.mem_kind, col.keep)


class PageTable:
    """a table's columns as encoded pages, cut into page sets (row ranges) that are pushed one after the other: what a scan operator is fed
    instead of a chunk (pipelines accept either)"""

    def __init__(self, columns: Dict[int, PageColumn], pages_per_set: int = 512):
        self.columns = columns
        n_pages = len(next(iter(columns.values())).pages)
        assert all(len(c.pages) == n_pages for c in columns.values()), "the columns' pages must cover the same row ranges"
        self.sets = [prepare({slot: slice_pages(c, a, min(a + pages_per_set, n_pages)) for slot, c in columns.items()}) for a in range(0, n_pages, pages_per_set)]
        self.n_rows = next(iter(columns.values())).n_rows()
        self.nbytes = sum(c.nbytes for c in columns.values())


def push_source(op, src):
    """push a chunk, or every page set of a PageTable, into a scan operator"""
    if isinstance(src, PageTable):
        for ps in src.sets:
            op.push_pages_prepared(ps)
    else:
        op.push_chunk(src)


# ---- the builder ----------------------------------------------------------------------------------------------------------------
def _bit_length(x):
    """significant bits of non-negative int64 values (bits() of frame_of_reference_coding.h:50-71)"""
    import torch
    out = torch.zeros_like(x)
    v = x.clone()
    for s in (32, 16, 8, 4, 2, 1):
        big = v >= (1 << s)
        out += big.to(torch.int64) * s
        v = torch.where(big, v >> s, v)
    return out + (v > 0).to(torch.int64)


def for_encode_column(values, page_rows: int = 1 << 16):
    """values: 1-D torch tensor of int8 / int16 / int32 / int64 (any device) -> (uint8 tensor with every page back to back, [(offset, bytes,
    n_rows)]). page_rows (a multiple of 128) values go into a page, as a PageBuilder fills pages up to data_page_size."""
    import torch
    assert page_rows % FRAME == 0 and values.dim() == 1
    width = values.element_size()
    mb = 8 if width == 8 else 4
    n = values.numel()
    dev = values.device
    if n == 0:
        return torch.tensor([FRAME, 0, 0, 0, 0], dtype=torch.uint8, device=dev), [(0, 5, 0)]
    n_frames = (n + FRAME - 1) // FRAME
    tail = n - (n_frames - 1) * FRAME  # values in the last frame
    v = values.to(torch.int64)
    if tail < FRAME:  # pad the last frame with its last value: min, max, order and deltas are unchanged, and only `tail` values are packed
        v = torch.cat([v, v[-1:].expand(FRAME - tail)])
    V = v.view(n_frames, FRAME)
    tmax = (1 << (8 * width - 1)) - 1
    half = tmax >> 1
    mask = -1 if width == 8 else (1 << (8 * width)) - 1
    asc = (V[:, 1:] >= V[:, :-1]).all(dim=1)
    mn, mx = V.min(dim=1).values, V.max(dim=1).values
    over_asc = (((V[:, 1:] >> 1) - (V[:, :-1] >> 1)) > half).any(dim=1)
    over_rng = ((mx >> 1) - (mn >> 1)) > half
    keep = torch.where(asc, over_asc, over_rng)
    fmt = torch.where(keep, torch.full_like(mn, 2), asc.to(torch.int64))
    # the numbers that are packed (as unsigned `width`-byte integers)
    d_asc = torch.cat([torch.zeros_like(V[:, :1]), V[:, 1:] - V[:, :-1]], dim=1)
    D = torch.where(keep[:, None], V, torch.where(asc[:, None], d_asc, V - mn[:, None])) & mask
    del d_asc
    top = torch.where(asc, D.max(dim=1).values, (mx - mn) & mask)
    bw = torch.where(keep, torch.full_like(mn, 8 * width), _bit_length(torch.where(keep, torch.zeros_like(top), top)))
    cnt = torch.full_like(bw, FRAME)
    cnt[-1] = tail
    D[-1, tail:] = 0  # bit_pack writes `tail` values: the rest of the last byte stays zero
    packed_len = (cnt * bw + 7) // 8
    # the reference reserves cnt * bit_width BYTES for an overflow frame (frame_of_reference_coding.cpp:176-180); the rest stays zero
    size = mb + torch.where(keep, cnt * bw, packed_len)
    # page layout: frames, then (format, bit width) pairs, u8 128, u32 count
    fpp = page_rows // FRAME
    n_pages = (n_frames + fpp - 1) // fpp
    page_of = torch.arange(n_frames, device=dev) // fpp
    frames_in = torch.bincount(page_of, minlength=n_pages)
    body = torch.zeros(n_pages, dtype=torch.int64, device=dev).index_add_(0, page_of, size)
    page_bytes = body + 2 * frames_in + 5
    page_start = torch.cumsum(page_bytes, 0) - page_bytes
    csum = torch.cumsum(size, 0) - size
    first_frame = torch.arange(n_pages, device=dev) * fpp
    frame_off = page_start[page_of] + csum - csum[first_frame][page_of]
    total = int(page_bytes.sum().item())
    out = torch.zeros(total + 16, dtype=torch.uint8, device=dev)
    # min values, little endian
    mnb = mn  # (narrow types store the sign-extended value in 4 bytes: put_fixed32_le)
    for b in range(mb):
        out[frame_off + b] = ((mnb >> (8 * b)) & 0xFF).to(torch.uint8)
    # packed bits, one group of frames per bit width
    for w in torch.unique(bw).tolist():
        if w == 0:
            continue
        sel = torch.nonzero(bw == w).flatten()
        for a in range(0, sel.numel(), 1 << 13):  # bounded temporaries: 8192 frames x 128 x w bits
            idx = sel[a:a + (1 << 13)]
            shifts = torch.arange(w - 1, -1, -1, device=dev)
            bits = ((D[idx][:, :, None] >> shifts) & 1).to(torch.uint8).view(idx.numel(), FRAME * w)
            by = (bits.view(idx.numel(), FRAME * w // 8, 8).to(torch.int32) * torch.tensor([128, 64, 32, 16, 8, 4, 2, 1], dtype=torch.int32, device=dev)).sum(dim=2).to(torch.uint8)
            ln = packed_len[idx]
            pos = torch.arange(FRAME * w // 8, device=dev)[None, :].expand(idx.numel(), -1)
            ok = pos < ln[:, None]
            dst = (frame_off[idx] + mb)[:, None] + pos
            out[dst[ok]] = by[ok]
    # footers
    meta0 = page_start + body
    rank = torch.arange(n_frames, device=dev) - first_frame[page_of]
    out[meta0[page_of] + 2 * rank] = fmt.to(torch.uint8)
    out[meta0[page_of] + 2 * rank + 1] = bw.to(torch.uint8)
    tailpos = meta0 + 2 * frames_in
    out[tailpos] = FRAME
    rows_in = torch.full((n_pages,), page_rows, dtype=torch.int64, device=dev)
    rows_in[-1] = n - (n_pages - 1) * page_rows
    for b in range(4):
        out[tailpos + 1 + b] = ((rows_in >> (8 * b)) & 0xFF).to(torch.uint8)
    pages = list(zip(page_start.tolist(), page_bytes.tolist(), rows_in.tolist()))
    return out[:total], pages


def plain_encode_column(values, page_rows: int = 1 << 16):
    """PlainPageBuilder (be/src/storage/rowset/plain_page.h:50-130): [u32 count][values] per page"""
    import torch
    width = values.element_size()
    n = values.numel()
    raw = values.contiguous().view(torch.uint8)
    parts, pages, off = [], [], 0
    for a in range(0, max(n, 1), page_rows):
        m = max(0, min(page_rows, n - a))
        hdr = torch.tensor(list(int(m).to_bytes(4, "little")), dtype=torch.uint8, device=values.device)
        parts += [hdr, raw[a * width:(a + m) * width]]
        pages.append((off, 4 + m * width, m))
        off += 4 + m * width
    return torch.cat(parts), pages
