// srb_pages.cuh — segment data pages of fixed-width columns decoded on the device, in front of the scan.
//
// Replaces, for the encodings below, what the scan does on the CPU before a predicate ever runs:
//   FrameOfReferencePageDecoder::next_batch -> ForDecoder<T>::get_batch / decode_current_frame / bit_unpack
//       be/src/storage/rowset/frame_of_reference_page.h:107-223, be/src/util/frame_of_reference_coding.cpp:243-410
//   PlainPageDecoder::next_batch                                be/src/storage/rowset/plain_page.h:133-276
//   (reached from SegmentIterator::_read -> ScalarColumnIterator::next_batch, be/src/storage/rowset/segment_iterator.cpp:1411-1760)
// FOR page (frame_of_reference_coding.h:78-101): frames of 128 values = [min: 4 / 8 / 16 bytes LE][128 x bit_width bits, MSB first];
// footer = (u8 storage_format, u8 bit_width) per frame, u8 128, u32 value count. storage_format 0: value - min; 1: value - previous value
// (non-decreasing frame, first delta 0); 2: the value itself in 8 * sizeof(T) bits. PLAIN page: [u32 count][values].
//
// The encoded bytes cross PCIe (5-6x fewer than the decoded columns of TPC-H Q1) and are expanded here at HBM speed:
//   k_page_dir     one CTA per page: reads the footer, prefix-sums the frame sizes -> one FrameDesc per frame
//   k_for_decode   one warp per frame: lane l decodes values l, l+32, l+64, l+96, so every store instruction writes 32 consecutive values;
//                  the delta format is a warp scan per 32 values with a running base
//   k_plain_decode one thread per value
#pragma once
#include "srb_dev.cuh"

namespace srb {

struct PageDesc {
    uint64_t off;    // first byte of the page inside the column's encoded buffer
    uint64_t row0;   // first output row of the page
    uint32_t bytes;  // page length
    uint32_t n_rows; // values the caller says the page holds (checked against the footer)
    uint32_t frame0; // index of the page's first frame in the frame directory
    uint32_t _pad;
};
struct FrameDesc {
    uint64_t off;  // first byte of the frame (its min value)
    uint64_t row0; // first output row
    uint32_t bw, fmt, cnt, _pad;
};
constexpr int PAGE_ERR_FOOTER = 1, PAGE_ERR_BOUNDS = 2;

SRB_D int for_min_bytes(int w) { return w == 16 ? 16 : (w == 8 ? 8 : 4); }

__global__ void __launch_bounds__(256) k_page_dir(const uint8_t* __restrict__ enc, const PageDesc* __restrict__ pages, const int width,
                                                  FrameDesc* __restrict__ dir, unsigned int* __restrict__ err) {
    __shared__ unsigned int wsum[8];
    __shared__ unsigned long long base_s;
    const PageDesc pg = pages[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (pg.bytes < 5) {
        if (tid == 0 && pg.n_rows) atomicOr(err, (unsigned)PAGE_ERR_FOOTER);
        return;
    }
    const uint8_t* p = enc + pg.off;
    const uint32_t fv = p[pg.bytes - 5];
    const uint32_t n = (uint32_t)p[pg.bytes - 4] | ((uint32_t)p[pg.bytes - 3] << 8) | ((uint32_t)p[pg.bytes - 2] << 16) | ((uint32_t)p[pg.bytes - 1] << 24);
    const uint32_t frames = (n + 127) / 128;
    if (n != pg.n_rows || (n && fv != 128) || (uint64_t)frames * 2 + 5 > pg.bytes) {
        if (tid == 0) atomicOr(err, (unsigned)PAGE_ERR_FOOTER);
        return;
    }
    const uint8_t* meta = p + pg.bytes - 5 - (size_t)frames * 2;
    const uint32_t body_end = pg.bytes - 5 - frames * 2;
    const int mb = for_min_bytes(width);
    if (tid == 0) base_s = 0;
    __syncthreads();
    for (uint32_t f0 = 0; f0 < frames; f0 += 256) {
        const uint32_t f = f0 + tid;
        uint32_t bw = 0, fmt = 0, size = 0;
        if (f < frames) {
            fmt = meta[2 * f];
            bw = meta[2 * f + 1];
            size = bw * 16 + mb; // ForDecoder::init: bit_width * 128 / 8 + sizeof(min), also for a short last frame
        }
        uint32_t incl = size;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) wsum[wid] = incl;
        __syncthreads();
        uint32_t wb = 0, tot = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            wb += w < wid ? wsum[w] : 0u;
            tot += wsum[w];
        }
        const unsigned long long base = base_s;
        if (f < frames) {
            const unsigned long long at = base + wb + incl - size;
            const uint32_t cnt = f == frames - 1 ? n - f * 128 : 128;
            const uint32_t used = mb + (cnt * bw + 7) / 8;
            if (bw > (uint32_t)width * 8 || fmt > 2 || at + used > body_end) atomicOr(err, (unsigned)PAGE_ERR_BOUNDS);
            dir[pg.frame0 + f] = FrameDesc{pg.off + at, pg.row0 + (uint64_t)f * 128, bw <= (uint32_t)width * 8 ? bw : 0u, fmt, at + used > body_end ? 0u : cnt, 0u};
        }
        __syncthreads();
        if (tid == 0) base_s = base + tot;
        __syncthreads();
    }
}

// bits [bitpos, bitpos + bw) (bw <= 64) of the big-endian bit stream at p
SRB_D uint64_t for_extract(const uint8_t* __restrict__ p, uint32_t bitpos, uint32_t bw) {
    if (bw == 0) return 0;
    const uint8_t* q = p + (bitpos >> 3);
    const uint32_t sh = bitpos & 7;
    uint64_t x = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) x = (x << 8) | q[i];
    uint64_t v = (x << sh) >> (64 - bw);
    if (sh + bw > 64) v |= (uint64_t)q[8] >> (8 - (sh + bw - 64));
    return v;
}

template <int W>
struct PageWord {
    typedef uint64_t type;
};
template <>
struct PageWord<16> {
    typedef u128 type;
};

SRB_D uint64_t pw_hi(uint64_t) { return 0; }
SRB_D uint64_t pw_hi(u128 v) { return (uint64_t)(v >> 64); }
template <class U>
SRB_D U pw_make(uint64_t hi, uint64_t lo);
template <>
SRB_D uint64_t pw_make<uint64_t>(uint64_t, uint64_t lo) { return lo; }
template <>
SRB_D u128 pw_make<u128>(uint64_t hi, uint64_t lo) { return ((u128)hi << 64) | lo; }

template <int W> // bytes per value: 1, 2, 4, 8, 16
__global__ void __launch_bounds__(256) k_for_decode(const uint8_t* __restrict__ enc, const FrameDesc* __restrict__ dir, const int64_t n_frames,
                                                     uint8_t* __restrict__ out) {
    typedef typename PageWord<W>::type U;
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t f = warp0; f < n_frames; f += n_warps) {
        const FrameDesc fd = dir[f];
        const uint8_t* p = enc + fd.off;
        U mn = 0;
        if (W == 16) {
            uint64_t lo = 0, hi = 0;
#pragma unroll
            for (int i = 7; i >= 0; i--) {
                lo = (lo << 8) | p[i];
                hi = (hi << 8) | p[8 + i];
            }
            mn = pw_make<U>(hi, lo);
        } else if (W == 8) {
            uint64_t lo = 0;
#pragma unroll
            for (int i = 7; i >= 0; i--) lo = (lo << 8) | p[i];
            mn = lo;
        } else {
            mn = (uint64_t)p[0] | ((uint64_t)p[1] << 8) | ((uint64_t)p[2] << 16) | ((uint64_t)p[3] << 24);
        }
        const uint8_t* body = p + (W == 16 ? 16 : (W == 8 ? 8 : 4));
        U run = mn; // delta format: the value before the current group of 32
#pragma unroll
        for (int t = 0; t < 4; t++) {
            const uint32_t j = (uint32_t)t * 32 + lane;
            const bool on = j < fd.cnt;
            U d = 0;
            if (on) {
                if (W == 16 && fd.bw > 64) {
                    const uint64_t hi = for_extract(body, j * fd.bw, fd.bw - 64);
                    const uint64_t lo = for_extract(body, j * fd.bw + (fd.bw - 64), 64);
                    d = pw_make<U>(hi, lo);
                } else {
                    d = for_extract(body, j * fd.bw, fd.bw);
                }
            }
            U v;
            if (fd.fmt == 1) { // value = previous value + delta: inclusive warp scan of the deltas on top of the running base
                U s = d;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    U y;
                    if (W == 16) {
                        const uint64_t ylo = __shfl_up_sync(0xffffffffu, (uint64_t)s, o), yhi = __shfl_up_sync(0xffffffffu, pw_hi(s), o);
                        y = pw_make<U>(yhi, ylo);
                    } else {
                        y = __shfl_up_sync(0xffffffffu, (uint64_t)s, o);
                    }
                    if (lane >= o) s += y;
                }
                v = run + s;
                U last;
                if (W == 16) {
                    const uint64_t llo = __shfl_sync(0xffffffffu, (uint64_t)v, 31), lhi = __shfl_sync(0xffffffffu, pw_hi(v), 31);
                    last = pw_make<U>(lhi, llo);
                } else {
                    last = __shfl_sync(0xffffffffu, (uint64_t)v, 31);
                }
                run = last;
            } else if (fd.fmt == 2) {
                v = d;
            } else {
                v = mn + d;
            }
            if (on) {
                const uint64_t r = fd.row0 + j;
                switch (W) {
                case 1: out[r] = (uint8_t)v; break;
                case 2: ((uint16_t*)out)[r] = (uint16_t)v; break;
                case 4: ((uint32_t*)out)[r] = (uint32_t)v; break;
                case 8: ((uint64_t*)out)[r] = (uint64_t)v; break;
                default: ((ulonglong2*)out)[r] = make_ulonglong2((uint64_t)v, pw_hi(v));
                }
            }
        }
    }
}

// PLAIN pages: [u32 count][values]; the values sit at an odd multiple of 4 bytes, so they are assembled from bytes
__global__ void __launch_bounds__(256) k_plain_decode(const uint8_t* __restrict__ enc, const PageDesc* __restrict__ pages, const int width,
                                                       uint8_t* __restrict__ out, unsigned int* __restrict__ err) {
    const PageDesc pg = pages[blockIdx.y];
    const uint8_t* p = enc + pg.off;
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const uint32_t n = pg.bytes >= 4 ? ((uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24)) : 0xffffffffu;
        if (n != pg.n_rows || (uint64_t)pg.n_rows * width + 4 != pg.bytes) atomicOr(err, (unsigned)PAGE_ERR_FOOTER);
    }
    if ((uint64_t)pg.n_rows * width + 4 > pg.bytes) return;
    const uint64_t total = (uint64_t)pg.n_rows * width;
    for (uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; b < total; b += (uint64_t)gridDim.x * blockDim.x)
        out[pg.row0 * width + b] = p[4 + b];
}

} // namespace srb
