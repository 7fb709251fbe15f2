// srb_hcpart.cuh — L2-blocked high-cardinality aggregation: partition the batch by table region before aggregating.
//
// With ~10^8 groups the group table is gigabytes and every row's probe + state update is a random DRAM access (a 32-byte
// sector read and written back for 16 useful bytes). The reference meets the same wall on CPU caches and answers with
// two-level / partitioned hash maps (exec/aggregate/agg_hash_map.h:92-98, aggregator.cpp:1132 try_convert_to_two_level_map;
// SURVEY §8 f1). Here the batch's (key, value) tuples are first scattered into P = 256 partitions, partition p holding exactly the rows
// whose home slot lies in the p-th 1/256 of the table. The unchanged aggregation kernel then sweeps the permuted tuples in
// order, so at any moment all CTAs update one or two ~32 MB table regions that stay resident in the 126 MB L2: probes and
// atomics complete in L2 and each table sector is read from and written to HBM once per batch instead of once per row.
//
// Passes: k_hc_hist (8 B/row read) -> k_hc_scan (256 counters) -> k_hc_scatter (16 B/row read, 16 B/row written, staged
// through shared memory so that a tile's rows of one partition leave as one contiguous run) -> aggregation over the copy.
#pragma once
#include "srb_rowpipe.cuh"

namespace srb {

constexpr int HC_PBITS = 8, HC_PARTS = 1 << HC_PBITS;
constexpr int HC_THREADS = 1024, HC_ROWS = 8, HC_TILE = HC_THREADS * HC_ROWS;

SRB_D int hc_part_of(uint64_t key, int seed, uint64_t mask, int shift) { return (int)(((phmap_mix8(key, seed) >> 7) & mask) >> shift); }

__global__ void __launch_bounds__(1024) k_hc_hist(const uint64_t* __restrict__ keys, int64_t n, int seed, uint64_t mask, int shift,
                                                  unsigned long long* __restrict__ hist) {
    __shared__ unsigned int h[HC_PARTS];
    for (int i = threadIdx.x; i < HC_PARTS; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) atomicAdd(&h[hc_part_of(keys[i], seed, mask, shift)], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < HC_PARTS; i += blockDim.x)
        if (h[i]) atomicAdd(&hist[i], (unsigned long long)h[i]);
}

// one block of HC_PARTS threads: cursor[p] = exclusive prefix of hist; hist is cleared for the next batch
__global__ void k_hc_scan(unsigned long long* __restrict__ hist, unsigned long long* __restrict__ cursor) {
    __shared__ unsigned long long s[HC_PARTS];
    const int t = threadIdx.x;
    s[t] = hist[t];
    __syncthreads();
    for (int o = 1; o < HC_PARTS; o <<= 1) {
        unsigned long long x = t >= o ? s[t - o] : 0;
        __syncthreads();
        s[t] += x;
        __syncthreads();
    }
    cursor[t] = s[t] - hist[t];
    hist[t] = 0;
}

template <int NV> // value columns carried with the key: 1 (phase 1: k, v) or 2 (phase 2: k, partial sum, partial count)
struct HcSmem {
    uint64_t k[HC_TILE];
    uint64_t v[NV][HC_TILE];
    uint8_t p[HC_TILE];
    unsigned int cnt[HC_PARTS], scan[HC_PARTS], start[HC_PARTS];
    unsigned long long gbase[HC_PARTS];
};

struct HcVals {
    const uint64_t* src[2];
    uint64_t* dst[2];
};
template <int NV>
__global__ void __launch_bounds__(HC_THREADS, 1)
        k_hc_scatter(const uint64_t* __restrict__ keys, const HcVals vals, int64_t n, int seed, uint64_t mask, int shift,
                     unsigned long long* __restrict__ cursor, uint64_t* __restrict__ pk) {
    extern __shared__ __align__(16) uint8_t raw[];
    HcSmem<NV>& sm = *reinterpret_cast<HcSmem<NV>*>(raw);
    const int tid = threadIdx.x;
    const int64_t n_tiles = (n + HC_TILE - 1) / HC_TILE;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t base = tile * HC_TILE;
        const int rows = (int)(n - base < HC_TILE ? n - base : HC_TILE);
        if (tid < HC_PARTS) sm.cnt[tid] = 0;
        __syncthreads();
        uint64_t k[HC_ROWS];
        int part[HC_ROWS];
        unsigned int rank[HC_ROWS];
#pragma unroll
        for (int r = 0; r < HC_ROWS; r++) {
            const int j = r * HC_THREADS + tid;
            if (j < rows) {
                k[r] = keys[base + j];
            }
        }
#pragma unroll
        for (int r = 0; r < HC_ROWS; r++) {
            const int j = r * HC_THREADS + tid;
            if (j < rows) {
                part[r] = hc_part_of(k[r], seed, mask, shift);
                rank[r] = atomicAdd(&sm.cnt[part[r]], 1u);
            }
        }
        __syncthreads();
        // exclusive scan of the tile's 256 counts (first 8 warps), and one global reservation per partition
        if (tid < HC_PARTS) {
            const unsigned int c = sm.cnt[tid];
            unsigned int x = c;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned int y = __shfl_up_sync(0xffffffffu, x, o);
                if ((tid & 31) >= o) x += y;
            }
            sm.scan[tid] = x; // inclusive within the warp
            sm.gbase[tid] = c ? atomicAdd(&cursor[tid], (unsigned long long)c) : 0ULL;
        }
        __syncthreads();
        if (tid < HC_PARTS) {
            unsigned int add = 0;
            for (int w = 0; w < (tid >> 5); w++) add += sm.scan[w * 32 + 31]; // totals of the preceding warps
            sm.start[tid] = sm.scan[tid] + add - sm.cnt[tid];
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < HC_ROWS; r++) {
            const int j = r * HC_THREADS + tid;
            if (j < rows) {
                const unsigned int pos = sm.start[part[r]] + rank[r];
                sm.k[pos] = k[r];
                sm.p[pos] = (uint8_t)part[r];
            }
        }
        // the value columns follow one at a time (their eight loads per thread in flight together): the keys alone had to be held
        // across the barriers above
#pragma unroll
        for (int c = 0; c < NV; c++) {
            uint64_t v[HC_ROWS];
#pragma unroll
            for (int r = 0; r < HC_ROWS; r++) {
                const int j = r * HC_THREADS + tid;
                if (j < rows) v[r] = vals.src[c][base + j];
            }
#pragma unroll
            for (int r = 0; r < HC_ROWS; r++) {
                const int j = r * HC_THREADS + tid;
                if (j < rows) sm.v[c][sm.start[part[r]] + rank[r]] = v[r];
            }
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < HC_ROWS; r++) {
            const int j = r * HC_THREADS + tid;
            if (j < rows) {
                const int p = sm.p[j];
                const unsigned long long dst = sm.gbase[p] + (unsigned long long)(j - sm.start[p]);
                pk[dst] = sm.k[j];
#pragma unroll
                for (int c = 0; c < NV; c++) vals.dst[c][dst] = sm.v[c][j];
            }
        }
        __syncthreads();
    }
}

} // namespace srb
