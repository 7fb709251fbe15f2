// srb_radix_agg.cuh — radix-partitioned aggregation for high-cardinality GROUP BY (k BIGINT -> SUM, COUNT; phase 1 and the phase-2 merge).
//
// Reference behaviour matched: the groups of AggHashMapWithOneNumberKey (exec/aggregate/agg_hash_map.h:228-339) with SUM / COUNT update
// (exprs/agg/sum.h:58-63, count.h:46-49) and merge (sum.h:113-117, count.h:92-96) states; the intent of the reference's two-level /
// partitioned tables (agg_hash_map.h:92-98, aggregator.cpp:1132): split the key space until a partition's table fits the nearest cache.
//
// Why: with ~10^8 groups a global table costs one random read-modify-write per row, and both shared-memory and L2 atomics top out near
// 0.4 lane-operations per cycle and SM (measured: 600 M shared atomics = 4.3 ms on a B200), so every atomic per row costs more than the
// whole HBM budget of the query. This path has ONE atomic per row (the partition cursor, in L2) and none after that:
//
//   k_rx_scatter    every row is stored into one of F = 2^10..2^17 partitions, chosen by the top bits of the table hash, at the position
//                   an L2 atomic on the partition's cursor returns. A partition's frontier is one cache line, all frontiers together a
//                   few MB of the 126 MB L2, so the 16-byte stores merge there and leave as full lines: 16 B/row read, 16 B/row written.
//   k_rx_aggregate  one persistent CTA per SM walks partitions (~4.5 k rows, ~750 distinct keys each at six rows per key): the rows arrive
//                   by one 1-D bulk copy (TMA, double buffered), are bucketed by 9 more hash bits with ballot-ranked counting (warp-private
//                   histograms, plain LDS/STS, no atomics), and every thread then folds ONE bucket (~9 rows, 1-2 keys) sequentially in
//                   registers. Groups leave densely packed through a shared-memory staging area: 16 B/row read, 24 B/group written.
//
// Partitions have a fixed capacity (mean + 6 sigma of a uniform hash + slack). A hot key overflows its partition: the kernels count the
// overflowing rows and the host then discards the attempt and runs the table-based path (srb_hcpart.cuh) instead — same result, slower.
#pragma once
#include "srb_rowpipe.cuh"
#include "srb_shapes.cuh"

namespace srb {

constexpr int RX_SB_BITS = 9, RX_SB = 1 << RX_SB_BITS; // buckets inside a partition
constexpr int RX_THREADS = 512, RX_WARPS = RX_THREADS / 32;
static_assert(RX_THREADS == RX_SB, "one thread per bucket in the scan and fold steps");
constexpr int RX_G = 8;                                // register groups per bucket; a bucket with more distinct keys takes the slow path
template <int NV>
struct RxGeom {
    static constexpr int CAP = NV == 1 ? 5120 : 3392;  // rows a partition buffer holds (two buffers of CAP rows must fit shared memory)
    static constexpr int MEAN = NV == 1 ? 4608 : 2944; // target rows per partition: MEAN + 6 sqrt(MEAN) + 64 <= CAP
    static constexpr size_t A_BYTES = (size_t)CAP * 16, B_BYTES = NV == 2 ? (size_t)CAP * 8 : 0;
    static constexpr size_t OFF_A = 0, OFF_B = 2 * A_BYTES, OFF_IDX = OFF_B + 2 * B_BYTES, OFF_DR = OFF_IDX + (size_t)CAP * 2,
                            OFF_HIST = OFF_DR + (size_t)CAP * 4, OFF_START = OFF_HIST + (size_t)RX_WARPS * RX_SB * 2,
                            OFF_MISC = (OFF_START + (size_t)(RX_SB + 1) * 4 + 127) & ~(size_t)127, BYTES = OFF_MISC + 256;
    static_assert(OFF_IDX % 16 == 0 && OFF_DR % 16 == 0 && OFF_HIST % 16 == 0 && OFF_START % 16 == 0 && BYTES <= 232448, "shared-memory layout");
};

struct RxCounters {
    unsigned long long groups;   // dense output rows produced so far
    unsigned long long overflow; // rows that did not fit their partition + groups that did not fit the output
};

SRB_D uint32_t rx_part(uint64_t key, int seed, int fbits) { return (uint32_t)(phmap_mix8(key, seed) >> (64 - fbits)); }
SRB_D uint32_t rx_bucket(uint64_t key, int seed, int fbits) { return (uint32_t)(phmap_mix8(key, seed) >> (64 - fbits - RX_SB_BITS)) & (RX_SB - 1); }

// rows -> partitions. NV = 1: (k, v) pairs; NV = 2: (k, partial sum) pairs + the partial counts in a second array.
template <int NV>
__global__ void __launch_bounds__(256) k_rx_scatter(const uint64_t* __restrict__ keys, const uint64_t* __restrict__ v0, const uint64_t* __restrict__ v1,
                                                     const int64_t n, const int seed, const int fbits, const uint32_t cap, uint32_t* __restrict__ cursor,
                                                     ulonglong2* __restrict__ outA, uint64_t* __restrict__ outB, RxCounters* __restrict__ ctr) {
    constexpr int R = 8;
    const int64_t tile_rows = 256 * R;
    unsigned int lost = 0;
    for (int64_t base = (int64_t)blockIdx.x * tile_rows; base < n; base += (int64_t)gridDim.x * tile_rows) {
        uint64_t k[R], a[R], b[R];
        uint32_t part[R], pos[R];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int64_t i = base + r * 256 + threadIdx.x;
            if (i < n) {
                k[r] = keys[i];
                a[r] = v0[i];
                if (NV == 2) b[r] = v1[i];
            }
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int64_t i = base + r * 256 + threadIdx.x;
            if (i < n) {
                part[r] = rx_part(k[r], seed, fbits);
                pos[r] = atomicAdd(&cursor[part[r]], 1u);
            }
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            const int64_t i = base + r * 256 + threadIdx.x;
            if (i < n) {
                if (pos[r] < cap) {
                    const size_t at = (size_t)part[r] * cap + pos[r];
                    outA[at] = make_ulonglong2(k[r], a[r]);
                    if (NV == 2) outB[at] = b[r];
                } else {
                    lost++;
                }
            }
        }
    }
    if (lost) atomicAdd(&ctr->overflow, (unsigned long long)lost);
}

struct RxOut {
    ulonglong2* keys;  // packed key records {key, 0}, as k_emit writes them
    uint64_t* sum;     // SUM state / result (BIGINT)
    uint8_t* sum_null; // may be nullptr
    uint64_t* cnt;     // COUNT state / result
    uint8_t* cnt_null; // may be nullptr
    unsigned long long row_base, capacity; // output rows [row_base, row_base + capacity)
};

template <int NV>
__global__ void __launch_bounds__(RX_THREADS, 1) k_rx_aggregate(const ulonglong2* __restrict__ partA, const uint64_t* __restrict__ partB,
                                                                 const uint32_t* __restrict__ cursor, const int n_parts, const uint32_t cap, const int seed,
                                                                 const int fbits, const RxOut out, RxCounters* __restrict__ ctr) {
    using G = RxGeom<NV>;
    extern __shared__ __align__(128) uint8_t smem[];
    uint16_t* idx = reinterpret_cast<uint16_t*>(smem + G::OFF_IDX);
    uint32_t* dr = reinterpret_cast<uint32_t*>(smem + G::OFF_DR);
    uint16_t* hist = reinterpret_cast<uint16_t*>(smem + G::OFF_HIST); // [warp][bucket]
    uint32_t* start = reinterpret_cast<uint32_t*>(smem + G::OFF_START);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + G::OFF_MISC);       // two mbarriers
    uint32_t* misc = reinterpret_cast<uint32_t*>(smem + G::OFF_MISC + 16); // [0..15] warp totals, [16] block total, [20..21] output base (u64)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = (1u << lane) - 1u;

    auto rows_of = [&](int p) -> uint32_t {
        const uint32_t c = cursor[p];
        return c < cap ? c : cap;
    };
    auto issue = [&](int p, int b) { // one elected thread: the partition's rows -> buffer b
        const uint32_t np = rows_of(p);
        const uint32_t bytesA = np * 16u, bytesB = NV == 2 ? ((np + 1u) & ~1u) * 8u : 0u;
        if (bytesA == 0) {
            mbar_arrive(&bar[b]);
            return;
        }
        mbar_expect_tx(&bar[b], bytesA + bytesB);
        tma_load_1d(smem + G::OFF_A + (size_t)b * G::A_BYTES, partA + (size_t)p * cap, bytesA, &bar[b]);
        if (NV == 2) tma_load_1d(smem + G::OFF_B + (size_t)b * G::B_BYTES, partB + (size_t)p * cap, bytesB, &bar[b]);
    };

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (tid == 0) {
        if ((int)blockIdx.x < n_parts) issue(blockIdx.x, 0);
        if ((int)(blockIdx.x + gridDim.x) < n_parts) issue(blockIdx.x + gridDim.x, 1);
    }
    uint32_t parity[2] = {0, 0};
    int b = 0;
    for (int p = blockIdx.x; p < n_parts; p += gridDim.x, b ^= 1) {
        const uint32_t np = rows_of(p);
        const ulonglong2* rowsA = reinterpret_cast<const ulonglong2*>(smem + G::OFF_A + (size_t)b * G::A_BYTES);
        const uint64_t* rowsB = reinterpret_cast<const uint64_t*>(smem + G::OFF_B + (size_t)b * G::B_BYTES);
        // ---- A: ballot-ranked counting by bucket, warp-private histograms (no atomics) -------------------------------------------
        {
            uint4* h4 = reinterpret_cast<uint4*>(hist);
            for (int i = tid; i < RX_WARPS * RX_SB * 2 / 16; i += RX_THREADS) h4[i] = make_uint4(0, 0, 0, 0);
        }
        mbar_wait(&bar[b], parity[b]);
        parity[b] ^= 1;
        __syncthreads();
        uint16_t* myhist = hist + warp * RX_SB;
        for (uint32_t i0 = 0; i0 < np; i0 += RX_THREADS) {
            const uint32_t i = i0 + tid;
            const bool on = i < np;
            uint32_t d = 0;
            if (on) d = rx_bucket(rowsA[i].x, seed, fbits);
            uint32_t m = __ballot_sync(0xffffffffu, on);
#pragma unroll
            for (int bit = 0; bit < RX_SB_BITS; bit++) {
                const uint32_t vote = __ballot_sync(0xffffffffu, (d >> bit) & 1u);
                m &= ((d >> bit) & 1u) ? vote : ~vote;
            }
            uint32_t basec = 0;
            if (on) basec = myhist[d];
            __syncwarp();
            if (on) {
                const uint32_t rank = __popc(m & lt);
                if (rank == 0) myhist[d] = (uint16_t)(basec + __popc(m));
                dr[i] = (d << 16) | (basec + rank);
            }
            __syncwarp();
        }
        __syncthreads();
        // ---- S: per bucket, exclusive prefix over the warps; then the buckets' first positions ------------------------------------
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < RX_WARPS; w++) {
            const uint32_t c = hist[w * RX_SB + tid];
            hist[w * RX_SB + tid] = (uint16_t)total;
            total += c;
        }
        uint32_t incl = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) misc[warp] = incl;
        __syncthreads();
        uint32_t wbase = 0;
#pragma unroll
        for (int w = 0; w < RX_WARPS; w++) wbase += w < warp ? misc[w] : 0u;
        const uint32_t s0 = wbase + incl - total; // first position of bucket `tid`
        start[tid] = s0;
        if (tid == RX_THREADS - 1) start[RX_SB] = s0 + total;
        __syncthreads();
        // ---- B: row numbers in bucket order -------------------------------------------------------------------------------------
        for (uint32_t i = tid; i < np; i += RX_THREADS) {
            const uint32_t v = dr[i], d = v >> 16;
            idx[start[d] + myhist[d] + (v & 0xffffu)] = (uint16_t)i;
        }
        __syncthreads();
        // ---- C: one thread folds one bucket in registers --------------------------------------------------------------------------
        uint64_t gk[RX_G], gs[RX_G], gc[RX_G];
        int ng = 0;
        bool slow = false;
        const uint32_t e0 = s0 + total;
        for (uint32_t q = s0; q < e0; q++) {
            const uint32_t i = idx[q];
            const ulonglong2 row = rowsA[i];
            const uint64_t c = NV == 2 ? rowsB[i] : 1ULL;
            int hit = -1;
#pragma unroll
            for (int g = 0; g < RX_G; g++)
                if (g < ng && gk[g] == row.x) hit = g;
            if (hit < 0) {
                if (ng == RX_G) {
                    slow = true;
                    break;
                }
                hit = ng++;
#pragma unroll
                for (int g = 0; g < RX_G; g++)
                    if (g == hit) {
                        gk[g] = row.x;
                        gs[g] = 0;
                        gc[g] = 0;
                    }
            }
#pragma unroll
            for (int g = 0; g < RX_G; g++)
                if (g == hit) {
                    gs[g] += row.y;
                    gc[g] += c;
                }
        }
        if (slow) { // more than RX_G distinct keys in one bucket (rare): quadratic pass over the bucket, groups straight to the output
            ng = 0;
            for (uint32_t q = s0; q < e0; q++) {
                const uint64_t key = rowsA[idx[q]].x;
                bool seen = false;
                for (uint32_t r = s0; r < q && !seen; r++) seen = rowsA[idx[r]].x == key;
                if (seen) continue;
                uint64_t s = 0, c = 0;
                for (uint32_t r = q; r < e0; r++) {
                    const uint32_t i = idx[r];
                    if (rowsA[i].x == key) {
                        s += rowsA[i].y;
                        c += NV == 2 ? rowsB[i] : 1ULL;
                    }
                }
                const unsigned long long at = atomicAdd(&ctr->groups, 1ULL);
                if (at < out.capacity) {
                    const unsigned long long o = out.row_base + at;
                    out.keys[o] = make_ulonglong2(key, 0ULL);
                    out.sum[o] = s;
                    out.cnt[o] = c;
                    if (out.sum_null) out.sum_null[o] = 0;
                    if (out.cnt_null) out.cnt_null[o] = 0;
                } else {
                    atomicAdd(&ctr->overflow, 1ULL);
                }
            }
        }
        // ---- D: groups leave densely packed through a staging area laid over the partition's (consumed) rows ----------------------
        uint32_t gin = (uint32_t)ng;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, gin, o);
            if (lane >= o) gin += y;
        }
        __syncthreads(); // every thread is done reading the rows (and misc[] of the scan above)
        if (lane == 31) misc[warp] = gin;
        __syncthreads();
        uint32_t gbase = 0, gtotal = 0;
#pragma unroll
        for (int w = 0; w < RX_WARPS; w++) {
            gbase += w < warp ? misc[w] : 0u;
            gtotal += misc[w];
        }
        const uint32_t goff = gbase + gin - (uint32_t)ng;
        if (tid == 0) {
            const unsigned long long at = gtotal ? atomicAdd(&ctr->groups, (unsigned long long)gtotal) : 0ULL;
            *reinterpret_cast<unsigned long long*>(&misc[20]) = at;
        }
        uint64_t* stK = reinterpret_cast<uint64_t*>(smem + G::OFF_A + (size_t)b * G::A_BYTES); // CAP x 8 bytes
        uint64_t* stS = stK + G::CAP;                                                            // CAP x 8 bytes (A_BYTES = CAP x 16)
        uint64_t* stC64 = reinterpret_cast<uint64_t*>(smem + G::OFF_B + (size_t)b * G::B_BYTES); // NV == 2
        uint32_t* stC32 = dr;                                                                    // NV == 1: counts <= CAP
#pragma unroll
        for (int g = 0; g < RX_G; g++)
            if (g < ng) {
                stK[goff + g] = gk[g];
                stS[goff + g] = gs[g];
                if (NV == 2)
                    stC64[goff + g] = gc[g];
                else
                    stC32[goff + g] = (uint32_t)gc[g];
            }
        __syncthreads();
        const unsigned long long at0 = *reinterpret_cast<unsigned long long*>(&misc[20]);
        unsigned int lost = 0;
        for (uint32_t j = tid; j < gtotal; j += RX_THREADS) {
            const unsigned long long at = at0 + j;
            if (at < out.capacity) {
                const unsigned long long o = out.row_base + at;
                out.keys[o] = make_ulonglong2(stK[j], 0ULL);
                out.sum[o] = stS[j];
                out.cnt[o] = NV == 2 ? stC64[j] : (uint64_t)stC32[j];
                if (out.sum_null) out.sum_null[o] = 0;
                if (out.cnt_null) out.cnt_null[o] = 0;
            } else {
                lost++;
            }
        }
        if (lost) atomicAdd(&ctr->overflow, (unsigned long long)lost);
        __syncthreads(); // buffer b is free: refill it with the partition two rounds ahead
        if (tid == 0) {
            const int nxt = p + 2 * (int)gridDim.x;
            if (nxt < n_parts) {
                fence_proxy_async(); // the staging stores (generic proxy) are ordered before the bulk copy's writes (async proxy)
                issue(nxt, b);
            }
        }
    }
}

// the groups of an earlier radix batch -> the group table (a later batch goes through the table, and a key may occur in both)
__global__ void __launch_bounds__(256) k_rx_ingest(const __grid_constant__ DProg prog, const TableView tv, const ulonglong2* __restrict__ keys,
                                                    const uint64_t* __restrict__ sum, const uint64_t* __restrict__ cnt, const int64_t n,
                                                    unsigned long long* __restrict__ inserted, unsigned long long* __restrict__ failed) {
    const DProg& p = prog;
    const DAgg& gs = p.aggs[0];
    const DAgg& gc = p.aggs[1];
    unsigned long long ins = 0, bad = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t key = keys[i].x;
        bool fresh = false;
        uint8_t* slot = table_find_or_insert(tv, key, 0ULL, phmap_mix8(key, p.phase_seed), &fresh);
        if (!slot) {
            bad++;
            continue;
        }
        ins += fresh ? 1 : 0;
        atomicAdd((unsigned long long*)(slot + gs.off_lo), (unsigned long long)sum[i]);
        if (gs.off_n >= 0) atomicAdd((unsigned long long*)(slot + gs.off_n), (unsigned long long)cnt[i]);
        atomicAdd((unsigned long long*)(slot + gc.off_n), (unsigned long long)cnt[i]);
    }
    if (ins) atomicAdd(inserted, ins);
    if (bad) atomicAdd(failed, bad);
}

} // namespace srb
