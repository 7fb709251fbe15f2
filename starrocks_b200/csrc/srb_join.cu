// srb_join.cu — hash join build / probe / late-materialised output (K6 / K7 / K8) and the operator pair around them.
//
// Stands in for, and keeps the call contract of:
//   HashJoinBuildOperator::push_chunk / set_finishing         be/src/exec/pipeline/hashjoin/hash_join_build_operator.cpp:40-213
//   HashJoinProbeOperator::push_chunk / pull_chunk / need_input / has_output   .../hash_join_probe_operator.cpp:55-104
//   JoinHashTable::append_chunk (row 0 = dummy, 1-based indexes)               be/src/exec/join_hash_map.cpp:613-642
//   JoinBuildFunc::prepare / construct_hash_table               be/src/exec/join_hash_map.tpp:29-91
//   JoinProbeFunc::lookup_init + _do_probe_from_ht (inner)      be/src/exec/join_hash_map.tpp:283-309,1066-1143
//   serialize_fixed_size_key_column (multi-column fixed keys)   be/src/exec/join_hash_map.h:391-406
//   _probe_output / _build_output (append_selective by index)   be/src/exec/join_hash_map.tpp:534-550,624-640
//
// Table layout is the reference's: first[bucket_size] + next[rows+1] of uint32, bucket = JoinKeyHash(key), bucket_size
// from calc_bucket_size, build rows numbered from 1 with 0 = end of chain, NULL keys never inserted. What changes is
// the execution: the build links rows with one atomicExch per row instead of a serial loop; the probe runs one thread
// per probe row over the whole pushed range (thousands of outstanding loads replace the coroutine interleaving of
// join_hash_map.tpp:1146-1184) in two passes — count, then warp-aggregated append of (probe_index, build_index) pairs —
// and payload columns are gathered only for matched rows.
#include <algorithm>
#include <cstring>
#include <memory>

#include "srb_dev.cuh"
#include "srb_host.h"

namespace srb {

constexpr int JMAX_KEYS = 4, JMAX_OUT = 16;

struct JKeyCol {
    const uint8_t* data;
    const uint8_t* nulls;
    int32_t width;
    int32_t _pad;
};
struct JPackArgs {
    int32_t n_keys;
    int32_t width; // packed width: 4, 8 or 16
    JKeyCol cols[JMAX_KEYS];
};

// packs the key columns of rows [0,n) column by column at increasing byte offsets, zero padded to `width`
__global__ void __launch_bounds__(256) k_join_pack(const __grid_constant__ JPackArgs a, int64_t n, uint8_t* __restrict__ out,
                                                    uint8_t* __restrict__ out_null) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        uint64_t w[2] = {0, 0};
        int off = 0;
        uint8_t isnull = 0;
        for (int k = 0; k < a.n_keys; k++) {
            const JKeyCol& c = a.cols[k];
            const uint8_t* p = c.data + (size_t)r * c.width;
            for (int b = 0; b < c.width; b++, off++) w[off >> 3] |= (uint64_t)p[b] << (8 * (off & 7));
            if (c.nulls && c.nulls[r]) isnull = 1;
        }
        if (a.width == 4)
            ((uint32_t*)out)[r] = (uint32_t)w[0];
        else if (a.width == 8)
            ((uint64_t*)out)[r] = w[0];
        else
            ((ulonglong2*)out)[r] = make_ulonglong2(w[0], w[1]);
        out_null[r] = isnull;
    }
}

struct JTable {
    const uint8_t* keys; // packed build keys, row i (1-based) at index i-1
    const uint8_t* key_null;
    uint32_t* first;
    uint32_t* next;
    uint32_t bucket_size, log_bucket, rows;
    int32_t width;
};

SRB_D uint32_t jbucket(const JTable& t, uint64_t lo, uint64_t hi) {
    if (t.width == 4) return join_key_hash32((uint32_t)lo, t.log_bucket);
    if (t.width == 8) return join_key_hash64(lo, t.log_bucket);
    return join_key_hash128(lo, hi, t.bucket_size);
}
SRB_D void jload(const uint8_t* keys, int width, int64_t i, uint64_t* lo, uint64_t* hi) {
    *hi = 0;
    if (width == 4)
        *lo = ((const uint32_t*)keys)[i];
    else if (width == 8)
        *lo = ((const uint64_t*)keys)[i];
    else {
        ulonglong2 v = ((const ulonglong2*)keys)[i];
        *lo = v.x;
        *hi = v.y;
    }
}

// K6: next[i] = exchange(first[bucket], i) — the chained table of join_hash_map.tpp:59-91 built in parallel
__global__ void __launch_bounds__(256) k_join_build(JTable t) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x + 1; i <= t.rows; i += gridDim.x * blockDim.x) {
        if (t.key_null[i - 1]) continue;
        uint64_t lo, hi;
        jload(t.keys, t.width, i - 1, &lo, &hi);
        uint32_t b = jbucket(t, lo, hi);
        t.next[i] = atomicExch(&t.first[b], i);
    }
}

// K7: mode 0 counts matches (block-reduced into *total); mode 1 appends (probe_index, build_index) pairs
// Outer / right-side joins (join_hash_map.tpp:1187-1986): keep_left = a probe row without a match still yields one output row with build
// index 0 (LEFT / FULL OUTER); matched = one byte per build row (1-based) set when the row matches (RIGHT / FULL OUTER, RIGHT SEMI / ANTI:
// HashTableProbeState::build_match_index); pairs = 0: only the flags are wanted (RIGHT SEMI / ANTI), nothing is counted or appended.
template <int MODE>
__global__ void __launch_bounds__(256) k_join_probe(JTable t, const uint8_t* __restrict__ pkeys, const uint8_t* __restrict__ pnull, int64_t n,
                                                     int semi, unsigned long long* __restrict__ total, uint32_t* __restrict__ out_probe,
                                                     uint32_t* __restrict__ out_build, int keep_left = 0, uint8_t* __restrict__ matched = nullptr,
                                                     int pairs = 1) {
    unsigned long long mine = 0;
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t n_round = (n + 31) & ~(int64_t)31; // keep warps converged for the ballots of mode 1
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_round; r += stride) {
        uint32_t cnt = 0;
        uint32_t bi = 0;
        uint64_t lo = 0, hi = 0;
        if (r < n && !pnull[r]) {
            jload(pkeys, t.width, r, &lo, &hi);
            bi = t.first[jbucket(t, lo, hi)];
        }
        if (MODE == 0) {
            while (bi) {
                uint64_t blo, bhi;
                jload(t.keys, t.width, bi - 1, &blo, &bhi);
                if (blo == lo && bhi == hi) {
                    cnt++;
                    if (matched) matched[bi] = 1;
                    if (semi) break;
                }
                bi = t.next[bi];
            }
            if (keep_left && r < n && cnt == 0) cnt = 1;
            mine += pairs ? cnt : 0;
        } else {
            // walk the chain; every match takes a ticket from a warp-aggregated atomic
            bool ever = false;
            for (;;) {
                bool hit = false;
                while (bi) {
                    uint64_t blo, bhi;
                    jload(t.keys, t.width, bi - 1, &blo, &bhi);
                    if (blo == lo && bhi == hi) {
                        hit = true;
                        break;
                    }
                    bi = t.next[bi];
                }
                unsigned m = __ballot_sync(0xffffffffu, hit);
                if (!m) break;
                unsigned long long base = 0;
                int leader = __ffs(m) - 1;
                if (lane == leader) base = atomicAdd(total, (unsigned long long)__popc(m));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (hit) {
                    unsigned long long pos = base + __popc(m & ((1u << lane) - 1));
                    out_probe[pos] = (uint32_t)r;
                    out_build[pos] = bi; // 1-based, as the reference's build_index
                    bi = semi ? 0 : t.next[bi];
                    ever = true;
                }
            }
            if (keep_left) { // the probe rows nothing matched: one output row each, build index 0 = "no build row"
                const bool lone = r < n && !ever;
                const unsigned m = __ballot_sync(0xffffffffu, lone);
                if (m) {
                    unsigned long long base = 0;
                    const int leader = __ffs(m) - 1;
                    if (lane == leader) base = atomicAdd(total, (unsigned long long)__popc(m));
                    base = __shfl_sync(0xffffffffu, base, leader);
                    if (lone) {
                        const unsigned long long pos = base + __popc(m & ((1u << lane) - 1));
                        out_probe[pos] = (uint32_t)r;
                        out_build[pos] = 0;
                    }
                }
            }
        }
    }
    if (MODE == 0) {
        for (int o = 16; o > 0; o >>= 1) mine += __shfl_down_sync(0xffffffffu, mine, o);
        if (lane == 0 && mine) atomicAdd(total, mine);
    }
}

// K8: late materialisation — payload columns gathered for matched rows only
struct GatherCol {
    const uint8_t* src;
    uint8_t* dst;
    const uint8_t* src_null;
    uint8_t* dst_null;
    int32_t width;
    int32_t from_build; // index array to use
};
struct GatherArgs {
    int32_t n_cols;
    int32_t _pad;
    GatherCol cols[JMAX_OUT];
};
__global__ void __launch_bounds__(256) k_join_gather(const __grid_constant__ GatherArgs g, int64_t n, const uint32_t* __restrict__ probe_idx,
                                                      const uint32_t* __restrict__ build_idx) {
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
        // probe_idx == nullptr / build index 0: that side is missing in this output row (outer joins): its columns are NULL
        const uint32_t pi = probe_idx ? probe_idx[j] : 0xFFFFFFFFu, bi = build_idx[j] - 1;
        for (int c = 0; c < g.n_cols; c++) {
            const GatherCol& gc = g.cols[c];
            const size_t s = gc.from_build ? bi : pi;
            if (s == 0xFFFFFFFFu) {
                switch (gc.width) {
                case 1: gc.dst[j] = 0; break;
                case 2: ((uint16_t*)gc.dst)[j] = 0; break;
                case 4: ((uint32_t*)gc.dst)[j] = 0; break;
                case 8: ((uint64_t*)gc.dst)[j] = 0; break;
                default: ((ulonglong2*)gc.dst)[j] = make_ulonglong2(0, 0);
                }
                if (gc.dst_null) gc.dst_null[j] = 1;
                continue;
            }
            switch (gc.width) {
            case 1: gc.dst[j] = gc.src[s]; break;
            case 2: ((uint16_t*)gc.dst)[j] = ((const uint16_t*)gc.src)[s]; break;
            case 4: ((uint32_t*)gc.dst)[j] = ((const uint32_t*)gc.src)[s]; break;
            case 8: ((uint64_t*)gc.dst)[j] = ((const uint64_t*)gc.src)[s]; break;
            default: ((ulonglong2*)gc.dst)[j] = ((const ulonglong2*)gc.src)[s];
            }
            if (gc.dst_null) gc.dst_null[j] = gc.src_null ? gc.src_null[s] : 0;
        }
    }
}

// other-join-conjuncts of an INNER join (HashJoiner::_process_other_conjunct, hash_joiner.cpp:488-507): the (probe, build) pairs whose
// columns satisfy every comparison survive; a NULL on either side fails it. Pairs are compacted with one atomic per warp step.
struct ConjCol {
    const uint8_t* data;
    const uint8_t* nulls;
    int32_t width;
    int32_t from_build;
};
struct ConjArgs {
    int32_t n;
    int32_t _pad;
    ConjCol l[8], r[8];
    int32_t op[8];
};
SRB_D i128 conj_load(const ConjCol& c, size_t row) {
    switch (c.width) {
    case 1: return (i128)((const int8_t*)c.data)[row];
    case 2: return (i128)((const int16_t*)c.data)[row];
    case 4: return (i128)((const int32_t*)c.data)[row];
    case 8: return (i128)((const int64_t*)c.data)[row];
    default: {
        const ulonglong2 v = ((const ulonglong2*)c.data)[row];
        return make_i128(v.x, (int64_t)v.y);
    }
    }
}
__global__ void __launch_bounds__(256) k_join_conjuncts(const __grid_constant__ ConjArgs a, int64_t m, const uint32_t* __restrict__ pi_in,
                                                         const uint32_t* __restrict__ bi_in, uint32_t* __restrict__ pi_out, uint32_t* __restrict__ bi_out,
                                                         unsigned long long* __restrict__ total) {
    const int lane = threadIdx.x & 31;
    const int64_t rounds = (m + 31) & ~(int64_t)31;
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < rounds; j += (int64_t)gridDim.x * blockDim.x) {
        bool keep = j < m;
        uint32_t pi = 0, bi = 0;
        if (keep) {
            pi = pi_in[j];
            bi = bi_in[j];
            for (int c = 0; c < a.n && keep; c++) {
                const size_t lr = a.l[c].from_build ? bi - 1 : pi, rr = a.r[c].from_build ? bi - 1 : pi;
                if ((a.l[c].nulls && a.l[c].nulls[lr]) || (a.r[c].nulls && a.r[c].nulls[rr])) {
                    keep = false;
                    break;
                }
                const i128 x = conj_load(a.l[c], lr), y = conj_load(a.r[c], rr);
                switch (a.op[c]) {
                case 1: keep = x == y; break;
                case 2: keep = x != y; break;
                case 3: keep = x > y; break;
                case 4: keep = x >= y; break;
                case 5: keep = x < y; break;
                default: keep = x <= y;
                }
            }
        }
        const unsigned msk = __ballot_sync(0xffffffffu, keep);
        if (!msk) continue;
        unsigned long long base = 0;
        const int leader = __ffs(msk) - 1;
        if (lane == leader) base = atomicAdd(total, (unsigned long long)__popc(msk));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (keep) {
            const unsigned long long at = base + __popc(msk & ((1u << lane) - 1));
            pi_out[at] = pi;
            bi_out[at] = bi;
        }
    }
}

// the build rows (1-based) whose matched flag equals `want`, in any order (_search_ht_remain, join_hash_map.tpp)
__global__ void __launch_bounds__(256) k_join_remainder(const uint8_t* __restrict__ matched, uint32_t rows, int want, unsigned long long* __restrict__ total,
                                                         uint32_t* __restrict__ out_build) {
    const int lane = threadIdx.x & 31;
    const uint32_t rounds = (rows + 31) & ~31u;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < rounds; j += gridDim.x * blockDim.x) {
        const bool take = j < rows && (matched[j + 1] != 0) == (want != 0);
        const unsigned m = __ballot_sync(0xffffffffu, take);
        if (!m) continue;
        unsigned long long base = 0;
        const int leader = __ffs(m) - 1;
        if (lane == leader) base = atomicAdd(total, (unsigned long long)__popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (take && out_build) out_build[base + __popc(m & ((1u << lane) - 1))] = j + 1;
    }
}

// ---- the tiled probe: at most one match per probe row ---------------------------------------------------------------------------
// Joins against a key (every TPC-H / SSB join of the configs: the build side is a primary key or a filtered slice of one) and
// LEFT SEMI joins produce at most one output row per probe row. Then the probe is two streaming passes over the probe rows:
//   pass A  per row, the build index of its match (0 = none) into match[], and the number of matches per 2048-row tile;
//           a thread walks the chains of its 8 rows round by round, so 8 independent loads are in flight per thread;
//   scan    exclusive prefix of the tile counts (the output offset of every tile, and the output size for the host);
//   pass B  per tile, matched rows ranked in row order and the output columns written straight from the probe columns
//           (row r) and the build columns (match[r]) — no (probe_index, build_index) arrays unless the plan asks for them,
//           no separate gather kernel, and the output keeps the probe row order like the reference's probe loop.
// k_join_unique decides after the build whether the table qualifies for INNER joins.
constexpr int JT_THREADS = 256, JT_ROWS = 8, JT_TILE = JT_THREADS * JT_ROWS;

// sets *dup when two build rows carry the same key (checked per row against the rest of its chain)
__global__ void __launch_bounds__(256) k_join_unique(JTable t, unsigned int* __restrict__ dup) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x + 1; i <= t.rows; i += gridDim.x * blockDim.x) {
        if (t.key_null[i - 1]) continue;
        uint64_t lo, hi;
        jload(t.keys, t.width, i - 1, &lo, &hi);
        for (uint32_t bi = t.next[i]; bi; bi = t.next[bi]) {
            uint64_t blo, bhi;
            jload(t.keys, t.width, bi - 1, &blo, &bhi);
            if (blo == lo && bhi == hi) {
                *dup = 1u;
                return;
            }
        }
    }
}

template <int W>
SRB_D void jload_w(const uint8_t* keys, int64_t i, uint64_t* lo, uint64_t* hi) {
    *hi = 0;
    if constexpr (W == 4)
        *lo = ((const uint32_t*)keys)[i];
    else if constexpr (W == 8)
        *lo = ((const uint64_t*)keys)[i];
    else {
        const ulonglong2 v = ((const ulonglong2*)keys)[i];
        *lo = v.x;
        *hi = v.y;
    }
}
template <int W>
SRB_D uint32_t jbucket_w(const JTable& t, uint64_t lo, uint64_t hi) {
    if constexpr (W == 4) return join_key_hash32((uint32_t)lo, t.log_bucket);
    if constexpr (W == 8) return join_key_hash64(lo, t.log_bucket);
    return join_key_hash128(lo, hi, t.bucket_size);
}

template <int W>
__global__ void __launch_bounds__(JT_THREADS) k_join_probe_first(JTable t, const uint8_t* __restrict__ pkeys, const uint8_t* __restrict__ pnull, int64_t n,
                                                                  uint32_t* __restrict__ match, uint32_t* __restrict__ tile_counts, int anti) {
    __shared__ uint32_t warp_cnt[JT_THREADS / 32];
    const int64_t n_tiles = (n + JT_TILE - 1) / JT_TILE;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        uint64_t lo[JT_ROWS], hi[JT_ROWS];
        uint32_t bi[JT_ROWS], m[JT_ROWS];
        bool any = false;
#pragma unroll
        for (int j = 0; j < JT_ROWS; j++) {
            const int64_t r = tile * JT_TILE + j * JT_THREADS + threadIdx.x;
            bi[j] = m[j] = 0;
            lo[j] = hi[j] = 0;
            if (r < n && !(pnull && pnull[r])) {
                jload_w<W>(pkeys, r, &lo[j], &hi[j]);
                bi[j] = t.first[jbucket_w<W>(t, lo[j], hi[j])];
            }
            any = any || bi[j] != 0;
        }
        while (any) { // one chain step of every row per round
            any = false;
#pragma unroll
            for (int j = 0; j < JT_ROWS; j++) {
                if (bi[j]) {
                    uint64_t blo, bhi;
                    jload_w<W>(t.keys, bi[j] - 1, &blo, &bhi);
                    if (blo == lo[j] && (W < 16 || bhi == hi[j])) {
                        m[j] = bi[j];
                        bi[j] = 0;
                    } else {
                        bi[j] = t.next[bi[j]];
                    }
                    any = any || bi[j] != 0;
                }
            }
        }
        uint32_t cnt = 0;
#pragma unroll
        for (int j = 0; j < JT_ROWS; j++) {
            const int64_t r = tile * JT_TILE + j * JT_THREADS + threadIdx.x;
            if (anti) m[j] = r < n && m[j] == 0 ? 1u : 0u; // LEFT ANTI keeps exactly the rows without a match (NULL keys never match)
            if (anti == 2 && m[j] && pnull && pnull[r]) m[j] = 0; // NULL-aware (NOT IN): a NULL probe key is never "not in" a non-empty set
            if (r < n) match[r] = m[j];
            cnt += m[j] ? 1u : 0u;
        }
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, o);
        if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = cnt;
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t s = 0;
            for (int w = 0; w < JT_THREADS / 32; w++) s += warp_cnt[w];
            tile_counts[tile] = s;
        }
        __syncthreads();
    }
}

// does any build row carry a NULL key? (NULL-aware LEFT ANTI: then NOT IN is never true, hash_joiner.h:386-396)
__global__ void __launch_bounds__(256) k_join_any_null(const uint8_t* __restrict__ nulls, int64_t n, unsigned int* __restrict__ flag) {
    bool any = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) any = any || nulls[i] != 0;
    if (__any_sync(0xffffffffu, any) && (threadIdx.x & 31) == 0) atomicOr(flag, 1u);
}

// LEFT ANTI against an empty build side: every probe row is kept
__global__ void __launch_bounds__(JT_THREADS) k_join_all_rows(int64_t n, uint32_t* __restrict__ match, uint32_t* __restrict__ tile_counts) {
    const int64_t n_tiles = (n + JT_TILE - 1) / JT_TILE;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int j = 0; j < JT_ROWS; j++) {
            const int64_t r = tile * JT_TILE + j * JT_THREADS + threadIdx.x;
            if (r < n) match[r] = 1u;
        }
        if (threadIdx.x == 0) tile_counts[tile] = (uint32_t)((n - tile * JT_TILE) < JT_TILE ? (n - tile * JT_TILE) : JT_TILE);
    }
}

// exclusive prefix of the tile counts (one block; the count of tiles is rows / 2048) and the grand total
__global__ void __launch_bounds__(1024) k_join_scan(const uint32_t* __restrict__ counts, uint64_t* __restrict__ offsets, int64_t n_tiles,
                                                     unsigned long long* __restrict__ total) {
    __shared__ uint64_t warp_sum[32];
    __shared__ uint64_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t base = 0; base < n_tiles; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const uint64_t v = i < n_tiles ? counts[i] : 0;
        uint64_t x = v;
        for (int o = 1; o < 32; o <<= 1) {
            const uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_sum[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint64_t w = warp_sum[lane];
            for (int o = 1; o < 32; o <<= 1) {
                const uint64_t y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            warp_sum[lane] = w; // inclusive over warps
        }
        __syncthreads();
        const uint64_t before = carry + (warp ? warp_sum[warp - 1] : 0) + (x - v);
        if (i < n_tiles) offsets[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

// pass B: rank the tile's matched rows in row order and write the output columns (and, when asked, the index pairs).
// One barrier per tile: the eight slabs' ballots are taken first, their per-warp counts go to shared memory together, and every
// thread derives the output positions of its eight rows from them. Columns are then copied one at a time with the element width
// a compile-time constant and all of a thread's loads issued before its first store (probe columns: row r, coalesced with holes;
// build columns: row match[r]-1, random), so eight independent loads per thread are in flight instead of one.
template <class T, bool FROM_BUILD>
SRB_D void emit_col(const T* __restrict__ src, T* __restrict__ dst, int64_t row0, const uint32_t (&m)[JT_ROWS], const uint32_t (&rel)[JT_ROWS]) {
    constexpr int U = sizeof(T) > 8 ? 4 : 8;
#pragma unroll
    for (int j0 = 0; j0 < JT_ROWS; j0 += U) {
        T v[U] = {};
#pragma unroll
        for (int u = 0; u < U; u++)
            if (m[j0 + u]) v[u] = FROM_BUILD ? src[m[j0 + u] - 1] : src[row0 + (j0 + u) * JT_THREADS];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (m[j0 + u]) dst[rel[j0 + u]] = v[u];
    }
}
template <bool FROM_BUILD>
SRB_D void emit_col_w(const GatherCol& gc, uint64_t base, int64_t row0, const uint32_t (&m)[JT_ROWS], const uint32_t (&rel)[JT_ROWS]) {
    switch (gc.width) {
    case 1: emit_col<uint8_t, FROM_BUILD>((const uint8_t*)gc.src, (uint8_t*)gc.dst + base, row0, m, rel); break;
    case 2: emit_col<uint16_t, FROM_BUILD>((const uint16_t*)gc.src, (uint16_t*)gc.dst + base, row0, m, rel); break;
    case 4: emit_col<uint32_t, FROM_BUILD>((const uint32_t*)gc.src, (uint32_t*)gc.dst + base, row0, m, rel); break;
    case 8: emit_col<uint64_t, FROM_BUILD>((const uint64_t*)gc.src, (uint64_t*)gc.dst + base, row0, m, rel); break;
    default: emit_col<ulonglong2, FROM_BUILD>((const ulonglong2*)gc.src, (ulonglong2*)gc.dst + base, row0, m, rel);
    }
    if (gc.src_null) emit_col<uint8_t, FROM_BUILD>(gc.src_null, gc.dst_null + base, row0, m, rel);
}

__global__ void __launch_bounds__(JT_THREADS, 2) k_join_emit(const __grid_constant__ GatherArgs g, int64_t n, const uint32_t* __restrict__ match,
                                                              const uint64_t* __restrict__ tile_offsets, uint32_t* __restrict__ out_probe,
                                                              uint32_t* __restrict__ out_build) {
    constexpr int NW = JT_THREADS / 32;
    __shared__ uint32_t cnt[JT_ROWS][NW];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t n_tiles = (n + JT_TILE - 1) / JT_TILE;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint64_t base = tile_offsets[tile];
        const int64_t row0 = tile * JT_TILE + threadIdx.x;
        uint32_t m[JT_ROWS], rel[JT_ROWS], bal[JT_ROWS];
#pragma unroll
        for (int j = 0; j < JT_ROWS; j++) {
            const int64_t r = row0 + j * JT_THREADS;
            m[j] = r < n ? match[r] : 0;
        }
#pragma unroll
        for (int j = 0; j < JT_ROWS; j++) { // slab j = 256 consecutive rows
            bal[j] = __ballot_sync(0xffffffffu, m[j] != 0);
            if (lane == 0) cnt[j][warp] = __popc(bal[j]);
        }
        __syncthreads();
        uint32_t run = 0; // matched rows of the tile before slab j
#pragma unroll
        for (int j = 0; j < JT_ROWS; j++) {
            uint32_t before = 0, slab = 0;
#pragma unroll
            for (int w = 0; w < NW; w++) {
                const uint32_t c = cnt[j][w];
                before += w < warp ? c : 0;
                slab += c;
            }
            rel[j] = run + before + __popc(bal[j] & ((1u << lane) - 1));
            run += slab;
        }
        __syncthreads(); // cnt[] is rewritten by the next tile
        if (run == 0) continue;
        for (int c = 0; c < g.n_cols; c++) {
            const GatherCol& gc = g.cols[c];
            if (gc.from_build) emit_col_w<true>(gc, base, row0, m, rel);
            else emit_col_w<false>(gc, base, row0, m, rel);
        }
        if (out_probe) {
#pragma unroll
            for (int j = 0; j < JT_ROWS; j++) {
                if (m[j]) {
                    out_probe[base + rel[j]] = (uint32_t)(row0 + j * JT_THREADS);
                    out_build[base + rel[j]] = m[j];
                }
            }
        }
    }
}

// ---- runtime filter from the build side ---------------------------------------------------------------------------------------------
// HashJoinBuildOperator::set_finishing -> HashJoiner::create_runtime_filters (hash_join_build_operator.cpp:97-208, hash_joiner.cpp:516-611)
// -> RuntimeBloomFilter::insert per build key (runtime_filter.h:1200-1205) + min/max (RuntimeFilterHelper). One thread per build
// row: eight atomicOr into the key's bucket, warp-reduced min / max.
struct RfMeta {
    long long min, max;
    unsigned long long inserted;
    unsigned int has_null;
    unsigned int _pad;
};
template <int W>
__global__ void __launch_bounds__(256) k_rf_build(const uint8_t* __restrict__ keys, const uint8_t* __restrict__ nulls, int64_t n, RfView v,
                                                   RfMeta* __restrict__ meta) {
    long long mn = INT64_MAX, mx = INT64_MIN;
    unsigned long long ins = 0;
    unsigned int any_null = 0;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        if (nulls && nulls[r]) {
            any_null = 1;
            continue;
        }
        long long x;
        if constexpr (W == 1) x = ((const int8_t*)keys)[r];
        else if constexpr (W == 2) x = ((const int16_t*)keys)[r];
        else if constexpr (W == 4) x = ((const int32_t*)keys)[r];
        else x = ((const long long*)keys)[r];
        const uint64_t hash = rf_hash(x, v.mix); // std::hash of an integer = its sign-extended value, then phmap_mix<8> (identity by default)
        uint32_t* b = v.dir + (size_t)((uint32_t)hash & v.directory_mask) * 8;
        const uint32_t key = (uint32_t)(hash >> v.log_num_buckets);
#pragma unroll
        for (int i = 0; i < 8; i++) atomicOr(&b[i], rf_word_mask(key, i));
        mn = x < mn ? x : mn;
        mx = x > mx ? x : mx;
        ins++;
    }
    for (int o = 16; o > 0; o >>= 1) {
        long long a = __shfl_down_sync(0xffffffffu, mn, o), c = __shfl_down_sync(0xffffffffu, mx, o);
        mn = a < mn ? a : mn;
        mx = c > mx ? c : mx;
        ins += __shfl_down_sync(0xffffffffu, ins, o);
        any_null |= __shfl_down_sync(0xffffffffu, any_null, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (ins) {
            atomicMin(&meta->min, mn);
            atomicMax(&meta->max, mx);
            atomicAdd(&meta->inserted, ins);
        }
        if (any_null) atomicOr(&meta->has_null, 1u);
    }
}

// RuntimeBitsetFilter: bit (x - min) for every non-NULL build key (Bitset::insert, exprs/runtime_filter.cpp:287-292)
template <int W>
__global__ void __launch_bounds__(256) k_rf_bits_build(const uint8_t* __restrict__ keys, const uint8_t* __restrict__ nulls, int64_t n, long long mn,
                                                        uint32_t* __restrict__ bits) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        if (nulls && nulls[r]) continue;
        long long x;
        if constexpr (W == 1) x = ((const int8_t*)keys)[r];
        else if constexpr (W == 2) x = ((const int16_t*)keys)[r];
        else if constexpr (W == 4) x = ((const int32_t*)keys)[r];
        else x = ((const long long*)keys)[r];
        const uint64_t b = (uint64_t)x - (uint64_t)mn;
        atomicOr(&bits[b >> 5], 1u << (b & 31));
    }
}

// =================================================================================================================
struct GrowCol { // device column that grows by appends
    DevBuf data, nulls;
    int type = 0, precision = 0, scale = 0, width = 0;
    bool nullable = false, known = false;
};

class JoinOp : public srb_op {
public:
    JoinOp(srb_ctx* c, const srb_join_plan* p) : srb_op(c) {
        if (!p) throw Error(SRB_INVALID_ARGUMENT, "null plan");
        const int jt = p->join_type;
        if (jt != SRB_JOIN_INNER && jt != SRB_JOIN_LEFT_SEMI && jt != SRB_JOIN_LEFT_ANTI && jt != SRB_JOIN_LEFT_OUTER && jt != SRB_JOIN_RIGHT_OUTER &&
            jt != SRB_JOIN_FULL_OUTER && jt != SRB_JOIN_RIGHT_SEMI && jt != SRB_JOIN_RIGHT_ANTI && jt != SRB_JOIN_NULL_AWARE_LEFT_ANTI)
            throw Error(SRB_NOT_IMPLEMENTED, fmt("join type %d", p->join_type));
        keep_left_ = jt == SRB_JOIN_LEFT_OUTER || jt == SRB_JOIN_FULL_OUTER;
        track_right_ = jt == SRB_JOIN_RIGHT_OUTER || jt == SRB_JOIN_FULL_OUTER || jt == SRB_JOIN_RIGHT_SEMI || jt == SRB_JOIN_RIGHT_ANTI;
        right_only_ = jt == SRB_JOIN_RIGHT_SEMI || jt == SRB_JOIN_RIGHT_ANTI;
        right_want_ = jt == SRB_JOIN_RIGHT_SEMI ? 1 : 0;
        if (p->n_keys < 1 || p->n_keys > JMAX_KEYS) throw Error(SRB_NOT_IMPLEMENTED, fmt("join with %d key columns", p->n_keys));
        if (p->n_build_out < 0 || p->n_probe_out < 0 || p->n_build_out + p->n_probe_out > JMAX_OUT)
            throw Error(SRB_NOT_IMPLEMENTED, "too many join output columns");
        null_aware_ = p->join_type == SRB_JOIN_NULL_AWARE_LEFT_ANTI;
        anti_ = p->join_type == SRB_JOIN_LEFT_ANTI || null_aware_;
        semi_ = p->join_type == SRB_JOIN_LEFT_SEMI || anti_; // an existence test either way: at most one output row per probe row
        bkeys_.assign(p->build_key_slots, p->build_key_slots + p->n_keys);
        pkeys_.assign(p->probe_key_slots, p->probe_key_slots + p->n_keys);
        bout_.assign(p->build_out_slots, p->build_out_slots + p->n_build_out);
        pout_.assign(p->probe_out_slots, p->probe_out_slots + p->n_probe_out);
        if (semi_) bout_.clear(); // LEFT SEMI emits probe columns only
        if (right_only_) pout_.clear(); // RIGHT SEMI / ANTI emit build columns only
        emit_indexes_ = p->emit_indexes != 0 && !anti_ && !keep_left_ && !track_right_;
        if (p->n_conjuncts < 0 || p->n_conjuncts > 8 || (p->n_conjuncts && !p->conjuncts)) throw Error(SRB_INVALID_ARGUMENT, "bad other-join-conjunct list (at most 8)");
        if (p->n_conjuncts && jt != SRB_JOIN_INNER) throw Error(SRB_NOT_IMPLEMENTED, "other-join-conjuncts: INNER joins only");
        conj_.assign(p->conjuncts, p->conjuncts + p->n_conjuncts);
        for (const srb_join_conjunct& cj : conj_)
            if (cj.op < SRB_PRED_EQ || cj.op > SRB_PRED_LE) throw Error(SRB_INVALID_ARGUMENT, "other-join-conjunct: comparison operators only");
        chunk_size_ = c->cfg.chunk_size > 0 ? c->cfg.chunk_size : 4096;
        bcols_.resize(bkeys_.size() + bout_.size());
        total_dev_.ensure(c, 8);
        total_host_.ensure(8);
    }
    ~JoinOp() override { cudaStreamSynchronize(stream); }

    // ---- build side ---------------------------------------------------------------------------------------------
    void push_chunk(const srb_chunk* ch) override {
        if (finishing) throw Error(SRB_INTERNAL_ERROR, "push_chunk after set_finishing");
        if (!ch || ch->n_rows < 0) throw Error(SRB_INVALID_ARGUMENT, "bad chunk");
        if (ch->n_rows == 0) { // an empty chunk still tells the column types (an outer join against nothing emits typed NULL columns)
            for (size_t i = 0; i < bcols_.size(); i++) {
                const int slot = i < bkeys_.size() ? bkeys_[i] : bout_[i - bkeys_.size()];
                for (int c = 0; c < ch->n_cols; c++)
                    if (ch->slot_ids[c] == slot) describe(bcols_[i], &ch->cols[c], slot);
            }
            return;
        }
        if ((uint64_t)build_rows_ + (uint64_t)ch->n_rows >= ((uint64_t)1 << 32) - 1)
            throw Error(SRB_NOT_IMPLEMENTED, "build side exceeds 2^32 rows (uint32 first/next, join_hash_map.h:101-120)");
        for (size_t i = 0; i < bcols_.size(); i++) {
            int slot = i < bkeys_.size() ? bkeys_[i] : bout_[i - bkeys_.size()];
            append_col(bcols_[i], find_or_throw(ch, slot), slot, build_rows_, ch->n_rows);
        }
        build_rows_ += ch->n_rows;
        rows_in += ch->n_rows;
    }

    void set_finishing() override {
        if (finishing) return;
        finishing = true;
        // packed key width from the declared key column widths (join_hash_map.h:391-406)
        int total = 0;
        for (size_t k = 0; k < bkeys_.size(); k++) {
            if (!bcols_[k].known) { // empty build side: widths come from the probe side later
                total = -1;
                break;
            }
            total += bcols_[k].width;
        }
        if (total > 16) throw Error(SRB_NOT_IMPLEMENTED, "packed join key wider than 16 bytes");
        if (total > 0) {
            width_ = total <= 4 ? 4 : (total <= 8 ? 8 : 16);
            build_table();
        }
        build_done_ = true;
    }
    bool has_output() override { return false; }
    bool need_input() override { return !finishing; }
    bool is_finished() override { return finishing; }
    srb_status pull_chunk(srb_chunk*, bool) override { return SRB_END_OF_FILE; }

    // ---- probe side ----------------------------------------------------------------------------------------------
    bool probe_need_input() override { return build_done_ && !probe_finishing_ && out_pos_ >= out_rows_; }
    bool probe_has_output() override { return out_pos_ < out_rows_; }
    bool probe_is_finished() override { return probe_finishing_ && out_pos_ >= out_rows_; }
    void probe_set_finishing() override {
        if (probe_finishing_) return;
        probe_finishing_ = true;
        if (!track_right_ || !build_done_ || build_rows_ == 0) return;
        if (out_pos_ < out_rows_) throw Error(SRB_INTERNAL_ERROR, "probe set_finishing while output is pending");
        // the build side's remainder (_search_ht_remain): rows that never matched (RIGHT / FULL OUTER, RIGHT ANTI) or did (RIGHT SEMI)
        out_rows_ = out_pos_ = 0;
        for (size_t i = 0; i < pout_.size(); i++)
            if (pkeys_.size() + i >= pcols_.size() || !pcols_[pkeys_.size() + i].known)
                throw Error(SRB_INTERNAL_ERROR, "outer join: no probe chunk described the probe columns (push an empty probe chunk first)");
        SRB_CUDA(cudaMemsetAsync(total_dev_.p, 0, 8, stream));
        const int grid = grid_for(build_rows_);
        k_join_remainder<<<grid, 256, 0, stream>>>(matched_.as<uint8_t>(), (uint32_t)build_rows_, right_want_, total_dev_.as<unsigned long long>(), nullptr);
        SRB_CUDA(cudaMemcpyAsync(total_host_.p, total_dev_.p, 8, cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaMemsetAsync(total_dev_.p, 0, 8, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        const int64_t m = (int64_t)*total_host_.as<unsigned long long>();
        launches++;
        if (m == 0) return;
        obuild_.ensure(ctx, (size_t)m * 4);
        k_join_remainder<<<grid, 256, 0, stream>>>(matched_.as<uint8_t>(), (uint32_t)build_rows_, right_want_, total_dev_.as<unsigned long long>(), obuild_.as<uint32_t>());
        std::vector<const uint8_t*> none(pcols_.size(), nullptr);
        GatherArgs ga = describe_outputs(m, none, none);
        if (ga.n_cols) k_join_gather<<<grid_for(m), 256, 0, stream>>>(ga, m, nullptr, obuild_.as<uint32_t>());
        launches += 2;
        SRB_CUDA(cudaGetLastError());
        SRB_CUDA(cudaStreamSynchronize(stream));
        rows_selected += m;
        out_rows_ = m;
    }

    void probe_push_chunk(const srb_chunk* ch) override {
        if (!build_done_) throw Error(SRB_INTERNAL_ERROR, "probe before the build side finished (need_input() is false)");
        if (probe_finishing_) throw Error(SRB_INTERNAL_ERROR, "probe push after set_finishing");
        if (out_pos_ < out_rows_) throw Error(SRB_INTERNAL_ERROR, "probe push while output is pending (need_input() is false)");
        if (!ch || ch->n_rows < 0) throw Error(SRB_INVALID_ARGUMENT, "bad chunk");
        if (ch->n_rows >= ((int64_t)1 << 32)) throw Error(SRB_INVALID_ARGUMENT, "chunk exceeds 2^32 rows");
        out_rows_ = out_pos_ = 0;
        probe_rows_in_ += ch->n_rows;
        const int64_t n = ch->n_rows;
        // stage probe key + output columns on the device (device columns are used in place)
        pcols_.resize(pkeys_.size() + pout_.size());
        std::vector<const uint8_t*> data(pcols_.size()), nulls(pcols_.size());
        int total = 0;
        for (size_t i = 0; i < pcols_.size(); i++) {
            int slot = i < pkeys_.size() ? pkeys_[i] : pout_[i - pkeys_.size()];
            const srb_column* c = find_or_throw(ch, slot);
            GrowCol& g = pcols_[i];
            describe(g, c, slot);
            if (c->mem_kind == SRB_MEM_DEVICE) {
                data[i] = (const uint8_t*)c->data;
                nulls[i] = c->nulls;
            } else {
                g.data.ensure(ctx, (size_t)n * g.width);
                if (n) SRB_CUDA(cudaMemcpyAsync(g.data.p, c->data, (size_t)n * g.width, cudaMemcpyHostToDevice, stream));
                data[i] = g.data.as<uint8_t>();
                nulls[i] = nullptr;
                if (c->nulls) {
                    g.nulls.ensure(ctx, (size_t)n);
                    if (n) SRB_CUDA(cudaMemcpyAsync(g.nulls.p, c->nulls, (size_t)n, cudaMemcpyHostToDevice, stream));
                    nulls[i] = g.nulls.as<uint8_t>();
                }
            }
            if (i < pkeys_.size()) total += g.width;
        }
        if (n == 0 || (build_rows_ == 0 && !anti_ && !keep_left_)) return;
        if (null_aware_ && build_rows_ > 0 && build_has_null_) return; // x NOT IN (..., NULL, ...) is never true (hash_joiner.h:386-396)
        if (build_rows_ == 0 && anti_) { // LEFT ANTI against nothing: every probe row
            probe_tiled(n, data, nulls);
            return;
        }
        int pw = total <= 4 ? 4 : (total <= 8 ? 8 : 16);
        if (build_rows_ == 0 && width_ == 0) { // LEFT / FULL OUTER against a build side that never described its keys: an empty table
            width_ = pw;
            build_table();
        }
        if (total > 16 || pw != width_) throw Error(SRB_INVALID_ARGUMENT, "probe key width does not match the build key width");
        const bool outer = keep_left_ || track_right_;
        if (((unique_ && !force_general_) || anti_) && !outer && conj_.empty()) { // (conjuncts filter (probe, build) pairs: the general path has them)
            probe_tiled(n, data, nulls);
            return;
        }
        // pack probe keys
        JPackArgs pa{};
        pa.n_keys = (int)pkeys_.size();
        pa.width = width_;
        for (size_t k = 0; k < pkeys_.size(); k++) pa.cols[k] = JKeyCol{data[k], nulls[k], pcols_[k].width, 0};
        ppacked_.ensure(ctx, (size_t)n * width_);
        pnull_.ensure(ctx, (size_t)n);
        int grid = grid_for(n);
        k_join_pack<<<grid, 256, 0, stream>>>(pa, n, ppacked_.as<uint8_t>(), pnull_.as<uint8_t>());
        // pass 1: count
        JTable t = table();
        SRB_CUDA(cudaMemsetAsync(total_dev_.p, 0, 8, stream));
        uint8_t* matched = track_right_ ? matched_.as<uint8_t>() : nullptr;
        k_join_probe<0><<<grid, 256, 0, stream>>>(t, ppacked_.as<uint8_t>(), pnull_.as<uint8_t>(), n, semi_ ? 1 : 0,
                                                   total_dev_.as<unsigned long long>(), nullptr, nullptr, keep_left_ ? 1 : 0, matched, right_only_ ? 0 : 1);
        SRB_CUDA(cudaMemcpyAsync(total_host_.p, total_dev_.p, 8, cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaMemsetAsync(total_dev_.p, 0, 8, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        launches += 2;
        int64_t m = (int64_t)*total_host_.as<unsigned long long>();
        rows_selected += m;
        if (m == 0) return;
        if (m >= ((int64_t)1 << 32)) throw Error(SRB_NOT_IMPLEMENTED, "join output of one probe chunk exceeds 2^32 rows; push smaller chunks");
        // pass 2: append pairs
        oprobe_.ensure(ctx, (size_t)m * 4);
        obuild_.ensure(ctx, (size_t)m * 4);
        k_join_probe<1><<<grid, 256, 0, stream>>>(t, ppacked_.as<uint8_t>(), pnull_.as<uint8_t>(), n, semi_ ? 1 : 0,
                                                   total_dev_.as<unsigned long long>(), oprobe_.as<uint32_t>(), obuild_.as<uint32_t>(), keep_left_ ? 1 : 0);
        launches++;
        if (!conj_.empty()) m = filter_pairs(m, data, nulls);
        if (m == 0) return;
        // K8 gather
        GatherArgs ga = describe_outputs(m, data, nulls);
        if (ga.n_cols) {
            k_join_gather<<<grid_for(m), 256, 0, stream>>>(ga, m, oprobe_.as<uint32_t>(), obuild_.as<uint32_t>());
            launches++;
        }
        SRB_CUDA(cudaGetLastError());
        SRB_CUDA(cudaStreamSynchronize(stream));
        out_rows_ = m;
    }

    static void swap_bufs(DevBuf& a, DevBuf& b) {
        std::swap(a.ctx, b.ctx);
        std::swap(a.p, b.p);
        std::swap(a.bytes, b.bytes);
    }
    // other-join-conjuncts over the m pairs in oprobe_ / obuild_: the surviving pairs replace them; returns their number
    int64_t filter_pairs(int64_t m, const std::vector<const uint8_t*>& data, const std::vector<const uint8_t*>& nulls) {
        ConjArgs ca{};
        ca.n = (int)conj_.size();
        auto resolve = [&](int slot, ConjCol& out, int* type, int* prec, int* scale) {
            for (size_t i = 0; i < pkeys_.size() + pout_.size(); i++) {
                const int s = i < pkeys_.size() ? pkeys_[i] : pout_[i - pkeys_.size()];
                if (s != slot) continue;
                out = ConjCol{data[i], nulls[i], pcols_[i].width, 0};
                *type = pcols_[i].type, *prec = pcols_[i].precision, *scale = pcols_[i].scale;
                return;
            }
            for (size_t i = 0; i < bkeys_.size() + bout_.size(); i++) {
                const int s = i < bkeys_.size() ? bkeys_[i] : bout_[i - bkeys_.size()];
                if (s != slot) continue;
                const GrowCol& g = bcols_[i];
                out = ConjCol{g.data.as<uint8_t>(), g.nullable ? g.nulls.as<uint8_t>() : nullptr, g.width, 1};
                *type = g.type, *prec = g.precision, *scale = g.scale;
                return;
            }
            throw Error(SRB_INVALID_ARGUMENT, fmt("other-join-conjunct: slot %d is neither a key nor an output column of the join", slot));
        };
        for (int c = 0; c < ca.n; c++) {
            int lt, lp, ls, rt, rp, rs;
            resolve(conj_[c].left_slot, ca.l[c], &lt, &lp, &ls);
            resolve(conj_[c].right_slot, ca.r[c], &rt, &rp, &rs);
            if (lt != rt || ls != rs || lt == SRB_TYPE_DOUBLE || is_binary(lt))
                throw Error(SRB_NOT_IMPLEMENTED, "other-join-conjunct: both sides must have the same integer / DATE / DECIMAL type and scale");
            ca.op[c] = conj_[c].op;
        }
        oprobe2_.ensure(ctx, (size_t)m * 4);
        obuild2_.ensure(ctx, (size_t)m * 4);
        SRB_CUDA(cudaMemsetAsync(total_dev_.p, 0, 8, stream));
        k_join_conjuncts<<<grid_for(m), 256, 0, stream>>>(ca, m, oprobe_.as<uint32_t>(), obuild_.as<uint32_t>(), oprobe2_.as<uint32_t>(), obuild2_.as<uint32_t>(),
                                                          total_dev_.as<unsigned long long>());
        SRB_CUDA(cudaMemcpyAsync(total_host_.p, total_dev_.p, 8, cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        launches++;
        swap_bufs(oprobe_, oprobe2_);
        swap_bufs(obuild_, obuild2_);
        const int64_t kept = (int64_t)*total_host_.as<unsigned long long>();
        rows_selected -= m - kept;
        return kept;
    }

    // output columns for m rows: probe_out columns (from the pushed chunk) then build_out columns (from the build side)
    GatherArgs describe_outputs(int64_t m, const std::vector<const uint8_t*>& data, const std::vector<const uint8_t*>& nulls) {
        GatherArgs ga{};
        ocols_.resize(pout_.size() + bout_.size());
        for (size_t i = 0; i < ocols_.size(); i++) {
            bool fb = i >= pout_.size();
            const GrowCol& src = fb ? bcols_[bkeys_.size() + (i - pout_.size())] : pcols_[pkeys_.size() + i];
            if (!src.known) throw Error(SRB_INTERNAL_ERROR, "join output column of a side that never described its columns (push an empty chunk of that side first)");
            GrowCol& o = ocols_[i];
            o.type = src.type;
            o.precision = src.precision;
            o.scale = src.scale;
            o.width = src.width;
            o.known = true;
            const uint8_t* sdata = fb ? src.data.as<uint8_t>() : data[pkeys_.size() + i];
            const uint8_t* snull = fb ? (src.nullable ? src.nulls.as<uint8_t>() : nullptr) : nulls[pkeys_.size() + i];
            // a side that may be missing in an output row comes back nullable: build columns of LEFT / FULL OUTER, probe columns of
            // RIGHT / FULL OUTER
            o.nullable = snull != nullptr || (fb ? keep_left_ : (track_right_ && !right_only_));
            o.data.ensure(ctx, (size_t)m * o.width);
            if (o.nullable) o.nulls.ensure(ctx, (size_t)m);
            ga.cols[i] = GatherCol{sdata, o.data.as<uint8_t>(), snull, o.nullable ? o.nulls.as<uint8_t>() : nullptr, o.width, fb ? 1 : 0};
        }
        ga.n_cols = (int)ocols_.size();
        return ga;
    }

    // at most one match per probe row (unique build keys, or LEFT SEMI): two streaming passes, output in probe row order
    void probe_tiled(int64_t n, const std::vector<const uint8_t*>& data, const std::vector<const uint8_t*>& nulls) {
        const uint8_t* keys = nullptr;
        const uint8_t* knull = nullptr;
        const bool all_rows = anti_ && build_rows_ == 0;
        if (all_rows) {
        } else if (pkeys_.size() == 1 && pcols_[0].width == width_) { // a single 4/8/16-byte key column is its own packed form
            keys = data[0];
            knull = nulls[0];
        } else {
            JPackArgs pa{};
            pa.n_keys = (int)pkeys_.size();
            pa.width = width_;
            for (size_t k = 0; k < pkeys_.size(); k++) pa.cols[k] = JKeyCol{data[k], nulls[k], pcols_[k].width, 0};
            ppacked_.ensure(ctx, (size_t)n * width_);
            pnull_.ensure(ctx, (size_t)n);
            k_join_pack<<<grid_for(n), 256, 0, stream>>>(pa, n, ppacked_.as<uint8_t>(), pnull_.as<uint8_t>());
            launches++;
            keys = ppacked_.as<uint8_t>();
            knull = pnull_.as<uint8_t>();
        }
        const int64_t tiles = (n + JT_TILE - 1) / JT_TILE;
        const int grid = (int)std::min<int64_t>(tiles, (int64_t)ctx->n_sms * 8);
        match_.ensure(ctx, (size_t)n * 4);
        tcounts_.ensure(ctx, (size_t)tiles * 4);
        toffs_.ensure(ctx, (size_t)tiles * 8);
        uint32_t* match = match_.as<uint32_t>();
        uint32_t* tc = tcounts_.as<uint32_t>();
        const int anti = anti_ ? (null_aware_ && build_rows_ > 0 ? 2 : 1) : 0;
        if (all_rows) {
            k_join_all_rows<<<grid, JT_THREADS, 0, stream>>>(n, match, tc);
        } else {
            JTable t = table();
            if (width_ == 4)
                k_join_probe_first<4><<<grid, JT_THREADS, 0, stream>>>(t, keys, knull, n, match, tc, anti);
            else if (width_ == 8)
                k_join_probe_first<8><<<grid, JT_THREADS, 0, stream>>>(t, keys, knull, n, match, tc, anti);
            else
                k_join_probe_first<16><<<grid, JT_THREADS, 0, stream>>>(t, keys, knull, n, match, tc, anti);
        }
        k_join_scan<<<1, 1024, 0, stream>>>(tc, toffs_.as<uint64_t>(), tiles, total_dev_.as<unsigned long long>());
        SRB_CUDA(cudaMemcpyAsync(total_host_.p, total_dev_.p, 8, cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        launches += 2;
        const int64_t m = (int64_t)*total_host_.as<unsigned long long>();
        rows_selected += m;
        if (m == 0) return;
        GatherArgs ga = describe_outputs(m, data, nulls);
        uint32_t *op = nullptr, *ob = nullptr;
        if (emit_indexes_) {
            oprobe_.ensure(ctx, (size_t)m * 4);
            obuild_.ensure(ctx, (size_t)m * 4);
            op = oprobe_.as<uint32_t>();
            ob = obuild_.as<uint32_t>();
        }
        if (ga.n_cols || emit_indexes_) {
            k_join_emit<<<grid, JT_THREADS, 0, stream>>>(ga, n, match, toffs_.as<uint64_t>(), op, ob);
            launches++;
        }
        SRB_CUDA(cudaGetLastError());
        SRB_CUDA(cudaStreamSynchronize(stream));
        out_rows_ = m;
    }

    // partial runtime filter over build key column `key_index`; expected_rows sizes the directory (0: this op's build rows) so that
    // the partial filters of several ranks can be merged (SimdBlockFilter::merge needs equal sizes)
    srb_runtime_filter* build_runtime_filter(int key_index, int64_t expected_rows, int flags) override {
        const int mix = flags & SRB_RF_HASH_MIX;
        if (!build_done_) throw Error(SRB_INTERNAL_ERROR, "runtime filter requested before the build side finished");
        if (key_index < 0 || key_index >= (int)bkeys_.size()) throw Error(SRB_INVALID_ARGUMENT, "bad build key index");
        const GrowCol& kc = bcols_[key_index];
        if (build_rows_ > 0 && (is_binary(kc.type) || kc.type == SRB_TYPE_DOUBLE || kc.width > 8))
            throw Error(SRB_NOT_IMPLEMENTED, "runtime filters cover integer, DATE and DECIMAL32/64 keys");
        auto rf = std::make_unique<srb_runtime_filter>();
        rf->ctx = ctx;
        rf->key_type = kc.type;
        rf->key_width = kc.width;
        rf->mix = mix ? 1 : 0;
        rf->log_num_buckets = rf_log_num_buckets((uint64_t)(expected_rows > 0 ? expected_rows : build_rows_));
        rf->directory_mask = (uint32_t)(((uint64_t)1 << rf->log_num_buckets) - 1);
        rf->dir.ensure(ctx, rf->dir_bytes());
        SRB_CUDA(cudaMemsetAsync(rf->dir.p, 0, rf->dir_bytes(), stream));
        DevBuf meta;
        PinnedBuf hmeta;
        meta.ensure(ctx, sizeof(RfMeta));
        hmeta.ensure(sizeof(RfMeta));
        RfMeta* hm = hmeta.as<RfMeta>();
        *hm = RfMeta{INT64_MAX, INT64_MIN, 0, 0, 0};
        SRB_CUDA(cudaMemcpyAsync(meta.p, hm, sizeof(RfMeta), cudaMemcpyHostToDevice, stream));
        if (build_rows_ > 0) {
            const uint8_t* keys = kc.data.as<uint8_t>();
            const uint8_t* nulls = kc.nullable ? kc.nulls.as<uint8_t>() : nullptr;
            const int grid = grid_for(build_rows_);
            const RfView v = rf->view();
            switch (kc.width) {
            case 1: k_rf_build<1><<<grid, 256, 0, stream>>>(keys, nulls, build_rows_, v, meta.as<RfMeta>()); break;
            case 2: k_rf_build<2><<<grid, 256, 0, stream>>>(keys, nulls, build_rows_, v, meta.as<RfMeta>()); break;
            case 4: k_rf_build<4><<<grid, 256, 0, stream>>>(keys, nulls, build_rows_, v, meta.as<RfMeta>()); break;
            default: k_rf_build<8><<<grid, 256, 0, stream>>>(keys, nulls, build_rows_, v, meta.as<RfMeta>());
            }
            SRB_CUDA(cudaGetLastError());
            launches++;
        }
        SRB_CUDA(cudaMemcpyAsync(hm, meta.p, sizeof(RfMeta), cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        rf->min = hm->min;
        rf->max = hm->max;
        rf->inserted = (int64_t)hm->inserted;
        rf->has_null = hm->has_null != 0;
        // RuntimeFilterHelper::create_join_runtime_filter (runtime_filter_bank.cpp:160-205): a builder that holds the WHOLE build side (a
        // broadcast join: the caller says so) takes a bitset over [min, max] instead of the bloom filter when it is no larger than the bloom
        // filter's budget (one byte per build row), or at most 4x that and cache resident (the reference asks the CPU's L2 / L3; here: the
        // GPU's L2). The bloom directory stays (srb_runtime_filter_directory still reads it); the probe tests the bitset.
        if ((flags & SRB_RF_MAYBE_BITSET) && expected_rows <= 0 && rf->inserted > 0 && rf->min <= rf->max) {
            const uint64_t interval = (uint64_t)rf->max - (uint64_t)rf->min + 1;
            const uint64_t bitset_bytes = (interval + 7) / 8, bloom_bytes = (uint64_t)build_rows_;
            if (interval != 0 && (bitset_bytes <= bloom_bytes || (bitset_bytes <= bloom_bytes * 4 && bitset_bytes <= (uint64_t)ctx->l2_bytes))) {
                rf->is_bitset = true;
                rf->bits.ensure(ctx, rf->bits_bytes());
                SRB_CUDA(cudaMemsetAsync(rf->bits.p, 0, rf->bits_bytes(), stream));
                const uint8_t* keys = kc.data.as<uint8_t>();
                const uint8_t* nulls = kc.nullable ? kc.nulls.as<uint8_t>() : nullptr;
                const int grid = grid_for(build_rows_);
                switch (kc.width) {
                case 1: k_rf_bits_build<1><<<grid, 256, 0, stream>>>(keys, nulls, build_rows_, rf->min, rf->bits.as<uint32_t>()); break;
                case 2: k_rf_bits_build<2><<<grid, 256, 0, stream>>>(keys, nulls, build_rows_, rf->min, rf->bits.as<uint32_t>()); break;
                case 4: k_rf_bits_build<4><<<grid, 256, 0, stream>>>(keys, nulls, build_rows_, rf->min, rf->bits.as<uint32_t>()); break;
                default: k_rf_bits_build<8><<<grid, 256, 0, stream>>>(keys, nulls, build_rows_, rf->min, rf->bits.as<uint32_t>());
                }
                SRB_CUDA(cudaGetLastError());
                SRB_CUDA(cudaStreamSynchronize(stream));
                launches++;
            }
        }
        return rf.release();
    }

    srb_status probe_pull_chunk(srb_chunk* out, bool all) override { return pull(out, all, false); }
    srb_status probe_pull_device(srb_chunk* out) override { return pull(out, true, true); }

    void refresh_counters() override {}

private:
    srb_status pull(srb_chunk* out, bool all, bool device) {
        if (out_pos_ >= out_rows_) {
            // a harness pulling "all" gets one empty, typed chunk when nothing was ever produced; otherwise done
            if (!(all && !gave_empty_)) return SRB_END_OF_FILE;
        }
        int64_t n = out_rows_ - out_pos_;
        if (!all && n > chunk_size_) n = chunk_size_;
        gave_empty_ = true;
        hslots_.clear();
        cview_.clear();
        size_t nc = pout_.size() + bout_.size() + (emit_indexes_ ? 2 : 0);
        hdata_.resize(nc);
        hnull_.resize(nc);
        for (size_t i = 0; i < nc; i++) {
            srb_column c{};
            const uint8_t* ddata;
            const uint8_t* dnull = nullptr;
            int w;
            if (i < pout_.size() + bout_.size()) {
                if (i >= ocols_.size() || !ocols_[i].known) { // nothing was ever produced: describe from the inputs
                    const GrowCol* src = i < pout_.size() ? (pkeys_.size() + i < pcols_.size() ? &pcols_[pkeys_.size() + i] : nullptr)
                                                          : &bcols_[bkeys_.size() + (i - pout_.size())];
                    c.type = src && src->known ? src->type : SRB_TYPE_BIGINT; // no chunk ever described it (empty side)
                    c.precision = src ? src->precision : 0;
                    c.scale = src ? src->scale : 0;
                    w = type_width(c.type);
                    ddata = nullptr;
                } else {
                    const GrowCol& o = ocols_[i];
                    c.type = o.type;
                    c.precision = o.precision;
                    c.scale = o.scale;
                    w = o.width;
                    ddata = o.data.as<uint8_t>();
                    dnull = o.nullable ? o.nulls.as<uint8_t>() : nullptr;
                }
                hslots_.push_back(i < pout_.size() ? pout_[i] : bout_[i - pout_.size()]);
            } else {
                bool is_probe = i == pout_.size() + bout_.size();
                c.type = SRB_TYPE_INT;
                w = 4;
                ddata = n ? (is_probe ? oprobe_.as<uint8_t>() : obuild_.as<uint8_t>()) : nullptr;
                hslots_.push_back(is_probe ? -4 : -3); // HASH_JOIN_PROBE_INDEX_SLOT_ID / HASH_JOIN_BUILD_INDEX_SLOT_ID
            }
            if (device) {
                c.mem_kind = SRB_MEM_DEVICE;
                c.data = ddata ? ddata + (size_t)out_pos_ * w : nullptr;
                c.nulls = dnull ? dnull + out_pos_ : nullptr;
            } else {
                c.mem_kind = SRB_MEM_HOST;
                hdata_[i].resize((size_t)n * w + 16);
                if (n && ddata) SRB_CUDA(cudaMemcpyAsync(hdata_[i].data(), ddata + (size_t)out_pos_ * w, (size_t)n * w, cudaMemcpyDeviceToHost, stream));
                c.data = hdata_[i].data();
                if (dnull) {
                    hnull_[i].resize((size_t)n + 16);
                    if (n) SRB_CUDA(cudaMemcpyAsync(hnull_[i].data(), dnull + out_pos_, (size_t)n, cudaMemcpyDeviceToHost, stream));
                    c.nulls = hnull_[i].data();
                }
            }
            cview_.push_back(c);
        }
        SRB_CUDA(cudaStreamSynchronize(stream));
        out_pos_ += n;
        out->n_rows = n;
        out->n_cols = (int32_t)cview_.size();
        out->slot_ids = hslots_.data();
        out->cols = cview_.data();
        return SRB_OK;
    }

    static const srb_column* find_or_throw(const srb_chunk* ch, int slot) {
        for (int i = 0; i < ch->n_cols; i++)
            if (ch->slot_ids[i] == slot) return &ch->cols[i];
        throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d not in chunk", slot));
    }
    void describe(GrowCol& g, const srb_column* c, int slot) {
        int w = type_width(c->type);
        if (w == 0) throw Error(SRB_NOT_IMPLEMENTED, fmt("slot %d: binary columns are not supported by the join yet", slot));
        if (g.known && g.type != c->type) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d changed type", slot));
        g.type = c->type;
        g.precision = c->precision;
        g.scale = c->scale;
        g.width = w;
        g.known = true;
    }
    void append_col(GrowCol& g, const srb_column* c, int slot, int64_t have, int64_t n) {
        describe(g, c, slot);
        size_t need = (size_t)(have + n) * g.width;
        if (need > g.data.bytes) regrow(g.data, (size_t)have * g.width, std::max(need, g.data.bytes * 2));
        SRB_CUDA(cudaMemcpyAsync(g.data.as<uint8_t>() + (size_t)have * g.width, c->data, (size_t)n * g.width,
                                 c->mem_kind == SRB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, stream));
        if (c->nulls && !g.nullable) { // first null map: earlier rows were all non-null
            g.nullable = true;
            regrow(g.nulls, 0, (size_t)std::max<int64_t>(have + n, 1024));
            SRB_CUDA(cudaMemsetAsync(g.nulls.p, 0, (size_t)have, stream));
        }
        if (g.nullable) {
            if ((size_t)(have + n) > g.nulls.bytes) regrow(g.nulls, (size_t)have, std::max((size_t)(have + n), g.nulls.bytes * 2));
            if (c->nulls)
                SRB_CUDA(cudaMemcpyAsync(g.nulls.as<uint8_t>() + have, c->nulls, (size_t)n,
                                         c->mem_kind == SRB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, stream));
            else
                SRB_CUDA(cudaMemsetAsync(g.nulls.as<uint8_t>() + have, 0, (size_t)n, stream));
        }
        if (c->mem_kind != SRB_MEM_DEVICE) SRB_CUDA(cudaStreamSynchronize(stream)); // pageable source may be reused by the caller
    }
    void regrow(DevBuf& b, size_t used, size_t ncap) {
        DevBuf nb;
        nb.ensure(ctx, ncap);
        if (used) SRB_CUDA(cudaMemcpyAsync(nb.p, b.p, used, cudaMemcpyDeviceToDevice, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        std::swap(b.p, nb.p);
        std::swap(b.bytes, nb.bytes);
        nb.ctx = ctx;
        b.ctx = ctx;
    }
    int grid_for(int64_t n) const { return (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)ctx->n_sms * 8)); }

    void build_table() {
        const int64_t n = build_rows_;
        JPackArgs pa{};
        pa.n_keys = (int)bkeys_.size();
        pa.width = width_;
        for (size_t k = 0; k < bkeys_.size(); k++)
            pa.cols[k] = JKeyCol{bcols_[k].data.as<uint8_t>(), bcols_[k].nullable ? bcols_[k].nulls.as<uint8_t>() : nullptr, bcols_[k].width, 0};
        bpacked_.ensure(ctx, (size_t)std::max<int64_t>(n, 1) * width_);
        bnull_.ensure(ctx, (size_t)std::max<int64_t>(n, 1));
        bucket_size_ = calc_bucket_size((uint32_t)n + 1);
        log_bucket_ = (uint32_t)__builtin_ctz(bucket_size_);
        first_.ensure(ctx, (size_t)bucket_size_ * 4);
        next_.ensure(ctx, (size_t)(n + 1) * 4);
        if (track_right_) {
            matched_.ensure(ctx, (size_t)(n + 1));
            SRB_CUDA(cudaMemsetAsync(matched_.p, 0, (size_t)(n + 1), stream));
        }
        SRB_CUDA(cudaMemsetAsync(first_.p, 0, (size_t)bucket_size_ * 4, stream));
        SRB_CUDA(cudaMemsetAsync(next_.p, 0, (size_t)(n + 1) * 4, stream));
        if (n > 0) {
            k_join_pack<<<grid_for(n), 256, 0, stream>>>(pa, n, bpacked_.as<uint8_t>(), bnull_.as<uint8_t>());
            k_join_build<<<grid_for(n), 256, 0, stream>>>(table());
            launches += 2;
            if (null_aware_) {
                SRB_CUDA(cudaMemsetAsync(total_dev_.p, 0, 8, stream));
                k_join_any_null<<<grid_for(n), 256, 0, stream>>>(bnull_.as<uint8_t>(), n, total_dev_.as<unsigned int>());
                SRB_CUDA(cudaMemcpyAsync(total_host_.p, total_dev_.p, 8, cudaMemcpyDeviceToHost, stream));
                SRB_CUDA(cudaStreamSynchronize(stream));
                build_has_null_ = *total_host_.as<unsigned int>() != 0;
                launches++;
            }
            if (!semi_) { // does any key occur twice? (decides whether INNER probes may take the one-match-per-row path)
                SRB_CUDA(cudaMemsetAsync(total_dev_.p, 0, 8, stream));
                k_join_unique<<<grid_for(n), 256, 0, stream>>>(table(), total_dev_.as<unsigned int>());
                SRB_CUDA(cudaMemcpyAsync(total_host_.p, total_dev_.p, 8, cudaMemcpyDeviceToHost, stream));
                launches++;
            }
        }
        SRB_CUDA(cudaGetLastError());
        SRB_CUDA(cudaStreamSynchronize(stream));
        unique_ = semi_ || n == 0 || *total_host_.as<unsigned int>() == 0;
    }
    JTable table() {
        JTable t;
        t.keys = bpacked_.as<uint8_t>();
        t.key_null = bnull_.as<uint8_t>();
        t.first = first_.as<uint32_t>();
        t.next = next_.as<uint32_t>();
        t.bucket_size = bucket_size_;
        t.log_bucket = log_bucket_;
        t.rows = (uint32_t)build_rows_;
        t.width = width_;
        return t;
    }

    bool null_aware_ = false, build_has_null_ = false;
    bool anti_ = false, keep_left_ = false, track_right_ = false, right_only_ = false;
    int right_want_ = 0;
    DevBuf matched_; // one byte per build row (1-based): the row matched some probe row (RIGHT / FULL OUTER, RIGHT SEMI / ANTI)
    bool semi_ = false, emit_indexes_ = false, build_done_ = false, probe_finishing_ = false, gave_empty_ = false;
    std::vector<int32_t> bkeys_, pkeys_, bout_, pout_;
    std::vector<srb_join_conjunct> conj_; // other-join-conjuncts (INNER joins)
    DevBuf oprobe2_, obuild2_;
    int64_t chunk_size_;
    std::vector<GrowCol> bcols_, pcols_, ocols_;
    int64_t build_rows_ = 0, probe_rows_in_ = 0;
    int width_ = 0;
    uint32_t bucket_size_ = 0, log_bucket_ = 0;
    DevBuf bpacked_, bnull_, first_, next_, ppacked_, pnull_, oprobe_, obuild_, total_dev_, match_, tcounts_, toffs_;
    bool unique_ = false;                                               // no key occurs twice on the build side (or LEFT SEMI)
    bool force_general_ = getenv("SRB_JOIN_GENERAL") != nullptr;        // test hook: the chain-walking two-pass probe for every join
    PinnedBuf total_host_;
    int64_t out_rows_ = 0, out_pos_ = 0;
    std::vector<std::vector<uint8_t>> hdata_, hnull_;
    std::vector<int32_t> hslots_;
    std::vector<srb_column> cview_;
};

srb_op* make_join_op(srb_ctx* ctx, const srb_join_plan* plan) { return new JoinOp(ctx, plan); }

} // namespace srb
