// srb_scan_agg.cu — the fused scan-filter -> project -> aggregate operator (one driver's operator instance).
//
// Stands in for, and keeps the call contract of:
//   OlapChunkSource::_read_chunk_from_storage predicate part    be/src/exec/pipeline/scan/olap_chunk_source.cpp:586-624
//   Chunk::filter / ColumnHelper::t_filter_range                be/src/column/chunk.cpp:302, column/column_helper.h:468-577
//   AggregateBlockingSinkOperator::push_chunk / set_finishing   be/src/exec/pipeline/aggregate/aggregate_blocking_sink_operator.cpp:51-134
//   AggregateBlockingSourceOperator::pull_chunk                 be/src/exec/pipeline/aggregate/aggregate_blocking_source_operator.cpp:47-72
//   Aggregator::build_hash_map / compute_*_agg_states / convert_hash_map_to_chunk   be/src/exec/aggregator.cpp:833-959,1357-1514
//
// Design (B200-first): whole column ranges per launch instead of 4096-row chunks; predicates, expressions, key packing,
// table probe and state update fused in one kernel so no intermediate column ever touches HBM; the group table is an
// open-addressing table in HBM updated with 64-bit atomics (int128 sums as lo/hi words with carry); recognised program
// shapes run on statically specialised kernels with TMA-staged column tiles and register pre-aggregation
// (srb_shapes.cuh). There is no CPU path: every row is processed by a CUDA kernel.
#include <limits>
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <type_traits>

#include "srb_host.h"
#include "srb_lower.h"
#include "srb_rowpipe.cuh"
#include "srb_shapes.cuh"
#include "srb_jit.h"
#include "srb_radix_agg.cuh"
#include "srb_pages.cuh"
#include "srb_stream_policy.h"
#include "srb_hcpart.cuh"

namespace srb {

struct Counters {
    unsigned long long rows_selected;
    unsigned long long deferred;
    unsigned long long key_too_long;
    unsigned long long emitted;
    unsigned long long inserted;
    unsigned long long any_null;
};

// NullableColumn::has_null() (column/nullable_column.h:95-99): the reference keeps a flag beside every null map and takes the
// not-null branch of its aggregate / expression loops when it is clear. The C ABI passes plain pointers, so the flag is recomputed
// here: one pass over the 1 byte/row maps of a batch.
__global__ void __launch_bounds__(256) k_null_any(const uint8_t* __restrict__ nulls, int64_t n, unsigned long long* __restrict__ flag) {
    const int64_t lead = (int64_t)((16 - ((uintptr_t)nulls & 15)) & 15), head = lead < n ? lead : n;
    const int64_t quads = (n - head) / 16;
    const uint4* body = reinterpret_cast<const uint4*>(nulls + head);
    unsigned int any = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int64_t i = t0; i < quads; i += stride) {
        const uint4 q = body[i];
        any |= q.x | q.y | q.z | q.w;
    }
    for (int64_t i = t0; i < head; i += stride) any |= nulls[i];
    for (int64_t i = head + quads * 16 + t0; i < n; i += stride) any |= nulls[i];
    if (__any_sync(0xffffffffu, any != 0) && (threadIdx.x & 31) == 0) atomicOr(flag, 1ULL);
}

// ---- generic kernel: interprets the lowered program, one row per thread iteration ---------------------------
struct LocalAcc {
    uint64_t lo, hi, n;
};

SRB_D void update_agg_local(const DProg& p, const DAgg& g, LocalAcc& s, const Val* v, int64_t row) {
    if (g.fn == FN_COUNT) {
        if (g.merge) {
            if (!v[g.arg].null) s.n += v[g.arg].lo;
        } else if (g.arg < 0 || !v[g.arg].null)
            s.n++;
        return;
    }
    if (g.merge && g.fn == FN_AVG) {
        const ColBind& c = p.cols[g.blob_col];
        if (c.nulls && c.nulls[row]) return;
        uint32_t b = c.offsets[row], e = c.offsets[row + 1];
        const uint8_t* ptr = (const uint8_t*)c.data + b;
        if (e - b == 24) {
            u128 sum = (u128)make_i128(s.lo, (int64_t)s.hi) + (u128)make_i128(load_u64_le(ptr), (int64_t)load_u64_le(ptr + 8));
            s.lo = (uint64_t)sum;
            s.hi = (uint64_t)(sum >> 64);
            s.n += load_u64_le(ptr + 16);
        } else if (e - b == 16) {
            double d = __longlong_as_double((long long)s.lo) + __longlong_as_double((long long)load_u64_le(ptr));
            s.lo = (uint64_t)__double_as_longlong(d);
            s.n += load_u64_le(ptr + 8);
        }
        return;
    }
    const Val& x = v[g.arg];
    if (x.null) return;
    if (g.fn == FN_SUM || g.fn == FN_AVG) {
        if (g.is_double || g.avg_as_double) {
            double add = g.is_double ? __longlong_as_double((long long)x.lo) : (double)(int64_t)x.lo;
            s.lo = (uint64_t)__double_as_longlong(__longlong_as_double((long long)s.lo) + add);
        } else {
            u128 sum = (u128)make_i128(s.lo, (int64_t)s.hi) + (u128)make_i128(x.lo, x.hi);
            s.lo = (uint64_t)sum;
            s.hi = (uint64_t)(sum >> 64);
        }
    } else {
        const bool mn = g.fn == FN_MIN;
        if (g.is_double) {
            double c = __longlong_as_double((long long)s.lo), d = __longlong_as_double((long long)x.lo);
            if (s.n == 0 || (mn ? d < c : d > c)) s.lo = x.lo;
        } else {
            i128 c = make_i128(s.lo, (int64_t)s.hi), d = make_i128(x.lo, x.hi);
            if (s.n == 0 || (mn ? d < c : d > c)) {
                s.lo = x.lo;
                s.hi = (uint64_t)x.hi;
            }
        }
    }
    s.n++;
}

// merges one thread's private accumulator into the (single) global state block
SRB_D void flush_local(const DAgg& g, const LocalAcc& s, uint8_t* slot) {
    if (s.n == 0) return;
    if (g.fn == FN_COUNT) {
        atomicAdd((unsigned long long*)(slot + g.off_n), (unsigned long long)s.n);
        return;
    }
    if (g.fn == FN_SUM || g.fn == FN_AVG) {
        if (g.is_double || g.avg_as_double)
            atomicAdd((double*)(slot + g.off_lo), __longlong_as_double((long long)s.lo));
        else
            atomic_add128(slot, g.off_lo, g.off_hi, s.lo, s.hi);
    } else {
        const bool mn = g.fn == FN_MIN;
        if (g.is_double)
            atomicMax((unsigned long long*)(slot + g.off_lo), (unsigned long long)enc_f64(s.lo, mn));
        else if (g.wide)
            atomic_umax128(slot, g.off_lo, enc_i128(make_i128(s.lo, (int64_t)s.hi), mn));
        else
            atomicMax((unsigned long long*)(slot + g.off_lo), (unsigned long long)enc_i64((int64_t)s.lo, mn));
    }
    if (g.off_n >= 0) atomicAdd((unsigned long long*)(slot + g.off_n), (unsigned long long)s.n);
}

// Low-cardinality GROUP BY through the interpreter: a few dozen groups take every row's atomics, which then serialise in L2 on the
// same few addresses (9.6 M rows into 35 groups: 1.25 ms). When the table is small (sh_cap slots x aggregates <= GEN_SH_ENTRIES: 48 KB) and
// every aggregate is an integer SUM / AVG / COUNT, a CTA adds its rows into a shared-memory mirror of the table's states instead and
// flushes each touched state with one global atomic when it retires. Integer addition commutes: the result is bit-identical.
constexpr int GEN_SH_ENTRIES = 2048;

// x added to a 64-bit shared-memory word with two NATIVE 32-bit atomics; returns the carry out of the word. A 64-bit atomicAdd on shared
// memory is a compare-and-swap loop (SASS ATOMS.CAST.SPIN: with 32 lanes on a few dozen addresses every round lets one lane per address
// through — profiles/r2_interpreter_nullable_full.txt), a 32-bit one is a single ATOMS.ADD that the unit serialises itself. Every carry is
// handed on exactly once, so the multi-word sum is exact whatever the interleaving.
SRB_D unsigned int smem_add64(unsigned long long* w, unsigned long long x) {
    unsigned int* p = (unsigned int*)w;
    const unsigned int xl = (unsigned int)x, xh = (unsigned int)(x >> 32);
    unsigned int c = 0;
    if (xl) {
        const unsigned int o = atomicAdd(p, xl);
        c = (o + xl) < o ? 1u : 0u;
    }
    const unsigned int add = xh + c;
    unsigned int cout = (c && add == 0) ? 1u : 0u; // 0xFFFFFFFF + the carry: nothing to add here, one to hand on
    if (add) {
        const unsigned int o = atomicAdd(p + 1, add);
        cout |= (o + add) < o ? 1u : 0u;
    }
    return cout;
}

__global__ void __launch_bounds__(256) k_scan_agg_generic(const __grid_constant__ DProg prog, const TableView tv, const int64_t base,
                                                           const int64_t n, const int64_t bit_base, const uint32_t* __restrict__ defer_in,
                                                           uint32_t* __restrict__ defer_out, Counters* __restrict__ ctr, const int sh_cap) {
    const DProg& p = prog;
    Val v[MAX_NODES];
    LocalAcc acc[MAX_AGGS];
    const bool no_keys = p.n_keys == 0;
    extern __shared__ unsigned long long sh_acc[]; // {lo, hi, n} per (slot, aggregate); sh_cap * n_aggs * 24 bytes, none when sh_cap == 0
    if (sh_cap) {
        for (int i = threadIdx.x; i < sh_cap * p.n_aggs * 3; i += blockDim.x) sh_acc[i] = 0;
        __syncthreads();
    }
    if (no_keys)
        for (int a = 0; a < p.n_aggs; a++) acc[a] = LocalAcc{0, 0, 0};
    unsigned long long selected = 0, deferred = 0, too_long = 0, inserted = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int64_t bit = i + bit_base; // position of this row in the batch's deferral bitmaps
        if (defer_in && !((defer_in[bit >> 5] >> (bit & 31)) & 1u)) continue;
        const int64_t row = base + i;
        if (!eval_preds(p, row)) continue;
        if (!defer_in) selected++;
        for (int k = 0; k < p.n_nodes; k++) v[k] = eval_node(p.nodes[k], v, p.cols, row);
        if (no_keys) {
            for (int a = 0; a < p.n_aggs; a++) update_agg_local(p, p.aggs[a], acc[a], v, row);
            continue;
        }
        bool fresh;
        uint8_t* slot;
        if (p.key_words > 2) { // a wide key (17-64 bytes)
            const WideKey wk = pack_key_wide(p, v, row);
            if (wk.too_long) {
                too_long++;
                continue;
            }
            slot = table_find_or_insert_wide(tv, wk, hash_wide_key(wk, p.phase_seed), &fresh);
        } else {
            const PackedKey k = pack_key(p, v, row);
            if (k.too_long) {
                too_long++;
                continue;
            }
            slot = table_find_or_insert(tv, k.lo, k.hi, hash_packed_key(k, p.phase_seed), &fresh);
        }
        inserted += fresh ? 1 : 0;
        if (!slot) {
            atomicOr(&defer_out[bit >> 5], 1u << (bit & 31));
            deferred++;
            continue;
        }
        if (sh_cap) {
            const uint32_t idx = (uint32_t)(slot - tv.slots) / (uint32_t)p.slot_bytes;
            for (int a = 0; a < p.n_aggs; a++) {
                const DAgg& g = p.aggs[a];
                LocalAcc d{0, 0, 0};
                update_agg_local(p, g, d, v, row);
                if (d.n == 0) continue; // a NULL argument: nothing to add
                unsigned long long* e = sh_acc + (idx * p.n_aggs + a) * 3;
                if (g.fn == FN_MIN || g.fn == FN_MAX) { // the mirror keeps the table's own order-preserving encoding (identity: zero)
                    const bool mn = g.fn == FN_MIN;
                    const unsigned long long enc = (unsigned long long)(g.is_double ? enc_f64(d.lo, mn) : enc_i64((int64_t)d.lo, mn));
                    if (enc > *(volatile unsigned long long*)e) atomicMax(e, enc); // (the word only grows: a value that does not beat it needs no atomic)
                } else if (g.fn != FN_COUNT && (g.is_double || g.avg_as_double)) {
                    atomicAdd((double*)e, __longlong_as_double((long long)d.lo));
                } else if (d.lo | d.hi) { // 128-bit add word by word: the carry out of the low word travels with the high word's addend
                    const unsigned long long up = d.hi + smem_add64(e, (unsigned long long)d.lo);
                    if (up) smem_add64(e + 1, up);
                }
                smem_add64(e + 2, (unsigned long long)d.n);
            }
            continue;
        }
        for (int a = 0; a < p.n_aggs; a++) update_agg_global(p, p.aggs[a], slot, v, row);
    }
    if (no_keys)
        for (int a = 0; a < p.n_aggs; a++) flush_local(p.aggs[a], acc[a], tv.slots);
    if (sh_cap) {
        __syncthreads();
        for (int i = threadIdx.x; i < sh_cap * p.n_aggs; i += blockDim.x) {
            const LocalAcc s{sh_acc[i * 3], sh_acc[i * 3 + 1], sh_acc[i * 3 + 2]};
            const DAgg& g = p.aggs[i % p.n_aggs];
            uint8_t* slot = tv.slots + (size_t)(i / p.n_aggs) * p.slot_bytes;
            if (s.n && (g.fn == FN_MIN || g.fn == FN_MAX)) { // already encoded: fold it in as it is
                atomicMax((unsigned long long*)(slot + g.off_lo), (unsigned long long)s.lo);
                if (g.off_n >= 0) atomicAdd((unsigned long long*)(slot + g.off_n), (unsigned long long)s.n);
            } else {
                flush_local(g, s, slot); // a state no row touched has n == 0
            }
        }
    }
    // one atomic per warp for the counters
    for (int o = 16; o > 0; o >>= 1) {
        selected += __shfl_down_sync(0xffffffffu, selected, o);
        deferred += __shfl_down_sync(0xffffffffu, deferred, o);
        too_long += __shfl_down_sync(0xffffffffu, too_long, o);
        inserted += __shfl_down_sync(0xffffffffu, inserted, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (selected) atomicAdd(&ctr->rows_selected, selected);
        if (deferred) atomicAdd(&ctr->deferred, deferred);
        if (too_long) atomicAdd(&ctr->key_too_long, too_long);
        if (inserted) atomicAdd(&ctr->inserted, inserted);
    }
}

// ---- low-cardinality GROUP BY through the interpreter: thread-private accumulators ------------------------------------------------
// A handful of groups (TPC-H Q1's four) take every row: per-row atomics — on the table or on a shared mirror — serialise on the same few
// addresses (134 M rows of Q1 with VARCHAR(1) keys: 48 ms through the shared mirror; with one MIN / MAX, which the mirror does not
// cover, 1.06 s of global atomicMax). Here a CTA keeps a small directory of the table slots its rows have hit (<= LC_MAXG) and every
// THREAD its own accumulator {lo, hi, n} per (directory entry, aggregate) in shared memory, laid out [entry][aggregate][field][thread]:
// an update is three conflict-free LDS, the same update_agg_local the no-GROUP-BY path uses (every function, DOUBLE and MIN / MAX
// included), three STS — no atomics. At retirement the threads' accumulators are merged with warp shuffles and one lane per (entry,
// aggregate) folds the result into the table. Rows of a group beyond the directory update the table directly. Integer results are
// bit-identical; DOUBLE sums are added in a different order (1e-9 relative, as everywhere).
constexpr int LC_THREADS = 128, LC_MAXG = 8;
constexpr uint32_t LC_EMPTY = 0xFFFFFFFFu;

SRB_D void merge_local(const DAgg& g, LocalAcc& into, const LocalAcc& from) {
    if (from.n == 0) return;
    if (g.fn == FN_SUM || g.fn == FN_AVG) {
        if (g.is_double || g.avg_as_double) {
            into.lo = (uint64_t)__double_as_longlong(__longlong_as_double((long long)into.lo) + __longlong_as_double((long long)from.lo));
        } else {
            const u128 sum = (u128)make_i128(into.lo, (int64_t)into.hi) + (u128)make_i128(from.lo, (int64_t)from.hi);
            into.lo = (uint64_t)sum;
            into.hi = (uint64_t)(sum >> 64);
        }
    } else if (g.fn == FN_MIN || g.fn == FN_MAX) {
        const bool mn = g.fn == FN_MIN;
        bool better;
        if (g.is_double) {
            const double c = __longlong_as_double((long long)into.lo), d = __longlong_as_double((long long)from.lo);
            better = mn ? d < c : d > c;
        } else {
            const i128 c = make_i128(into.lo, (int64_t)into.hi), d = make_i128(from.lo, (int64_t)from.hi);
            better = mn ? d < c : d > c;
        }
        if (into.n == 0 || better) {
            into.lo = from.lo;
            into.hi = from.hi;
        }
    }
    into.n += from.n;
}

__global__ void __launch_bounds__(LC_THREADS) k_scan_agg_lowcard(const __grid_constant__ DProg prog, const TableView tv, const int64_t base, const int64_t n,
                                                                  const int64_t bit_base, uint32_t* __restrict__ defer_out, Counters* __restrict__ ctr,
                                                                  const int G) {
    const DProg& p = prog;
    extern __shared__ unsigned long long lc_acc[]; // [entry][aggregate][field][thread]
    __shared__ uint32_t dir[LC_MAXG];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < G * p.n_aggs * 3 * LC_THREADS; i += LC_THREADS) lc_acc[i] = 0;
    if (tid < LC_MAXG) dir[tid] = LC_EMPTY;
    __syncthreads();
    Val v[MAX_NODES];
    unsigned long long selected = 0, deferred = 0, too_long = 0, inserted = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + tid; i < n; i += stride) {
        const int64_t row = base + i;
        if (!eval_preds(p, row)) continue;
        selected++;
        for (int k = 0; k < p.n_nodes; k++) v[k] = eval_node(p.nodes[k], v, p.cols, row);
        PackedKey k = pack_key(p, v, row);
        if (k.too_long) {
            too_long++;
            continue;
        }
        bool fresh;
        uint8_t* slot = table_find_or_insert(tv, k.lo, k.hi, hash_packed_key(k, p.phase_seed), &fresh);
        inserted += fresh ? 1 : 0;
        if (!slot) {
            atomicOr(&defer_out[(i + bit_base) >> 5], 1u << ((i + bit_base) & 31));
            deferred++;
            continue;
        }
        const uint32_t idx = (uint32_t)((slot - tv.slots) / (uint32_t)p.slot_bytes);
        int e = -1;
        for (int j = 0; j < G && e < 0; j++) {
            uint32_t cur = ((volatile uint32_t*)dir)[j];
            if (cur == LC_EMPTY) cur = atomicCAS(&dir[j], LC_EMPTY, idx), cur = cur == LC_EMPTY ? idx : cur;
            if (cur == idx) e = j;
        }
        if (e < 0) { // more groups than directory entries: this row goes to the table directly
            for (int a = 0; a < p.n_aggs; a++) update_agg_global(p, p.aggs[a], slot, v, row);
            continue;
        }
        for (int a = 0; a < p.n_aggs; a++) {
            unsigned long long* w = lc_acc + (size_t)((e * p.n_aggs + a) * 3) * LC_THREADS + tid;
            LocalAcc d{w[0], w[LC_THREADS], w[2 * LC_THREADS]};
            update_agg_local(p, p.aggs[a], d, v, row);
            w[0] = d.lo;
            w[LC_THREADS] = d.hi;
            w[2 * LC_THREADS] = d.n;
        }
    }
    __syncthreads();
    // merge the threads' accumulators: warp w takes the (entry, aggregate) pairs w, w + 4, ...
    for (int pair = warp; pair < G * p.n_aggs; pair += LC_THREADS / 32) {
        const int e = pair / p.n_aggs, a = pair % p.n_aggs;
        const uint32_t idx = dir[e];
        if (idx == LC_EMPTY) continue;
        const DAgg& g = p.aggs[a];
        LocalAcc s{0, 0, 0};
        for (int t = lane; t < LC_THREADS; t += 32) {
            const unsigned long long* w = lc_acc + (size_t)(pair * 3) * LC_THREADS + t;
            merge_local(g, s, LocalAcc{w[0], w[LC_THREADS], w[2 * LC_THREADS]});
        }
        for (int o = 16; o > 0; o >>= 1) {
            LocalAcc y;
            y.lo = __shfl_down_sync(0xffffffffu, s.lo, o);
            y.hi = __shfl_down_sync(0xffffffffu, s.hi, o);
            y.n = __shfl_down_sync(0xffffffffu, s.n, o);
            merge_local(g, s, y);
        }
        if (lane == 0) flush_local(g, s, tv.slots + (size_t)idx * p.slot_bytes);
    }
    for (int o = 16; o > 0; o >>= 1) {
        selected += __shfl_down_sync(0xffffffffu, selected, o);
        deferred += __shfl_down_sync(0xffffffffu, deferred, o);
        too_long += __shfl_down_sync(0xffffffffu, too_long, o);
        inserted += __shfl_down_sync(0xffffffffu, inserted, o);
    }
    if (lane == 0) {
        if (selected) atomicAdd(&ctr->rows_selected, selected);
        if (deferred) atomicAdd(&ctr->deferred, deferred);
        if (too_long) atomicAdd(&ctr->key_too_long, too_long);
        if (inserted) atomicAdd(&ctr->inserted, inserted);
    }
}

// moves every ready slot of the old table into the (empty, larger) new one; keys are unique so no state merging
__global__ void k_rehash(const __grid_constant__ DProg prog, const uint8_t* __restrict__ old_slots, uint64_t old_cap, TableView nt) {
    const DProg& p = prog;
    const int sb = p.slot_bytes;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < old_cap; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint8_t* s = old_slots + i * sb;
        uint64_t m = *(const uint64_t*)s;
        if ((m & 3ULL) != SLOT_READY) continue;
        uint64_t idx = ((m & ~3ULL) >> 7) & nt.mask; // tag keeps hash bits 2..63, enough to recompute the home bucket
        for (;;) {
            uint8_t* d = nt.slots + idx * (uint64_t)sb;
            uint64_t old = atomicCAS((unsigned long long*)d, 0ULL, (unsigned long long)m);
            if (old == SLOT_EMPTY) { // the rehash runs alone on the stream: plain copies of the rest of the slot are safe
                for (int b = 8; b < sb; b += 8) *(uint64_t*)(d + b) = *(const uint64_t*)(s + b);
                break;
            }
            idx = (idx + 1) & nt.mask;
        }
    }
}

// ---- cross-rank merge of small group tables (ScanAggOp::set_finishing_merged) ---------------------------------------------------
// wire region of a rank: [u64 slots sent][u64 flag: 1 = table too large to send][slots x slot_bytes]
constexpr int MERGE_MAX_SLOTS = 1024;
__global__ void k_merge_pack(const uint8_t* __restrict__ slots, uint64_t cap, int slot_bytes, uint8_t* __restrict__ wire) {
    const bool fits = cap <= (uint64_t)MERGE_MAX_SLOTS;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        ((uint64_t*)wire)[0] = fits ? cap : 0;
        ((uint64_t*)wire)[1] = fits ? 0 : 1;
    }
    if (!fits) return;
    const uint64_t words = cap * (uint64_t)slot_bytes / 8;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (uint64_t)gridDim.x * blockDim.x)
        ((uint64_t*)(wire + 16))[i] = ((const uint64_t*)slots)[i];
}
// One block. The peers' tables are folded in one after the other: inside one table every key occurs once, so its slots go to distinct
// local slots and plain read-modify-writes are safe between the barriers. States merge as the reference's merge() does: sums and counts
// add (sum.h:113-117, avg.h:161-166, count.h:92-96), MIN / MAX keep the larger order-preserving encoding (maxmin.h).
__global__ void __launch_bounds__(256) k_merge_tables(const __grid_constant__ DProg prog, const TableView tv, const uint8_t* __restrict__ wire,
                                                       const size_t region, const int world, const int me, unsigned long long* __restrict__ inserted,
                                                       unsigned long long* __restrict__ failed) {
    const DProg& p = prog;
    const int sb = p.slot_bytes;
    for (int r = 0; r < world; r++) {
        const uint8_t* reg = wire + (size_t)r * region;
        if (((const uint64_t*)reg)[1] != 0) {
            if (threadIdx.x == 0) atomicAdd(failed, 1ULL);
            continue;
        }
        if (r == me) continue;
        const uint64_t cap = ((const uint64_t*)reg)[0];
        for (uint64_t i = threadIdx.x; i < cap; i += blockDim.x) {
            const uint8_t* s = reg + 16 + i * sb;
            const uint64_t m = *(const uint64_t*)s;
            if ((m & 3ULL) != SLOT_READY) continue;
            bool fresh = false;
            uint8_t* d = table_find_or_insert(tv, ((const uint64_t*)s)[1], p.key_words > 1 ? ((const uint64_t*)s)[2] : 0ULL, m & ~3ULL, &fresh);
            if (!d) {
                atomicAdd(failed, 1ULL);
                continue;
            }
            if (fresh) atomicAdd(inserted, 1ULL);
            for (int a = 0; a < p.n_aggs; a++) {
                const DAgg& g = p.aggs[a];
                if (g.off_n >= 0) *(uint64_t*)(d + g.off_n) += *(const uint64_t*)(s + g.off_n);
                if (g.fn == FN_COUNT || g.off_lo < 0) continue;
                if (g.fn == FN_MIN || g.fn == FN_MAX) {
                    if (g.wide && !g.is_double) {
                        const u128 x = ((u128) * (const uint64_t*)(s + g.off_lo + 8) << 64) | *(const uint64_t*)(s + g.off_lo);
                        const u128 c = ((u128) * (const uint64_t*)(d + g.off_lo + 8) << 64) | *(const uint64_t*)(d + g.off_lo);
                        if (x > c) {
                            *(uint64_t*)(d + g.off_lo) = (uint64_t)x;
                            *(uint64_t*)(d + g.off_lo + 8) = (uint64_t)(x >> 64);
                        }
                    } else if (*(const uint64_t*)(s + g.off_lo) > *(const uint64_t*)(d + g.off_lo)) {
                        *(uint64_t*)(d + g.off_lo) = *(const uint64_t*)(s + g.off_lo);
                    }
                } else if (g.is_double || g.avg_as_double) {
                    *(double*)(d + g.off_lo) += *(const double*)(s + g.off_lo);
                } else {
                    const uint64_t lo = *(const uint64_t*)(s + g.off_lo), old = *(uint64_t*)(d + g.off_lo);
                    *(uint64_t*)(d + g.off_lo) = old + lo;
                    if (g.off_hi >= 0) *(uint64_t*)(d + g.off_hi) += *(const uint64_t*)(s + g.off_hi) + (old + lo < old ? 1ULL : 0ULL);
                }
            }
        }
        __syncthreads();
    }
}

// ---- emit: table -> columnar result (K5), with AVG finalisation on the device --------------------------------
struct EmitAgg {
    void* val;      // out_width bytes per row (blob: the sum part)
    uint8_t* null;  // may be nullptr (COUNT)
    int64_t* cnt;   // blob only
    int32_t out_width;
    int32_t mode;   // 0 count, 1 raw state (sum/min/max/blob sum), 2 avg decimal, 3 avg date, 4 avg largeint->double, 5 avg double
    int32_t scale;  // avg decimal: argument scale
    int32_t decode; // MIN/MAX states are biased encodings (srb_rowpipe.cuh): 0 none, 1 max, 2 min; width/is_double pick the codec
};
struct EmitArgs {
    ulonglong2* keys;      // one record per result row: the packed key, 16 bytes (keys of one or two words) or key_rec_words * 8 (wide keys)
    int32_t key_rec_words; // 2, or the wide key's word count
    int32_t _pad;
    EmitAgg aggs[MAX_AGGS];
};

SRB_D void store_w(void* base, int64_t row, int w, uint64_t lo, uint64_t hi) {
    switch (w) {
    case 1: ((uint8_t*)base)[row] = (uint8_t)lo; break;
    case 2: ((uint16_t*)base)[row] = (uint16_t)lo; break;
    case 4: ((uint32_t*)base)[row] = (uint32_t)lo; break;
    case 8: ((uint64_t*)base)[row] = lo; break;
    default: ((ulonglong2*)base)[row] = make_ulonglong2(lo, hi);
    }
}

SRB_D void emit_row(const DProg& p, const uint8_t* s, const EmitArgs& ea, int64_t row);

// gathers the pieces of a small result (a row range of every output column) into one contiguous buffer, so that a pull
// costs one device->host copy instead of one per column
struct PackArgs {
    static constexpr int MAX = 40;
    int n;
    const uint8_t* src[MAX];
    uint32_t bytes[MAX];
    uint32_t dst_off[MAX];
};
__global__ void k_pack_result(const __grid_constant__ PackArgs a, uint8_t* __restrict__ dst) {
    for (int i = blockIdx.x; i < a.n; i += gridDim.x) {
        const uint8_t* s = a.src[i];
        uint8_t* d = dst + a.dst_off[i];
        for (uint32_t b = threadIdx.x; b < a.bytes[i]; b += blockDim.x) d[b] = s[b];
    }
}

// One output position per occupied slot. A block takes 8 x 256 slots at a time, counts its occupied ones, scans the
// counts and reserves the block's output range with ONE atomic (a ticket per warp made the single counter the bottleneck
// of a 10^8-group emit).
constexpr int EMIT_ROWS = 8;
__global__ void __launch_bounds__(256) k_emit(const __grid_constant__ DProg prog, const uint8_t* __restrict__ slots, uint64_t cap,
                                              const __grid_constant__ EmitArgs ea, Counters* __restrict__ ctr, const int64_t row_base) {
    const DProg& p = prog;
    const int sb = p.slot_bytes;
    __shared__ unsigned int warp_tot[8];
    __shared__ unsigned long long block_base;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint64_t chunk = (uint64_t)EMIT_ROWS * 256;
    for (uint64_t c0 = (uint64_t)blockIdx.x * chunk; c0 < cap; c0 += (uint64_t)gridDim.x * chunk) {
        unsigned int ready = 0;
        // 32-byte slots (k -> sum, count: the high-cardinality plans, ~10^8 of them): the whole slot is fetched with two 16-byte loads per
        // row, all of a thread's sixteen loads in flight together, and the row is emitted from registers — the table is read once
        ulonglong2 s0[EMIT_ROWS], s1[EMIT_ROWS];
        const bool whole = sb == 32;
        if (whole) {
#pragma unroll
            for (int r = 0; r < EMIT_ROWS; r++) {
                const uint64_t i = c0 + (uint64_t)r * 256 + tid;
                s0[r] = s1[r] = make_ulonglong2(0, 0);
                if (i < cap) {
                    const ulonglong2* q = reinterpret_cast<const ulonglong2*>(slots + i * 32);
                    s0[r] = __ldg(q);
                    s1[r] = __ldg(q + 1);
                }
            }
#pragma unroll
            for (int r = 0; r < EMIT_ROWS; r++)
                if ((s0[r].x & 3ULL) == SLOT_READY) ready |= 1u << r;
        } else {
#pragma unroll
            for (int r = 0; r < EMIT_ROWS; r++) {
                const uint64_t i = c0 + (uint64_t)r * 256 + tid;
                if (i < cap && ((*(const uint64_t*)(slots + i * sb)) & 3ULL) == SLOT_READY) ready |= 1u << r;
            }
        }
        // (Ranking the lanes of each slot row by ballot, so that every store instruction writes a dense run of output rows, was measured
        // slower: 28.5 vs 26.9 ms per high-cardinality step. A thread's own ready slots take consecutive rows.)
        const unsigned int mine = __popc(ready);
        unsigned int incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        if (tid == 0) {
            unsigned int t = 0;
            for (int w = 0; w < 8; w++) {
                const unsigned int x = warp_tot[w];
                warp_tot[w] = t;
                t += x;
            }
            block_base = t ? atomicAdd(&ctr->emitted, (unsigned long long)t) : 0ULL;
        }
        __syncthreads();
        int64_t row = row_base + (int64_t)(block_base + warp_tot[wid] + (incl - mine)); // row_base: pass-through rows of a streaming sink come first
#pragma unroll
        for (int r = 0; r < EMIT_ROWS; r++)
            if (ready & (1u << r)) {
                if (whole) {
                    const uint64_t copy[4] = {s0[r].x, s0[r].y, s1[r].x, s1[r].y};
                    emit_row(p, reinterpret_cast<const uint8_t*>(copy), ea, row++);
                } else {
                    emit_row(p, slots + (c0 + (uint64_t)r * 256 + tid) * sb, ea, row++);
                }
            }
        __syncthreads(); // warp_tot / block_base are rewritten by the next chunk
    }
}

// one aggregate of one output row from its state words; raw: MIN / MAX words are plain values (a pass-through row's own value),
// not the table's biased encodings
SRB_D void emit_agg(const DAgg& g, const EmitAgg& e, uint64_t n, uint64_t lo, uint64_t hi, int64_t row, bool raw) {
    bool isnull = e.null != nullptr && n == 0;
    if (e.null) e.null[row] = isnull ? 1 : 0;
    if (e.cnt) e.cnt[row] = (int64_t)n;
    if (isnull) {
        store_w(e.val, row, e.out_width, 0, 0);
        return;
    }
    switch (e.mode) {
    case 0: store_w(e.val, row, 8, n, 0); break;
    case 1:
        if (e.decode && !raw) {
            const bool mn = e.decode == 2;
            if (g.is_double)
                lo = dec_f64(lo, mn);
            else if (g.wide) {
                i128 x = dec_i128(((u128)hi << 64) | lo, mn);
                lo = lo64(x);
                hi = (uint64_t)hi64(x);
            } else
                lo = (uint64_t)dec_i64(lo, mn);
        }
        store_w(e.val, row, e.out_width, lo, hi);
        break;
    case 2: {
        i128 r = decimal_div_integer(make_i128(lo, (int64_t)hi), (i128)(int64_t)n, e.scale);
        store_w(e.val, row, 16, lo64(r), (uint64_t)hi64(r));
        break;
    }
    case 3: {
        i128 r = make_i128(lo, (int64_t)hi) / (i128)(int64_t)n; // DATE avg: truncating (avg.h:231)
        store_w(e.val, row, 4, lo64(r), 0);
        break;
    }
    case 4: {
        double r = (double)make_i128(lo, (int64_t)hi) / (double)(int64_t)n;
        store_w(e.val, row, 8, (uint64_t)__double_as_longlong(r), 0);
        break;
    }
    default: {
        double r = __longlong_as_double((long long)lo) / (double)(int64_t)n;
        store_w(e.val, row, 8, (uint64_t)__double_as_longlong(r), 0);
    }
    }
}

SRB_D void emit_row(const DProg& p, const uint8_t* s, const EmitArgs& ea, int64_t row) {
    if (p.n_keys) {
        const uint64_t* k = (const uint64_t*)s;
        if (p.key_words > 2) { // word 0 in pair A, word 1 in pair B, the rest behind pair B
            uint64_t* rec = (uint64_t*)ea.keys + (size_t)row * ea.key_rec_words;
            rec[0] = k[1];
            rec[1] = k[2];
            for (int i = 2; i < p.key_words; i++) rec[i] = k[4 + (i - 2)];
        } else {
            ea.keys[row] = make_ulonglong2(k[1], p.key_words > 1 ? k[2] : 0ULL);
        }
    }
    for (int a = 0; a < p.n_aggs; a++) {
        const DAgg& g = p.aggs[a];
        const uint64_t n = g.off_n >= 0 ? *(const uint64_t*)(s + g.off_n) : 1;
        const uint64_t lo = g.off_lo >= 0 ? *(const uint64_t*)(s + g.off_lo) : 0;
        const uint64_t hi = g.off_hi >= 0 ? *(const uint64_t*)(s + g.off_hi) : (uint64_t)((int64_t)lo >> 63);
        emit_agg(g, ea.aggs[a], n, lo, hi, row, false);
    }
}

// Streaming pre-aggregation (AggregateStreamingSinkOperator, pipeline/aggregate/aggregate_streaming_sink_operator.cpp:93-145):
// rows that are not aggregated leave in intermediate format, one output row each (Aggregator::output_chunk_by_streaming ->
// AggregateFunction::convert_to_serialize_format, exprs/agg/aggregate.h:337): keys as they are, every aggregate the serialized
// state of that single row. selective = 0: every selected row passes through (force_streaming); selective = 1: rows whose group
// is already in the table update it in place and only the others pass through (selective_preaggregation; the lookup never
// inserts). Output positions come from one atomic per warp step; the order of pass-through rows is not kept (phase 2 does not see it).
__global__ void __launch_bounds__(256) k_stream_rows(const __grid_constant__ DProg prog, const TableView tv, const int64_t base, const int64_t n,
                                                      const int selective, const __grid_constant__ EmitArgs ea, const int64_t row_base,
                                                      Counters* __restrict__ ctr) {
    const DProg& p = prog;
    Val v[MAX_NODES];
    unsigned long long selected = 0, too_long = 0, hits = 0;
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // the loop bound is warp-uniform (a warp's 32 rows are consecutive), so every lane reaches the ballot below in every iteration
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); i0 < n; i0 += stride) {
        const int64_t row = base + i0 + lane;
        bool pass = i0 + lane < n && eval_preds(p, row);
        PackedKey k{};
        WideKey wk;
        const bool wide = p.key_words > 2;
        if (pass) {
            selected++;
            for (int q = 0; q < p.n_nodes; q++) v[q] = eval_node(p.nodes[q], v, p.cols, row);
            bool too;
            if (wide) {
                wk = pack_key_wide(p, v, row);
                too = wk.too_long;
            } else {
                k = pack_key(p, v, row);
                too = k.too_long;
            }
            if (too) {
                too_long++;
                pass = false;
            }
        }
        if (pass && selective) {
            uint8_t* slot = wide ? table_find_wide(tv, wk, hash_wide_key(wk, p.phase_seed)) : table_find(tv, k.lo, k.hi, hash_packed_key(k, p.phase_seed));
            if (slot) {
                for (int a = 0; a < p.n_aggs; a++) update_agg_global(p, p.aggs[a], slot, v, row);
                hits++;
                pass = false;
            }
        }
        // the lanes that pass a row through reserve their output rows with one atomic
        const unsigned int m = __ballot_sync(0xffffffffu, pass);
        if (m == 0) continue;
        const int leader = __ffs(m) - 1;
        unsigned long long first = 0;
        if (lane == leader) first = atomicAdd(&ctr->emitted, (unsigned long long)__popc(m));
        first = __shfl_sync(0xffffffffu, first, leader);
        if (!pass) continue;
        const int64_t pos = row_base + (int64_t)first + __popc(m & ((1u << lane) - 1u));
        if (wide) {
            uint64_t* rec = (uint64_t*)ea.keys + (size_t)pos * ea.key_rec_words;
            for (int i = 0; i < p.key_words; i++) rec[i] = wk.w[i];
        } else {
            ea.keys[pos] = make_ulonglong2(k.lo, p.key_words > 1 ? k.hi : 0ULL);
        }
        for (int a = 0; a < p.n_aggs; a++) {
            const DAgg& g = p.aggs[a];
            LocalAcc d{0, 0, 0};
            update_agg_local(p, g, d, v, row); // the state of this one row
            const uint64_t cnt = g.off_n >= 0 ? d.n : 1;
            const uint64_t hi = g.off_hi >= 0 ? d.hi : (uint64_t)((int64_t)d.lo >> 63);
            emit_agg(g, ea.aggs[a], cnt, d.lo, hi, pos, true);
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        selected += __shfl_down_sync(0xffffffffu, selected, o);
        too_long += __shfl_down_sync(0xffffffffu, too_long, o);
        hits += __shfl_down_sync(0xffffffffu, hits, o);
    }
    if (lane == 0) {
        if (selected) atomicAdd(&ctr->rows_selected, selected);
        if (too_long) atomicAdd(&ctr->key_too_long, too_long);
        if (hits) atomicAdd(&ctr->inserted, hits); // this kernel never inserts: the field carries the hit count of the batch
    }
}

// ---- filter-only mode: predicates -> bitmap + per-tile popcounts (K1), then order-preserving compaction (K2) ------
constexpr int FTILE = 8192; // rows per tile = 256 mask words

__global__ void __launch_bounds__(256) k_filter_mask(const __grid_constant__ DProg prog, int64_t base, int64_t n, uint32_t* __restrict__ mask,
                                                      uint32_t* __restrict__ tile_counts) {
    const DProg& p = prog;
    __shared__ uint32_t wcount[8];
    const int64_t tile0 = (int64_t)blockIdx.x * FTILE;
    uint32_t mine = 0;
    for (int it = 0; it < FTILE / 256; it++) {
        int64_t i = tile0 + it * 256 + threadIdx.x;
        bool pass = i < n && eval_preds(p, base + i);
        uint32_t w = __ballot_sync(0xffffffffu, pass);
        if ((threadIdx.x & 31) == 0) {
            if (i < n) mask[i >> 5] = w;
            mine += __popc(w);
        }
    }
    if ((threadIdx.x & 31) == 0) wcount[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; w++) t += wcount[w];
        tile_counts[blockIdx.x] = t;
    }
}

// The same bitmap for predicates that lower to per-column ranges over NULL-free 4- / 8-byte columns (lower_ranges): every
// thread takes four consecutive rows per step with 128-bit loads (one LDG.128 per 4-byte column, two per 8-byte column), all
// loads of four steps are issued before the first result is used, and a warp assembles its four mask words with three
// shuffles. One subtract + one unsigned compare per row and range whatever the operators were.
struct FRange {
    const uint8_t* data; // already offset to the first row of the range being filtered; 16-byte aligned
    int32_t width;       // 4 or 8
    int32_t neg;
    int64_t lo;
    uint64_t span;
};
struct FRangeArgs {
    int32_t n;
    int32_t _pad;
    FRange r[ST_MAX_RANGES];
};

template <int I>
SRB_D uint32_t quad_ranges(const FRangeArgs& fa, int64_t row, uint32_t nib) {
    if constexpr (I < ST_MAX_RANGES) {
        if (I < fa.n) {
            const FRange& g = fa.r[I];
            uint32_t b;
            if (g.width == 4) {
                const uint4 v = __ldg((const uint4*)(g.data + row * 4));
                const uint32_t lo = (uint32_t)g.lo, span = (uint32_t)g.span;
                b = (uint32_t)(v.x - lo <= span) | ((uint32_t)(v.y - lo <= span) << 1) | ((uint32_t)(v.z - lo <= span) << 2) | ((uint32_t)(v.w - lo <= span) << 3);
            } else {
                const ulonglong2 v0 = __ldg((const ulonglong2*)(g.data + row * 8));
                const ulonglong2 v1 = __ldg((const ulonglong2*)(g.data + row * 8 + 16));
                const uint64_t lo = (uint64_t)g.lo;
                b = (uint32_t)(v0.x - lo <= g.span) | ((uint32_t)(v0.y - lo <= g.span) << 1) | ((uint32_t)(v1.x - lo <= g.span) << 2) |
                    ((uint32_t)(v1.y - lo <= g.span) << 3);
            }
            nib &= g.neg ? ~b : b;
            return quad_ranges<I + 1>(fa, row, nib);
        }
    }
    return nib;
}

SRB_D uint32_t quad_ranges_tail(const FRangeArgs& fa, int64_t row, int64_t n) { // the last, partial quad of the range
    uint32_t nib = 0;
    for (int j = 0; j < 4 && row + j < n; j++) {
        bool pass = true;
        for (int k = 0; k < fa.n; k++) {
            const FRange& g = fa.r[k];
            bool in;
            if (g.width == 4) in = (uint32_t)(((const uint32_t*)g.data)[row + j] - (uint32_t)g.lo) <= (uint32_t)g.span;
            else in = (uint64_t)(((const uint64_t*)g.data)[row + j] - (uint64_t)g.lo) <= g.span;
            pass = pass && (in != (g.neg != 0));
        }
        nib |= (uint32_t)pass << j;
    }
    return nib;
}

// a single range over a W-byte column: straight-line code, so the loads of all four steps are issued back to back
template <int W>
SRB_D uint32_t quad_one_range(const FRange& g, int64_t row) {
    uint32_t b;
    if constexpr (W == 4) {
        const uint4 v = __ldg((const uint4*)(g.data + row * 4));
        const uint32_t lo = (uint32_t)g.lo, span = (uint32_t)g.span;
        b = (uint32_t)(v.x - lo <= span) | ((uint32_t)(v.y - lo <= span) << 1) | ((uint32_t)(v.z - lo <= span) << 2) | ((uint32_t)(v.w - lo <= span) << 3);
    } else {
        const ulonglong2 v0 = __ldg((const ulonglong2*)(g.data + row * 8));
        const ulonglong2 v1 = __ldg((const ulonglong2*)(g.data + row * 8 + 16));
        const uint64_t lo = (uint64_t)g.lo;
        b = (uint32_t)(v0.x - lo <= g.span) | ((uint32_t)(v0.y - lo <= g.span) << 1) | ((uint32_t)(v1.x - lo <= g.span) << 2) | ((uint32_t)(v1.y - lo <= g.span) << 3);
    }
    return (g.neg ? ~b : b) & 0xFu;
}

// MODE 0: any conjunction of ranges; MODE 4 / 8: exactly one range over a 4- / 8-byte column
template <int MODE>
__global__ void __launch_bounds__(256) k_filter_mask_ranges(const __grid_constant__ FRangeArgs fa, int64_t n, uint32_t* __restrict__ mask,
                                                             uint32_t* __restrict__ tile_counts) {
    __shared__ uint32_t wcount[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t tile0 = (int64_t)blockIdx.x * FTILE;
    uint32_t mine = 0;
#pragma unroll 1
    for (int it = 0; it < FTILE / 1024; it += 4) {
        uint32_t nib[4];
#pragma unroll
        for (int u = 0; u < 4; u++) { // a warp covers 128 consecutive rows per step; steps of one block are 1024 rows apart
            const int64_t row = tile0 + (int64_t)(it + u) * 1024 + warp * 128 + lane * 4;
            if constexpr (MODE == 0) {
                nib[u] = row + 4 <= n ? quad_ranges<0>(fa, row, 0xFu) : (row < n ? quad_ranges_tail(fa, row, n) : 0u);
            } else {
                nib[u] = row + 4 <= n ? quad_one_range<MODE>(fa.r[0], row) : 0x10u; // 0x10: not a whole quad, settled below
            }
        }
        if constexpr (MODE != 0) {
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (nib[u] == 0x10u) {
                    const int64_t row = tile0 + (int64_t)(it + u) * 1024 + warp * 128 + lane * 4;
                    nib[u] = row < n ? quad_ranges_tail(fa, row, n) : 0u;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            uint32_t w = (nib[u] & 0xFu) << (4 * (lane & 7)); // word j of the step = lanes 8j .. 8j+7
            w |= __shfl_xor_sync(0xffffffffu, w, 1);
            w |= __shfl_xor_sync(0xffffffffu, w, 2);
            w |= __shfl_xor_sync(0xffffffffu, w, 4);
            const int64_t row = tile0 + (int64_t)(it + u) * 1024 + warp * 128 + lane * 4;
            if ((lane & 7) == 0 && row < n) {
                mask[row >> 5] = w;
                mine += __popc(w);
            }
        }
    }
    mine += __shfl_xor_sync(0xffffffffu, mine, 8);
    mine += __shfl_xor_sync(0xffffffffu, mine, 16);
    if (lane == 0) wcount[warp] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; w++) t += wcount[w];
        tile_counts[blockIdx.x] = t;
    }
}

// Runtime filter on the scan side (Operator::eval_runtime_bloom_filters, pipeline/operator.h:185; RuntimeBloomFilter::_t_evaluate,
// exprs/runtime_filter.h:1369-1422): rows still selected by the mask keep their bit only if the key lies in [min, max] and the
// bucket of its hash has all eight bits set; NULL keys keep it iff the build side inserted a NULL. Four rows per thread and step:
// the key loads of the four rows are issued together, then their 32-byte bucket loads (the directory of a build side of a few
// 10^7 keys is tens of MB: L2-resident), then the ballots rewrite the mask words and the tile's count.
// one 256-bit load (sm_100: LDG.E.256) fetches a whole bucket = one 32-byte sector, one L2 request per test
struct alignas(32) RfBucket {
    uint32_t w[8];
};
SRB_D RfBucket ld_bucket(const uint32_t* p) {
    RfBucket r;
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]), "=r"(r.w[7])
                 : "l"(p));
    return r;
}
struct RfArgs {
    RfView v;
    const uint8_t* data; // key column, already offset to the first row of the range
    const uint8_t* nulls;
};
// Partly selected tiles: the threads walk the list of selected row numbers four entries at a time — key loads together, bucket
// loads together, tests — and clear the bit of every row that fails. Entries past the end of the list repeat row 0 of the tile
// (always a valid address) and are ignored at the end, so that no load sits behind a predicate.
template <int W, bool MIX, bool NULLS, class KT>
SRB_D void rf_test_list(const RfArgs& a, const uint32_t* __restrict__ dir, const KT* __restrict__ keys, const uint8_t* __restrict__ nulls, uint32_t total,
                        const uint16_t* sidx, uint32_t* smask) {
    constexpr long long KMIN = (long long)std::numeric_limits<KT>::min(), KMAX = (long long)std::numeric_limits<KT>::max();
    const long long lo64 = a.v.min > KMIN ? a.v.min : KMIN, hi64 = a.v.max < KMAX ? a.v.max : KMAX;
    const bool empty = lo64 > hi64;
    const KT lo = empty ? (KT)1 : (KT)lo64, hi = empty ? (KT)0 : (KT)hi64;
    const uint32_t dmask = a.v.directory_mask;
    const int lnb = a.v.log_num_buckets;
    const bool null_pass = a.v.has_null != 0;
#pragma unroll 1
    for (uint32_t j0 = 0; j0 < total; j0 += 256 * 4) {
        uint32_t r[4];
        KT x[4];
        uint8_t nl[4] = {};
        uint64_t h[4];
        RfBucket b[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t j = j0 + u * 256 + threadIdx.x;
            r[u] = j < total ? (uint32_t)sidx[j] : 0xFFFFFFFFu;
        }
#pragma unroll
        for (int u = 0; u < 4; u++) x[u] = keys[r[u] == 0xFFFFFFFFu ? 0u : r[u]];
        if constexpr (NULLS) {
#pragma unroll
            for (int u = 0; u < 4; u++) nl[u] = nulls[r[u] == 0xFFFFFFFFu ? 0u : r[u]];
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            h[u] = MIX ? phmap_mix8((uint64_t)(long long)x[u], 1) : (uint64_t)(long long)x[u];
            b[u] = ld_bucket(dir + (size_t)((uint32_t)h[u] & dmask) * 8);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t key = (uint32_t)(h[u] >> lnb);
            uint32_t miss = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) miss |= ~b[u].w[i] & rf_word_mask(key, i);
            bool pass = miss == 0 && x[u] >= lo && x[u] <= hi;
            if constexpr (NULLS) pass = nl[u] ? null_pass : pass;
            if (!pass && r[u] != 0xFFFFFFFFu) atomicAnd(&smask[r[u] >> 5], ~(1u << (r[u] & 31)));
        }
    }
}

// Full tiles (every row still selected: the first filter of a scan without predicates) take a leaner body: the hash flavour and
// the presence of a null map are compile-time, the [min, max] test runs in the key's own width (the bounds are clamped to the
// key type's range on entry), bucket loads are unconditional (the masked index is always inside the directory) and a warp's 32
// verdicts are one ballot = one mask word.
template <int W, bool MIX, bool NULLS, class KT>
SRB_D void rf_test_full(const RfArgs& a, const uint32_t* __restrict__ dir, const KT* __restrict__ keys, const uint8_t* __restrict__ nulls, uint32_t* smask,
                        int lane) {
    constexpr long long KMIN = (long long)std::numeric_limits<KT>::min(), KMAX = (long long)std::numeric_limits<KT>::max();
    const long long lo64 = a.v.min > KMIN ? a.v.min : KMIN, hi64 = a.v.max < KMAX ? a.v.max : KMAX;
    const bool empty = lo64 > hi64; // nothing inserted, or every inserted key lies outside the probe key type's range
    const KT lo = empty ? (KT)1 : (KT)lo64, hi = empty ? (KT)0 : (KT)hi64;
    const uint32_t dmask = a.v.directory_mask;
    const int lnb = a.v.log_num_buckets;
    const bool null_pass = a.v.has_null != 0;
#pragma unroll 1
    for (uint32_t j0 = 0; j0 < (uint32_t)FTILE; j0 += 256 * 4) {
        KT x[4];
        uint8_t nl[4] = {};
        uint64_t h[4];
        RfBucket b[4];
#pragma unroll
        for (int u = 0; u < 4; u++) x[u] = keys[j0 + u * 256 + threadIdx.x];
        if constexpr (NULLS) {
#pragma unroll
            for (int u = 0; u < 4; u++) nl[u] = nulls[j0 + u * 256 + threadIdx.x];
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            h[u] = MIX ? phmap_mix8((uint64_t)(long long)x[u], 1) : (uint64_t)(long long)x[u];
            b[u] = ld_bucket(dir + (size_t)((uint32_t)h[u] & dmask) * 8);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t key = (uint32_t)(h[u] >> lnb);
            uint32_t miss = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) miss |= ~b[u].w[i] & rf_word_mask(key, i);
            bool pass = miss == 0 && x[u] >= lo && x[u] <= hi;
            if constexpr (NULLS) pass = nl[u] ? null_pass : pass;
            const uint32_t w = __ballot_sync(0xffffffffu, pass);
            if (lane == 0) smask[(j0 + u * 256 + threadIdx.x) >> 5] = w;
        }
    }
}

// RuntimeBitsetFilter (exprs/runtime_filter.h:1583-1700, ComposedRuntimeFilter's min / max in front of it): a row passes iff its key lies in
// [min, max] and bit (key - min) is set; NULL keys pass iff has_null. One word load and a shift instead of a 32-byte bucket and eight
// salted multiplies; the bitsets of dimension keys (a few hundred KB) stay in L1 / L2.
constexpr int RF_BITS_U = 8; // rows per thread and step of the bitset paths: eight independent key loads in flight (the test is a load and a shift: latency is all there is)
template <bool NULLS, class KT>
SRB_D void rf_bits_full(const RfArgs& a, const KT* __restrict__ keys, const uint8_t* __restrict__ nulls, uint32_t* smask, int lane) {
    constexpr long long KMIN = (long long)std::numeric_limits<KT>::min(), KMAX = (long long)std::numeric_limits<KT>::max();
    const long long lo64 = a.v.min > KMIN ? a.v.min : KMIN, hi64 = a.v.max < KMAX ? a.v.max : KMAX;
    const bool empty = lo64 > hi64;
    const KT lo = empty ? (KT)1 : (KT)lo64, hi = empty ? (KT)0 : (KT)hi64;
    const uint32_t* __restrict__ bits = a.v.bits;
    const long long mn = a.v.min;
    const bool null_pass = a.v.has_null != 0;
#pragma unroll 1
    for (uint32_t j0 = 0; j0 < (uint32_t)FTILE; j0 += 256 * RF_BITS_U) {
        KT x[RF_BITS_U];
        uint8_t nl[RF_BITS_U] = {};
        uint64_t idx[RF_BITS_U];
        uint32_t w[RF_BITS_U];
        bool in[RF_BITS_U];
#pragma unroll
        for (int u = 0; u < RF_BITS_U; u++) x[u] = keys[j0 + u * 256 + threadIdx.x];
        if constexpr (NULLS) {
#pragma unroll
            for (int u = 0; u < RF_BITS_U; u++) nl[u] = nulls[j0 + u * 256 + threadIdx.x];
        }
#pragma unroll
        for (int u = 0; u < RF_BITS_U; u++) {
            in[u] = x[u] >= lo && x[u] <= hi;
            idx[u] = in[u] ? (uint64_t)((long long)x[u] - mn) : 0ULL;
            w[u] = __ldg(bits + (idx[u] >> 5));
        }
#pragma unroll
        for (int u = 0; u < RF_BITS_U; u++) {
            bool pass = in[u] && ((w[u] >> (idx[u] & 31)) & 1u);
            if constexpr (NULLS) pass = nl[u] ? null_pass : pass;
            const uint32_t word = __ballot_sync(0xffffffffu, pass);
            if (lane == 0) smask[(j0 + u * 256 + threadIdx.x) >> 5] = word;
        }
    }
}
template <bool NULLS, class KT>
SRB_D void rf_bits_list(const RfArgs& a, const KT* __restrict__ keys, const uint8_t* __restrict__ nulls, uint32_t total, const uint16_t* sidx,
                        uint32_t* smask) {
    constexpr long long KMIN = (long long)std::numeric_limits<KT>::min(), KMAX = (long long)std::numeric_limits<KT>::max();
    const long long lo64 = a.v.min > KMIN ? a.v.min : KMIN, hi64 = a.v.max < KMAX ? a.v.max : KMAX;
    const bool empty = lo64 > hi64;
    const KT lo = empty ? (KT)1 : (KT)lo64, hi = empty ? (KT)0 : (KT)hi64;
    const uint32_t* __restrict__ bits = a.v.bits;
    const long long mn = a.v.min;
    const bool null_pass = a.v.has_null != 0;
#pragma unroll 1
    for (uint32_t j0 = 0; j0 < total; j0 += 256 * RF_BITS_U) {
        uint32_t r[RF_BITS_U], w[RF_BITS_U];
        KT x[RF_BITS_U];
        uint8_t nl[RF_BITS_U] = {};
        uint64_t idx[RF_BITS_U];
        bool in[RF_BITS_U];
#pragma unroll
        for (int u = 0; u < RF_BITS_U; u++) {
            const uint32_t j = j0 + u * 256 + threadIdx.x;
            r[u] = j < total ? (uint32_t)sidx[j] : 0xFFFFFFFFu;
        }
#pragma unroll
        for (int u = 0; u < RF_BITS_U; u++) x[u] = keys[r[u] == 0xFFFFFFFFu ? 0u : r[u]];
        if constexpr (NULLS) {
#pragma unroll
            for (int u = 0; u < RF_BITS_U; u++) nl[u] = nulls[r[u] == 0xFFFFFFFFu ? 0u : r[u]];
        }
#pragma unroll
        for (int u = 0; u < RF_BITS_U; u++) {
            in[u] = x[u] >= lo && x[u] <= hi;
            idx[u] = in[u] ? (uint64_t)((long long)x[u] - mn) : 0ULL;
            w[u] = __ldg(bits + (idx[u] >> 5));
        }
#pragma unroll
        for (int u = 0; u < RF_BITS_U; u++) {
            bool pass = in[u] && ((w[u] >> (idx[u] & 31)) & 1u);
            if constexpr (NULLS) pass = nl[u] ? null_pass : pass;
            if (!pass && r[u] != 0xFFFFFFFFu) atomicAnd(&smask[r[u] >> 5], ~(1u << (r[u] & 31)));
        }
    }
}

template <int W>
__global__ void __launch_bounds__(256, 4) k_filter_rf(const __grid_constant__ RfArgs a, int64_t n, uint32_t* __restrict__ mask,
                                                       uint32_t* __restrict__ tile_counts) {
    // Work in proportion to the rows still selected, not to the rows of the tile: the tile's 256 mask words are expanded into a
    // list of selected row numbers in shared memory (prefix sum of the popcounts), and the threads then walk that dense list four
    // entries at a time — key loads together, bucket loads together, tests — clearing the bit of every row that fails. A scan whose
    // earlier filters already dropped 80 % or 96 % of the rows (the star join's second and third filter) does 5x / 25x less work,
    // and no instruction is spent on predicated-off rows.
    using KT = typename std::conditional<W == 8, long long, typename std::conditional<W == 4, int, typename std::conditional<W == 2, short, signed char>::type>::type>::type;
    __shared__ uint32_t smask[FTILE / 32];
    __shared__ uint16_t sidx[FTILE];
    __shared__ uint32_t wsum[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t tile0 = (int64_t)blockIdx.x * FTILE;
    const int64_t w0 = tile0 >> 5, n_words = (n + 31) >> 5;
    const uint32_t word = w0 + threadIdx.x < n_words ? mask[w0 + threadIdx.x] : 0u; // bits past n were never set
    smask[threadIdx.x] = word;
    const uint32_t c = __popc(word);
    uint32_t inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const uint32_t t = wsum[w];
        before += w < warp ? t : 0;
        total += t;
    }
    if (total != FTILE) { // expansion: every thread lists the set bits of its own word
        uint32_t off = before + inc - c, bits = word;
        while (bits) {
            sidx[off++] = (uint16_t)(threadIdx.x * 32 + __ffs(bits) - 1);
            bits &= bits - 1;
        }
    }
    __syncthreads();
    const uint32_t* dir = a.v.dir;
    const KT* keys = (const KT*)a.data + tile0;
    const uint8_t* nulls = a.nulls ? a.nulls + tile0 : nullptr;
    if (a.v.bits) { // a bitset filter (uniform: one filter per launch)
        if (total == FTILE) {
            if (nulls) rf_bits_full<true, KT>(a, keys, nulls, smask, lane);
            else rf_bits_full<false, KT>(a, keys, nulls, smask, lane);
        } else {
            if (nulls) rf_bits_list<true, KT>(a, keys, nulls, total, sidx, smask);
            else rf_bits_list<false, KT>(a, keys, nulls, total, sidx, smask);
        }
    } else if (total == FTILE) { // every row of the tile is still selected: rows are their own list
        if (a.v.mix) {
            if (nulls) rf_test_full<W, true, true, KT>(a, dir, keys, nulls, smask, lane);
            else rf_test_full<W, true, false, KT>(a, dir, keys, nulls, smask, lane);
        } else {
            if (nulls) rf_test_full<W, false, true, KT>(a, dir, keys, nulls, smask, lane);
            else rf_test_full<W, false, false, KT>(a, dir, keys, nulls, smask, lane);
        }
    } else if (a.v.mix) {
        if (nulls) rf_test_list<W, true, true, KT>(a, dir, keys, nulls, total, sidx, smask);
        else rf_test_list<W, true, false, KT>(a, dir, keys, nulls, total, sidx, smask);
    } else {
        if (nulls) rf_test_list<W, false, true, KT>(a, dir, keys, nulls, total, sidx, smask);
        else rf_test_list<W, false, false, KT>(a, dir, keys, nulls, total, sidx, smask);
    }
    __syncthreads();
    const uint32_t now = smask[threadIdx.x];
    if (now != word) mask[w0 + threadIdx.x] = now;
    uint32_t kept = __popc(now);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kept += __shfl_down_sync(0xffffffffu, kept, o);
    if (lane == 0) wsum[warp] = kept;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; w++) t += wsum[w];
        tile_counts[blockIdx.x] = t;
    }
}

// Several runtime filters in ONE pass over a tile (opt-in: SRB_RF_FUSED=1, all key columns of one width): the tile's mask words stay
// in shared memory between the filters; after each filter the list of still-selected rows is rebuilt from them, so the second and
// third filter of a star join walk 20 % / 4 % of the rows without a kernel launch, a global mask round trip or a pass over the
// tiles of their own. Same tests, same result as the sequence of k_filter_rf launches.
constexpr int RF_FUSED_MAX = 4;
struct RfMultiArgs {
    int32_t n;
    RfArgs f[RF_FUSED_MAX];
};
template <int W>
__global__ void __launch_bounds__(256, 4) k_filter_rf_multi(const __grid_constant__ RfMultiArgs a, int64_t n, uint32_t* __restrict__ mask,
                                                             uint32_t* __restrict__ tile_counts) {
    using KT = typename std::conditional<W == 8, long long, typename std::conditional<W == 4, int, typename std::conditional<W == 2, short, signed char>::type>::type>::type;
    __shared__ uint32_t smask[FTILE / 32];
    __shared__ uint16_t sidx[FTILE];
    __shared__ uint32_t wsum[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t tile0 = (int64_t)blockIdx.x * FTILE;
    const int64_t w0 = tile0 >> 5, n_words = (n + 31) >> 5;
    const uint32_t word0 = w0 + threadIdx.x < n_words ? mask[w0 + threadIdx.x] : 0u;
    smask[threadIdx.x] = word0;
    __syncthreads();
    for (int fi = 0; fi < a.n; fi++) {
        const RfArgs& f = a.f[fi];
        // list of the rows still selected (prefix sum of the popcounts of this thread's word, as in k_filter_rf)
        const uint32_t word = smask[threadIdx.x];
        const uint32_t c = __popc(word);
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += y;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        uint32_t before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            const uint32_t t = wsum[w];
            before += w < warp ? t : 0;
            total += t;
        }
        if (total != FTILE) {
            uint32_t off = before + inc - c, bits = word;
            while (bits) {
                sidx[off++] = (uint16_t)(threadIdx.x * 32 + __ffs(bits) - 1);
                bits &= bits - 1;
            }
        }
        __syncthreads();
        const uint32_t* dir = f.v.dir;
        const KT* keys = (const KT*)f.data + tile0;
        const uint8_t* nulls = f.nulls ? f.nulls + tile0 : nullptr;
        if (total == FTILE) {
            if (f.v.mix) {
                if (nulls) rf_test_full<W, true, true, KT>(f, dir, keys, nulls, smask, lane);
                else rf_test_full<W, true, false, KT>(f, dir, keys, nulls, smask, lane);
            } else {
                if (nulls) rf_test_full<W, false, true, KT>(f, dir, keys, nulls, smask, lane);
                else rf_test_full<W, false, false, KT>(f, dir, keys, nulls, smask, lane);
            }
        } else if (total != 0) {
            if (f.v.mix) {
                if (nulls) rf_test_list<W, true, true, KT>(f, dir, keys, nulls, total, sidx, smask);
                else rf_test_list<W, true, false, KT>(f, dir, keys, nulls, total, sidx, smask);
            } else {
                if (nulls) rf_test_list<W, false, true, KT>(f, dir, keys, nulls, total, sidx, smask);
                else rf_test_list<W, false, false, KT>(f, dir, keys, nulls, total, sidx, smask);
            }
        }
        __syncthreads(); // smask is final for this filter; wsum / sidx are free again
    }
    const uint32_t now = smask[threadIdx.x];
    if (now != word0) mask[w0 + threadIdx.x] = now;
    uint32_t kept = __popc(now);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) kept += __shfl_down_sync(0xffffffffu, kept, o);
    if (lane == 0) wsum[warp] = kept;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; w++) t += wsum[w];
        tile_counts[blockIdx.x] = t;
    }
}

// exclusive scan of the tile counts (a few to ~10^5 entries): one block; every thread sums a contiguous run, the 1024 run
// totals are scanned with warp shuffles, then every thread writes its run's offsets
__global__ void __launch_bounds__(1024) k_scan_tiles(const uint32_t* __restrict__ counts, uint64_t* __restrict__ offsets, int64_t n_tiles,
                                                      Counters* __restrict__ ctr) {
    __shared__ uint64_t wsum[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t per = (n_tiles + 1023) / 1024;
    const int64_t a = (int64_t)threadIdx.x * per, b = a + per < n_tiles ? a + per : n_tiles;
    uint64_t run = 0;
    for (int64_t i = a; i < b; i++) run += counts[i];
    uint64_t inc = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint64_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint64_t v = wsum[lane], w = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t y = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += y;
        }
        wsum[lane] = w - v; // exclusive over warps
        if (lane == 31) ctr->rows_selected = w;
    }
    __syncthreads();
    uint64_t pos = wsum[warp] + inc - run;
    for (int64_t i = a; i < b; i++) {
        offsets[i] = pos;
        pos += counts[i];
    }
}

struct CompactCol {
    const uint8_t* src; // already offset to the batch base
    uint8_t* dst;       // already offset to the output base of this batch
    const uint8_t* src_null;
    uint8_t* dst_null;
    int32_t width;
    int32_t _pad;
};
struct CompactArgs {
    int32_t n_cols;
    int32_t _pad;
    CompactCol cols[MAX_COLS];
};

// One column of a warp's 32 mask words (1024 rows): lane l holds word l's bits and its first output position. Column outside,
// words inside: the element width is a compile-time constant of the loop, eight words' loads are in flight per lane before the
// first store (source and destination never alias), a set bit's lane reads its row and writes it at the word's position plus
// the number of set bits below it — rows keep their order, like Chunk::filter (column_helper.h:468-577).
template <class T>
SRB_D void compact_words(const T* __restrict__ src, T* __restrict__ dst, uint32_t my_bits, uint32_t my_pos, int lane) {
    const uint32_t below = (1u << lane) - 1u;
    constexpr int U = sizeof(T) > 8 ? 4 : 8;
#pragma unroll 1
    for (int w0 = 0; w0 < 32; w0 += U) {
        T v[U] = {};
        uint32_t at[U];
        bool take[U];
#pragma unroll
        for (int u = 0; u < U; u++) { // all U loads are issued before the first store needs its value
            const uint32_t bits = __shfl_sync(0xffffffffu, my_bits, w0 + u);
            const uint32_t pos = __shfl_sync(0xffffffffu, my_pos, w0 + u);
            take[u] = (bits >> lane) & 1u;
            at[u] = pos + __popc(bits & below);
            if (take[u]) v[u] = src[(w0 + u) * 32 + lane];
        }
#pragma unroll
        for (int u = 0; u < U; u++)
            if (take[u]) dst[at[u]] = v[u];
    }
}

// A sparse tile (at most a quarter of its rows kept: the scan behind a selective runtime filter) is copied through the list of its
// kept row numbers instead: work in proportion to the kept rows, consecutive threads write consecutive output rows.
template <class T>
SRB_D void compact_list(const T* __restrict__ src, T* __restrict__ dst, const uint16_t* sidx, uint32_t total) {
    for (uint32_t j0 = 0; j0 < total; j0 += 256 * 4) {
        T v[4] = {};
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t j = j0 + u * 256 + threadIdx.x;
            if (j < total) v[u] = src[sidx[j]];
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t j = j0 + u * 256 + threadIdx.x;
            if (j < total) dst[j] = v[u];
        }
    }
}

__global__ void __launch_bounds__(256, 4) k_compact(const __grid_constant__ CompactArgs ca, int64_t n, const uint32_t* __restrict__ mask,
                                                  const uint64_t* __restrict__ tile_offsets) {
    __shared__ uint32_t wtot[8];
    __shared__ uint16_t sidx[FTILE / 4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile0 = (int64_t)blockIdx.x * FTILE;
    const int64_t w0 = tile0 >> 5;
    const int64_t n_words = (n + 31) >> 5;
    // rows past n inside the last word were never set (the mask kernels write `i < n && pass`)
    const uint32_t word = (w0 + threadIdx.x) < n_words ? mask[w0 + threadIdx.x] : 0;
    const uint32_t c = __popc(word);
    uint32_t inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const uint32_t t = wtot[w];
        before += w < warp ? t : 0;
        total += t;
    }
    const uint32_t my_pos = before + inc - c; // first output position of this thread's word, relative to the tile
    if (total == 0) return;
    const uint64_t out0 = tile_offsets[blockIdx.x];
    if (total <= FTILE / 4) { // block-uniform
        uint32_t off = my_pos, bits = word;
        while (bits) {
            sidx[off++] = (uint16_t)(threadIdx.x * 32 + __ffs(bits) - 1);
            bits &= bits - 1;
        }
        __syncthreads();
        if (total * (uint32_t)ca.n_cols <= 1024u) {
            // a handful of rows (a star join's scan keeps 1-2 % of a tile): one (column, row) per thread, so the loads of ALL columns are in
            // flight together instead of one column's round trip after the other
            const uint32_t work = total * (uint32_t)ca.n_cols;
            for (uint32_t i = threadIdx.x; i < work; i += 256) {
                const uint32_t cidx = i / total, j = i - cidx * total;
                const CompactCol& cc = ca.cols[cidx];
                const uint32_t r = sidx[j];
                switch (cc.width) {
                case 1: ((uint8_t*)cc.dst)[out0 + j] = ((const uint8_t*)cc.src)[tile0 + r]; break;
                case 2: ((uint16_t*)cc.dst)[out0 + j] = ((const uint16_t*)cc.src)[tile0 + r]; break;
                case 4: ((uint32_t*)cc.dst)[out0 + j] = ((const uint32_t*)cc.src)[tile0 + r]; break;
                case 8: ((uint64_t*)cc.dst)[out0 + j] = ((const uint64_t*)cc.src)[tile0 + r]; break;
                default: ((ulonglong2*)cc.dst)[out0 + j] = ((const ulonglong2*)cc.src)[tile0 + r];
                }
            }
            for (int cidx = 0; cidx < ca.n_cols; cidx++)
                if (ca.cols[cidx].src_null) compact_list(ca.cols[cidx].src_null + tile0, ca.cols[cidx].dst_null + out0, sidx, total);
            return;
        }
        for (int cidx = 0; cidx < ca.n_cols; cidx++) {
            const CompactCol& cc = ca.cols[cidx];
            switch (cc.width) {
            case 1: compact_list((const uint8_t*)cc.src + tile0, (uint8_t*)cc.dst + out0, sidx, total); break;
            case 2: compact_list((const uint16_t*)cc.src + tile0, (uint16_t*)cc.dst + out0, sidx, total); break;
            case 4: compact_list((const uint32_t*)cc.src + tile0, (uint32_t*)cc.dst + out0, sidx, total); break;
            case 8: compact_list((const uint64_t*)cc.src + tile0, (uint64_t*)cc.dst + out0, sidx, total); break;
            default: compact_list((const ulonglong2*)cc.src + tile0, (ulonglong2*)cc.dst + out0, sidx, total);
            }
            if (cc.src_null) compact_list(cc.src_null + tile0, cc.dst_null + out0, sidx, total);
        }
        return;
    }
    if (__ballot_sync(0xffffffffu, word != 0) == 0) return;
    const int64_t row0 = tile0 + (int64_t)warp * 1024;
    for (int cidx = 0; cidx < ca.n_cols; cidx++) {
        const CompactCol& cc = ca.cols[cidx];
        switch (cc.width) {
        case 1: compact_words((const uint8_t*)cc.src + row0, (uint8_t*)cc.dst + out0, word, my_pos, lane); break;
        case 2: compact_words((const uint16_t*)cc.src + row0, (uint16_t*)cc.dst + out0, word, my_pos, lane); break;
        case 4: compact_words((const uint32_t*)cc.src + row0, (uint32_t*)cc.dst + out0, word, my_pos, lane); break;
        case 8: compact_words((const uint64_t*)cc.src + row0, (uint64_t*)cc.dst + out0, word, my_pos, lane); break;
        default: compact_words((const ulonglong2*)cc.src + row0, (ulonglong2*)cc.dst + out0, word, my_pos, lane);
        }
        if (cc.src_null) compact_words(cc.src_null + row0, cc.dst_null + out0, word, my_pos, lane);
    }
}

// packed 16-byte group keys -> one device column per (fixed-width) key, for device-resident result chunks
struct UnpackKey {
    uint8_t* dst;
    uint8_t* dst_null;
    int32_t width;
    int32_t nullable;
};
struct UnpackArgs {
    int32_t n_keys;
    int32_t _pad;
    UnpackKey keys[MAX_KEYS];
};
__global__ void __launch_bounds__(256) k_unpack_keys(const __grid_constant__ UnpackArgs ua, const uint64_t* __restrict__ packed, int rec_words, int64_t n) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t* k = packed + (size_t)r * rec_words;
        int off = 0;
        for (int i = 0; i < ua.n_keys; i++) {
            const UnpackKey& u = ua.keys[i];
            auto byte_at = [&](int o) -> uint64_t { return (k[o >> 3] >> (8 * (o & 7))) & 0xffULL; };
            if (u.nullable) u.dst_null[r] = (uint8_t)byte_at(off++);
            uint64_t lo = 0, hi = 0;
            for (int b = 0; b < u.width; b++) {
                uint64_t v = byte_at(off + b);
                if (b < 8)
                    lo |= v << (8 * b);
                else
                    hi |= v << (8 * (b - 8));
            }
            off += u.width;
            store_w(u.dst, r, u.width, lo, hi);
        }
    }
}

// =================================================================================================================
// host operator
// =================================================================================================================
struct HostCol { // result column handed out by pull_chunk (host memory owned by the op)
    int type = 0, precision = 0, scale = 0;
    std::vector<uint8_t> data, nulls;
    std::vector<uint32_t> offsets;
    bool has_nulls = false;
};

// SRB_TRACE=1: host-side wall time of the stages of a push / set_finishing on stderr (a diagnostic, off by default)
struct HostTrace {
    bool on = enabled();
    static bool enabled() { // read once per process: a trace object is made per push / launch
        static const bool e = getenv("SRB_TRACE") != nullptr;
        return e;
    }
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    void mark(const char* what) {
        if (!on) return;
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[srb trace] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

struct Stage { // device staging of host chunks: one buffer per bound column
    DevBuf data[MAX_COLS], nulls[MAX_COLS], offsets[MAX_COLS];
    DevBuf enc[MAX_COLS], pdesc[MAX_COLS], fdir[MAX_COLS]; // encoded pages: their bytes, page descriptors and frame directory
    std::vector<PageDesc> hdesc[MAX_COLS];                   // host copies of the page descriptors (alive until the stage is reused)
    int64_t rows = 0;
    int64_t bin_bytes[MAX_COLS] = {0};
    cudaEvent_t copied = nullptr;
};

class ScanAggOp : public srb_op {
public:
    ScanAggOp(srb_ctx* c, const srb_scan_agg_plan* plan) : srb_op(c), plan_(plan) {
        filter_only_ = plan_.aggs.empty() && plan_.key_nodes.empty();
        staging_rows_ = c->cfg.staging_rows > 0 ? c->cfg.staging_rows : (int64_t)4 << 20;
        chunk_size_ = c->cfg.chunk_size > 0 ? c->cfg.chunk_size : 4096;
        ctr_dev_.ensure(c, sizeof(Counters));
        ctr_host_.ensure(sizeof(Counters));
        SRB_CUDA(cudaMemsetAsync(ctr_dev_.p, 0, sizeof(Counters), stream));
    }
    ~ScanAggOp() override {
        cudaStreamSynchronize(stream);
        if (copy_stream_) {
            cudaStreamSynchronize(copy_stream_);
            ctx->give_stream(copy_stream_);
        }
        for (auto& s : stage_) ctx->give_event(s.copied, false);
        ctx->give_event(kernel_done_, false);
        ctx->give_event(ev0_, true);
        ctx->give_event(ev1_, true);
    }

    void push_chunk(const srb_chunk* ch) override {
        if (finishing) throw Error(SRB_INTERNAL_ERROR, "push_chunk after set_finishing");
        if (cancelled) throw Error(SRB_CANCELLED, "operator cancelled");
        if (!ch || ch->n_rows < 0) throw Error(SRB_INVALID_ARGUMENT, "bad chunk");
        if (ch->n_rows == 0) return; // empty chunks are dropped by the driver (pipeline_driver.cpp:340-342)
        if (ch->n_rows >= ((int64_t)1 << 32)) throw Error(SRB_INVALID_ARGUMENT, "chunk exceeds Column::MAX_CAPACITY_LIMIT rows");
        if (!lowered_) lower(ch);
        const srb_column* cols[MAX_COLS];
        bool any_dev = false, any_host = false;
        for (size_t i = 0; i < L_.cols.size(); i++) {
            const srb_column* c = find_slot(ch, L_.cols[i].slot);
            if (!c) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d not in chunk", L_.cols[i].slot));
            if (c->type != L_.cols[i].type) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d changed type", L_.cols[i].slot));
            if (c->nulls && !L_.cols[i].nullable)
                throw Error(SRB_NOT_IMPLEMENTED, fmt("slot %d: null map appeared after the first chunk had none", L_.cols[i].slot));
            if (!c->data && !is_binary(c->type)) throw Error(SRB_INVALID_ARGUMENT, "null data pointer");
            (c->mem_kind == SRB_MEM_DEVICE ? any_dev : any_host) = true;
            cols[i] = c;
        }
        if (any_dev && any_host) throw Error(SRB_NOT_IMPLEMENTED, "chunk mixes device and host columns");
        rows_in += ch->n_rows;
        if (any_dev || L_.cols.empty()) {
            flush_stage(); // keep row order between staged host rows and this device range
            ColBind b[MAX_COLS];
            for (size_t i = 0; i < L_.cols.size(); i++) {
                b[i] = L_.prog.cols[i];
                b[i].data = cols[i]->data;
                b[i].nulls = L_.cols[i].nullable ? cols[i]->nulls : nullptr;
                b[i].offsets = cols[i]->offsets;
            }
            process_device_range(b, ch->n_rows);
        } else {
            append_host(cols, ch->n_rows);
        }
    }

    // Encoded data pages of one row range -> the current stage's column buffers, decoded on the device; then the stage is launched
    // like a stage of host rows. The copy of this call's pages overlaps the kernels of the previous call's rows.
    void push_pages(int32_t n_cols, const int32_t* slot_ids, const srb_page_column* pc, int64_t n_rows) override {
        if (finishing) throw Error(SRB_INTERNAL_ERROR, "push_pages after set_finishing");
        if (cancelled) throw Error(SRB_CANCELLED, "operator cancelled");
        if (n_cols < 0 || (n_cols && (!slot_ids || !pc)) || n_rows < 0) throw Error(SRB_INVALID_ARGUMENT, "bad page set");
        if (n_rows == 0) return;
        if (n_rows >= ((int64_t)1 << 31)) throw Error(SRB_INVALID_ARGUMENT, "more than 2^31 rows in one page set");
        if (!lowered_) { // the plan is lowered against the page set's column types (fixed width, NOT NULL)
            std::vector<srb_column> desc((size_t)n_cols);
            for (int i = 0; i < n_cols; i++) {
                desc[(size_t)i] = srb_column{};
                desc[(size_t)i].type = pc[i].type;
                desc[(size_t)i].precision = pc[i].precision;
                desc[(size_t)i].scale = pc[i].scale;
            }
            srb_chunk first{};
            first.n_rows = n_rows;
            first.n_cols = n_cols;
            first.slot_ids = slot_ids;
            first.cols = desc.data();
            lower(&first);
        }
        if (stage_[cur_stage_].rows) flush_stage(); // host rows pushed earlier keep their place in front of these
        HostTrace ptr_;
        Stage& st = stage_[cur_stage_];
        if (!copy_stream_ && !(copy_stream_ = ctx->take_stream())) throw Error(SRB_INTERNAL_ERROR, "cudaStreamCreate failed");
        if (!st.copied && !(st.copied = ctx->take_event(false))) throw Error(SRB_INTERNAL_ERROR, "cudaEventCreate failed");
        if (!page_err_.p) {
            page_err_.ensure(ctx, 8);
            SRB_CUDA(cudaMemsetAsync(page_err_.p, 0, 8, stream));
        }
        struct Work {
            int col, width, enc;
            int64_t frames, n_pages;
            const uint8_t* bytes;
        };
        std::vector<Work> work;
        for (size_t i = 0; i < L_.cols.size(); i++) {
            int k = -1;
            for (int j = 0; j < n_cols; j++)
                if (slot_ids[j] == L_.cols[i].slot) k = j;
            if (k < 0) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d not in the page set", L_.cols[i].slot));
            const srb_page_column& c = pc[k];
            const int w = type_width(c.type);
            if (c.type != L_.cols[i].type) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d changed type", L_.cols[i].slot));
            if (w == 0 || L_.cols[i].nullable) throw Error(SRB_NOT_IMPLEMENTED, "encoded pages of binary or nullable columns");
            if (c.encoding != SRB_ENC_FOR && c.encoding != SRB_ENC_PLAIN) throw Error(SRB_NOT_IMPLEMENTED, fmt("page encoding %d", c.encoding));
            if (c.encoding == SRB_ENC_FOR && c.type == SRB_TYPE_DOUBLE) throw Error(SRB_INVALID_ARGUMENT, "frame-of-reference pages hold integers");
            if (!c.data || c.bytes < 0 || c.n_pages < 0 || (c.n_pages && !c.pages)) throw Error(SRB_INVALID_ARGUMENT, "bad page column");
            // page descriptors: row positions and frame numbers are prefix sums over the pages
            std::vector<PageDesc>& hdesc_ = st.hdesc[i];
            hdesc_.resize((size_t)c.n_pages);
            int64_t rows = 0, frames = 0, lo = c.bytes, hi = 0; // [lo, hi): the bytes this page set touches (a host copy moves only those)
            for (int p = 0; p < c.n_pages; p++) {
                const srb_page& g = c.pages[p];
                if (g.offset < 0 || g.bytes < 0 || g.n_rows < 0 || g.offset + g.bytes > c.bytes || g.bytes >= ((int64_t)1 << 32))
                    throw Error(SRB_INVALID_ARGUMENT, "page outside its column's bytes");
                lo = std::min(lo, g.offset);
                hi = std::max(hi, g.offset + g.bytes);
            }
            if (hi < lo) lo = hi = 0;
            const int64_t rebase = c.mem_kind != SRB_MEM_DEVICE ? lo : 0;
            for (int p = 0; p < c.n_pages; p++) {
                const srb_page& g = c.pages[p];
                hdesc_[(size_t)p] = PageDesc{(uint64_t)(g.offset - rebase), (uint64_t)rows, (uint32_t)g.bytes, (uint32_t)g.n_rows, (uint32_t)frames, 0u};
                rows += g.n_rows;
                frames += (g.n_rows + 127) / 128;
            }
            if (rows != n_rows) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d: pages hold %lld rows, the page set %lld", L_.cols[i].slot, (long long)rows, (long long)n_rows));
            st.data[i].ensure(ctx, (size_t)n_rows * w);
            const uint8_t* bytes = (const uint8_t*)c.data;
            if (c.mem_kind != SRB_MEM_DEVICE) {
                // decoders may read a few bytes past a frame: the copy is padded. Sized like the decoded rows (an encoded page set is smaller,
                // bar a few pathological frames), so that page sets of slightly different encoded sizes reuse the buffer instead of
                // reallocating it (cudaFree + cudaMalloc in the middle of the copy queue: 0.2 ms per page set)
                st.enc[i].ensure(ctx, std::max((size_t)(hi - lo) + 64, (size_t)n_rows * w + 4096));
                if (hi > lo) SRB_CUDA(cudaMemcpyAsync(st.enc[i].p, (const uint8_t*)c.data + lo, (size_t)(hi - lo), cudaMemcpyHostToDevice, copy_stream_));
                bytes = st.enc[i].as<uint8_t>();
            }
            st.pdesc[i].ensure(ctx, std::max<size_t>(1, (size_t)c.n_pages) * sizeof(PageDesc));
            if (c.n_pages)
                SRB_CUDA(cudaMemcpyAsync(st.pdesc[i].p, hdesc_.data(), (size_t)c.n_pages * sizeof(PageDesc), cudaMemcpyHostToDevice, copy_stream_));
            if (c.encoding == SRB_ENC_FOR) st.fdir[i].ensure(ctx, std::max<size_t>(1, (size_t)frames) * sizeof(FrameDesc));
            work.push_back(Work{(int)i, w, c.encoding, frames, c.n_pages, bytes});
        }
        SRB_CUDA(cudaEventRecord(st.copied, copy_stream_));
        SRB_CUDA(cudaStreamWaitEvent(stream, st.copied, 0));
        ptr_.mark("push_pages: copies enqueued");
        // The previous page set's batch is checked NOW, while only its own kernels are on the compute stream: waiting after this set's
        // decode kernels were enqueued made the host wait for this set's copy as well, so the next set's copy was never queued behind it
        // (one decode time of idle PCIe per page set)
        wait_pending();
        ptr_.mark("push_pages: previous batch");
        unsigned int* err = page_err_.as<unsigned int>();
        for (const Work& wk : work) {
            if (wk.n_pages == 0) continue;
            uint8_t* out = st.data[wk.col].as<uint8_t>();
            const PageDesc* pd = st.pdesc[wk.col].as<PageDesc>();
            if (wk.enc == SRB_ENC_PLAIN) {
                k_plain_decode<<<dim3(64, (unsigned)wk.n_pages), 256, 0, stream>>>(wk.bytes, pd, wk.width, out, err);
                launches++;
                continue;
            }
            FrameDesc* fd = st.fdir[wk.col].as<FrameDesc>();
            k_page_dir<<<(int)wk.n_pages, 256, 0, stream>>>(wk.bytes, pd, wk.width, fd, err);
            const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((wk.frames + 7) / 8, (int64_t)ctx->n_sms * 16));
            switch (wk.width) {
            case 1: k_for_decode<1><<<grid, 256, 0, stream>>>(wk.bytes, fd, wk.frames, out); break;
            case 2: k_for_decode<2><<<grid, 256, 0, stream>>>(wk.bytes, fd, wk.frames, out); break;
            case 4: k_for_decode<4><<<grid, 256, 0, stream>>>(wk.bytes, fd, wk.frames, out); break;
            case 8: k_for_decode<8><<<grid, 256, 0, stream>>>(wk.bytes, fd, wk.frames, out); break;
            default: k_for_decode<16><<<grid, 256, 0, stream>>>(wk.bytes, fd, wk.frames, out);
            }
            launches += 2;
        }
        SRB_CUDA(cudaGetLastError());
        rows_in += n_rows;
        st.rows = n_rows;
        pages_pushed_ = true;
        ptr_.mark("push_pages: decodes enqueued");
        flush_stage(); // checks the previous batch, launches this one, switches stages
        ptr_.mark("push_pages: flush_stage");
        // (corrupt pages are reported by set_finishing: asking the device here would stall the copy queue behind this call's own copies)
    }
    // the decode kernels flag footers / frame sizes that do not fit their page (Status::Corruption in FrameOfReferencePageDecoder::init)
    void check_page_errors(bool wait) {
        if (!pages_pushed_) return;
        if (wait) SRB_CUDA(cudaStreamSynchronize(stream));
        unsigned int e[2] = {0, 0};
        SRB_CUDA(cudaMemcpyAsync(e, page_err_.p, 8, cudaMemcpyDeviceToHost, copy_stream_));
        SRB_CUDA(cudaStreamSynchronize(copy_stream_));
        if (e[0]) throw Error(SRB_CORRUPTION, e[0] & PAGE_ERR_FOOTER ? "corrupt data page: footer / header does not match the page" : "corrupt data page: frames do not fit the page");
    }

    // Operator::set_runtime_bloom_filters / eval_runtime_bloom_filters (pipeline/operator.h:185): a scan-filter operator narrows its
    // selection with the join's runtime filter on column `slot` before compacting. Attach before the first chunk is pushed.
    void add_runtime_filter(const srb_runtime_filter* rf, int slot) override {
        if (!rf) throw Error(SRB_INVALID_ARGUMENT, "null runtime filter");
        if (!filter_only_) throw Error(SRB_NOT_IMPLEMENTED, "runtime filters attach to scan-filter operators (plans without aggregates)");
        if (lowered_) throw Error(SRB_INTERNAL_ERROR, "runtime filters must be attached before the first chunk is pushed");
        rfs_.push_back(RfBind{rf, slot, -1});
    }

    void set_finishing() override {
        if (finishing) return;
        finishing = true;
        if (cancelled) {
            finished = true;
            return;
        }
        flush_stage();
        wait_pending();
        if (!lowered_) { // no chunk ever arrived: lower against an empty description so that result types exist
            lower(nullptr);
        }
        check_page_errors(true);
        produce_output();
    }
    // set_finishing on every rank at once + the gather exchange + the merge phase, on the device (srb_op_set_finishing_merged)
    void set_finishing_merged(srb_comm* comm) override {
        if (!comm) throw Error(SRB_INVALID_ARGUMENT, "null communicator");
        if (finishing) throw Error(SRB_INTERNAL_ERROR, "set_finishing_merged after set_finishing");
        if (filter_only_ || stream_mode_ >= 0) throw Error(SRB_INVALID_ARGUMENT, "set_finishing_merged is for blocking aggregate operators");
        if (plan_.phase != SRB_AGG_UPDATE_FINALIZE) throw Error(SRB_INVALID_ARGUMENT, "set_finishing_merged finalises: the plan's phase must be UPDATE_FINALIZE");
        finishing = true;
        if (cancelled) {
            finished = true;
            return;
        }
        flush_stage();
        wait_pending();
        if (!lowered_) lower(nullptr);
        check_page_errors(true);
        if (radix_rows_ > 0) radix_ingest();
        ensure_table();
        const DProg& P = L_.prog;
        if (P.key_words > 2) throw Error(SRB_NOT_IMPLEMENTED, "set_finishing_merged with group keys beyond 16 bytes: gather the serialized states instead");
        const int W = comm_world(comm), me = comm_rank(comm);
        // room for the peers' groups (every rank of a low-cardinality plan holds about the same few groups; a full table fails the merge)
        if (P.n_keys && (occupied_ + 16) * 4 > cap_ && cap_ < (uint64_t)MERGE_MAX_SLOTS) grow_table(std::min<uint64_t>((occupied_ + 16) * 4, MERGE_MAX_SLOTS));
        const size_t region = 16 + (size_t)MERGE_MAX_SLOTS * (size_t)P.slot_bytes;
        merge_send_.ensure(ctx, region);
        merge_recv_.ensure(ctx, region * (size_t)W);
        rx_ctr_.ensure(ctx, sizeof(RxCounters));
        rx_hctr_.ensure(sizeof(RxCounters));
        SRB_CUDA(cudaMemsetAsync(rx_ctr_.p, 0, sizeof(RxCounters), stream));
        k_merge_pack<<<8, 256, 0, stream>>>(slots_.as<uint8_t>(), cap_, P.slot_bytes, merge_send_.as<uint8_t>());
        comm_all_gather_device(comm, merge_send_.p, merge_recv_.p, region, stream);
        RxCounters* dc = rx_ctr_.as<RxCounters>();
        k_merge_tables<<<1, 256, 0, stream>>>(P, view(), merge_recv_.as<uint8_t>(), region, W, me, &dc->groups, &dc->overflow);
        launches += 2;
        SRB_CUDA(cudaGetLastError());
        // emit straight away: the number of merged-in groups arrives with the emit's own counters (one host wait for both)
        RxCounters* hc = rx_hctr_.as<RxCounters>();
        SRB_CUDA(cudaMemcpyAsync(hc, rx_ctr_.p, sizeof(RxCounters), cudaMemcpyDeviceToHost, stream));
        produce_output(P.n_keys ? (uint64_t)std::min<uint64_t>(cap_, occupied_ + (uint64_t)(W - 1) * MERGE_MAX_SLOTS) : 1, false);
        if (hc->overflow) throw Error(SRB_NOT_IMPLEMENTED, "set_finishing_merged: a rank's group table is too large (or full) for the state-level gather");
        occupied_ += hc->groups;
    }
    bool has_output() override { return finishing && !finished && out_pos_ < out_rows_; }
    bool need_input() override { return !finishing; }
    bool is_finished() override { return finished || (finishing && out_pos_ >= out_rows_); }

    srb_status pull_chunk(srb_chunk* out, bool all) override {
        if (cancelled) throw Error(SRB_CANCELLED, "operator cancelled");
        if (!finishing) throw Error(SRB_INTERNAL_ERROR, "pull_chunk before set_finishing (blocking aggregate has no output yet)");
        if (out_pos_ >= out_rows_ && !(out_rows_ == 0 && !pulled_empty_ && all)) {
            finished = true;
            return SRB_END_OF_FILE;
        }
        int64_t n = out_rows_ - out_pos_;
        if (!all && n > chunk_size_) n = chunk_size_;
        materialize(out_pos_, n);
        out_pos_ += n;
        pulled_empty_ = true;
        fill_chunk(out, n);
        if (out_pos_ >= out_rows_) finished = true;
        return SRB_OK;
    }

    void refresh_counters() override { rows_selected = (int64_t)selected_total_; }
    void profile(double* kernel_ms, int64_t* batches, int64_t* static_launches, int32_t* shape) override {
        *kernel_ms = kernel_ms_;
        *batches = timed_batches_;
        *static_launches = static_launches_;
        *shape = jit_launches_ ? SHAPE_JIT : shape_;
    }

    // all remaining rows as device-resident columns (operator chaining on the GPU)
    srb_status pull_device(srb_chunk* out) override {
        if (cancelled) throw Error(SRB_CANCELLED, "operator cancelled");
        if (!finishing) throw Error(SRB_INTERNAL_ERROR, "pull before set_finishing");
        if (out_pos_ >= out_rows_ && pulled_empty_) {
            finished = true;
            return SRB_END_OF_FILE;
        }
        const int64_t n = out_rows_ - out_pos_;
        hslots_.clear();
        cview_.clear();
        if (filter_only_) {
            for (size_t k = 0; k < out_col_idx_.size(); k++) {
                const ColMeta& cm = L_.cols[out_col_idx_[k]];
                srb_column c{};
                c.type = cm.type;
                c.precision = cm.precision;
                c.scale = cm.scale;
                c.mem_kind = SRB_MEM_DEVICE;
                int w = type_width(cm.type);
                c.data = k < fout_.size() && fout_[k].data.p ? fout_[k].data.as<uint8_t>() + (size_t)out_pos_ * w : nullptr;
                c.nulls = cm.nullable && k < fout_.size() && fout_[k].nulls.p ? fout_[k].nulls.as<uint8_t>() + out_pos_ : nullptr;
                cview_.push_back(c);
                hslots_.push_back(cm.slot);
            }
        } else {
            device_result_columns(n);
        }
        out_pos_ = out_rows_;
        pulled_empty_ = true;
        finished = true;
        out->n_rows = n;
        out->n_cols = (int32_t)cview_.size();
        out->slot_ids = hslots_.data();
        out->cols = cview_.data();
        return SRB_OK;
    }

private:
    // aggregate result as device columns: keys unpacked by a kernel, aggregate values as emitted
    void device_result_columns(int64_t n);

    // ---- lowering / table ------------------------------------------------------------------------------------
    void lower(const srb_chunk* first) {
        srb_chunk empty{};
        std::vector<srb_column> fake;
        std::vector<int32_t> fake_slots;
        if (!first) {
            // synthesize column descriptions from the plan's own SLOT_REF / predicate types
            auto add = [&](int slot, int type, int prec, int scale) {
                for (int s : fake_slots)
                    if (s == slot) return;
                srb_column c{};
                c.type = type;
                c.precision = prec;
                c.scale = scale;
                fake.push_back(c);
                fake_slots.push_back(slot);
            };
            for (auto& e : plan_.nodes)
                if (e.op == SRB_EXPR_SLOT_REF) add(e.slot_id, e.type, e.precision, e.scale);
            for (auto& p : plan_.preds) add(p.slot_id, p.type, 0, 0);
            for (int s : plan_.out_slots) add(s, SRB_TYPE_BIGINT, 0, 0);
            empty.n_cols = (int)fake.size();
            empty.slot_ids = fake_slots.data();
            empty.cols = fake.data();
            first = &empty;
        }
        L_ = lower_plan(plan_, first);
        if (filter_only_) {
            for (int s : plan_.out_slots) {
                bool found = false;
                for (size_t i = 0; i < L_.cols.size(); i++) found |= L_.cols[i].slot == s;
                if (!found) {
                    const srb_column* c = find_slot(first, s);
                    if (!c) throw Error(SRB_INVALID_ARGUMENT, fmt("output slot %d not in chunk", s));
                    if ((int)L_.cols.size() >= MAX_COLS) throw Error(SRB_NOT_IMPLEMENTED, "too many columns");
                    int idx = (int)L_.cols.size();
                    L_.cols.push_back({s, c->type, c->precision, c->scale, c->nulls != nullptr});
                    L_.prog.cols[idx].width = type_width(c->type);
                    L_.prog.cols[idx].kind = is_binary(c->type) ? CK_BINARY : (c->type == SRB_TYPE_DOUBLE ? CK_DOUBLE : CK_INT);
                    L_.prog.n_cols = idx + 1;
                }
            }
            for (auto& r : rfs_) { // the key column of every runtime filter is bound like an output column
                for (size_t i = 0; i < L_.cols.size(); i++)
                    if (L_.cols[i].slot == r.slot) r.col = (int)i;
                if (r.col < 0) {
                    const srb_column* c = find_slot(first, r.slot);
                    if (!c) throw Error(SRB_INVALID_ARGUMENT, fmt("runtime filter slot %d not in chunk", r.slot));
                    if ((int)L_.cols.size() >= MAX_COLS) throw Error(SRB_NOT_IMPLEMENTED, "too many columns");
                    r.col = (int)L_.cols.size();
                    L_.cols.push_back({r.slot, c->type, c->precision, c->scale, c->nulls != nullptr});
                    L_.prog.cols[r.col].width = type_width(c->type);
                    L_.prog.cols[r.col].kind = CK_INT;
                    L_.prog.n_cols = r.col + 1;
                }
                const int t = L_.cols[r.col].type;
                if (is_binary(t) || t == SRB_TYPE_DOUBLE || type_width(t) > 8 || (r.rf->key_width && type_width(t) != r.rf->key_width))
                    throw Error(SRB_INVALID_ARGUMENT, fmt("runtime filter on slot %d: column type does not match the build key", r.slot));
            }
            for (int s : plan_.out_slots)
                for (size_t i = 0; i < L_.cols.size(); i++)
                    if (L_.cols[i].slot == s) {
                        if (is_binary(L_.cols[i].type)) throw Error(SRB_NOT_IMPLEMENTED, "filter output of binary columns");
                        out_col_idx_.push_back((int)i);
                    }
        } else {
            init_table();
            shape_ = getenv("SRB_NO_BUILTIN_SHAPES") ? -1 : match_static_shape(L_.prog); // hook: A/B of a built-in shape against its run-time compiled twin
            // no built-in shape: can the static kernel be compiled for this plan at run time (srb_jit.h)?
            // Two variants: [0] for batches without null maps (or whose maps hold no NULL: they are stripped), [1] with the maps of the
            // columns that arrived nullable compiled in.
            if (shape_ < 0 && jit_mode_ && !force_generic_) {
                bool nullable[MAX_COLS] = {};
                bool any = false;
                for (size_t i = 0; i < L_.cols.size() && i < (size_t)MAX_COLS; i++) any = (nullable[i] = L_.cols[i].nullable) || any;
                const int hint = (int)std::min<int64_t>(plan_.group_hint, 1 << 30);
                JitShape js;
                if (jit_shape_of(L_.prog, hint, nullptr, js)) {
                    jit_src_[0] = jit_source(js);
                    jit_gmax_[0] = js.gmax;
                }
                if (any && jit_shape_of(L_.prog, hint, nullable, js)) {
                    jit_src_[1] = jit_source(js);
                    jit_gmax_[1] = js.gmax;
                }
                for (int v = 0; v < 2; v++)
                    if (!jit_src_[v].empty()) jit_[v] = JitCache::get().peek(jit_src_[v]); // an earlier operator may have compiled it already
            }
        }
        lowered_ = true;
    }

    void init_table() {
        const DProg& P = L_.prog;
        if (P.n_keys == 0) {
            cap_ = 1;
        } else {
            uint64_t want = plan_.group_hint > 0 ? 64 : 1024; // a hinted small table stays small enough for the interpreter's shared mirror
            if (plan_.group_hint > 0) {
                uint64_t h = (uint64_t)plan_.group_hint * 2;
                while (want < h) want <<= 1;
            }
            cap_ = want;
        }
        occupied_ = 0;
        table_ready_ = false;
        // a large hinted table (gigabytes to allocate and clear) is only built when a batch goes through it: the radix path does not
        if ((size_t)cap_ * (size_t)P.slot_bytes < ((size_t)64 << 20)) ensure_table();
    }
    void ensure_table() {
        if (table_ready_) return;
        alloc_table(slots_, cap_);
        if (L_.prog.n_keys == 0) { // the single state block of a no-group-by aggregation: zero-filled states are valid, mark it taken
            static const uint64_t ready = SLOT_READY;
            SRB_CUDA(cudaMemcpyAsync(slots_.p, &ready, 8, cudaMemcpyHostToDevice, stream));
        }
        table_ready_ = true;
    }
    void alloc_table(DevBuf& buf, uint64_t cap) {
        size_t bytes = (size_t)cap * (size_t)L_.prog.slot_bytes;
        buf.ensure(ctx, bytes);
        SRB_CUDA(cudaMemsetAsync(buf.p, 0, bytes, stream));
    }
    TableView view() {
        TableView t;
        t.slots = slots_.as<uint8_t>();
        t.mask = cap_ - 1;
        t.slot_bytes = L_.prog.slot_bytes;
        t.key_words = L_.prog.key_words;
        return t;
    }
    void grow_table(uint64_t min_cap) {
        if (!table_ready_) { // nothing to rehash: size the first allocation directly
            while (cap_ < min_cap) cap_ <<= 1;
            ensure_table();
            return;
        }
        uint64_t ncap = cap_ * 2;
        while (ncap < min_cap) ncap <<= 1;
        DevBuf nb;
        alloc_table(nb, ncap);
        TableView nt;
        nt.slots = nb.as<uint8_t>();
        nt.mask = ncap - 1;
        nt.slot_bytes = L_.prog.slot_bytes;
        nt.key_words = L_.prog.key_words;
        int blocks = (int)std::min<uint64_t>((cap_ + 255) / 256, (uint64_t)ctx->n_sms * 16);
        k_rehash<<<blocks, 256, 0, stream>>>(L_.prog, slots_.as<uint8_t>(), cap_, nt);
        launches++;
        SRB_CUDA(cudaStreamSynchronize(stream));
        std::swap(slots_.p, nb.p);
        std::swap(slots_.bytes, nb.bytes);
        nb.ctx = ctx;
        slots_.ctx = ctx;
        cap_ = ncap;
    }

    // ---- batches ---------------------------------------------------------------------------------------------
    static constexpr int64_t BATCH = (int64_t)1 << 31;     // filter batches: one mask + count pass, one exactly sized growth, one compaction
    static constexpr int64_t AGG_BATCH = (int64_t)1 << 30; // aggregation batches: the host waits for the counters once per batch (a static launch covers <= 2^30 rows)

    void process_device_range(const ColBind* cols, int64_t n) {
        if (filter_only_) {
            filter_range(cols, n);
            return;
        }
        if (stream_mode_ >= 0) {
            stream_range(cols, n);
            return;
        }
        // The L2-blocked path wants the LARGEST batch memory allows: the table region that is live at any moment is
        // (rows in flight / rows per batch) x table bytes, so more rows per batch means a smaller, L2-resident window and
        // more updates per table sector while it is cached.
        if (radix_applies(n) && radix_batch(cols, n)) return;
        int64_t batch = AGG_BATCH;
        if (hc_partition_applies(n)) {
            size_t free_b = 0, total_b = 0;
            cudaMemGetInfo(&free_b, &total_b);
            free_b += (size_t)ctx->pool_bytes; // pooled buffers are reusable
            batch = (int64_t)1 << 30;
            while (batch > AGG_BATCH && (size_t)std::min<int64_t>(batch, n) * 17 + ((size_t)1 << 30) > free_b) batch >>= 1;
        }
        HostTrace tr;
        for (int64_t off = 0; off < n; off += batch) {
            int64_t m = std::min<int64_t>(batch, n - off);
            launch_batch(cols, off, m);
            tr.mark("launch_batch");
            check_batch(cols, off, m);
            tr.mark("check_batch");
        }
    }

    // the interpreter's shared-memory mirror of a small table (see k_scan_agg_generic): every state but 128-bit MIN / MAX
    int generic_shared_cap() const {
        const DProg& P = L_.prog;
        if (!P.n_keys || no_shared_agg_ || cap_ * (uint64_t)P.n_aggs > (uint64_t)GEN_SH_ENTRIES) return 0;
        for (int a = 0; a < P.n_aggs; a++) {
            const DAgg& g = P.aggs[a];
            if ((g.fn == FN_MIN || g.fn == FN_MAX) && g.wide && !g.is_double) return 0; // 128-bit MIN / MAX has no shared-memory atomic
            if (g.merge && g.fn == FN_AVG && (g.is_double || g.avg_as_double)) return 0;
        }
        return (int)cap_;
    }

    // directory entries of the thread-private low-cardinality kernel (k_scan_agg_lowcard), 0 = not applicable: a GROUP BY whose table is
    // small (hinted: at most 64 slots), a batch large enough to amortise the per-CTA merge, and room for >= 2 entries per CTA
    int lowcard_entries(int64_t n) const {
        const DProg& P = L_.prog;
        if (!P.n_keys || P.key_words > 2 || no_lowcard_ || plan_.group_hint <= 0 || n < ((int64_t)1 << 18) || P.n_aggs < 1) return 0;
        const size_t per_entry = (size_t)P.n_aggs * 3 * LC_THREADS * 8;
        const int fit = std::min((int)((size_t)(110 * 1024) / per_entry), LC_MAXG); // two CTAs per SM
        // the planner expects no more groups than a CTA's directory holds (more groups: the shared mirror of k_scan_agg_generic does better)
        return fit >= 2 && plan_.group_hint <= fit ? fit : 0;
    }

    void bind_prog(const ColBind* cols) {
        for (int i = 0; i < L_.prog.n_cols; i++) {
            L_.prog.cols[i].data = cols[i].data;
            L_.prog.cols[i].nulls = cols[i].nulls;
            L_.prog.cols[i].offsets = cols[i].offsets;
        }
    }

    int grid_for(int64_t n) const {
        int64_t blocks = (n + 255) / 256;
        int64_t maxb = (int64_t)ctx->n_sms * 8;
        return (int)std::max<int64_t>(1, std::min(blocks, maxb));
    }

    // ---- radix-partitioned aggregation (srb_radix_agg.cuh): first batch of a high-cardinality k -> SUM, COUNT plan -------------
    bool hc_cols_ok() const {
        const DProg& P = L_.prog;
        return shape_ == 2 ? (P.n_cols == 2 && P.nodes[0].col != P.nodes[1].col)
                           : (shape_ == 4 && P.n_cols == 3 && P.nodes[0].col != P.nodes[1].col && P.nodes[0].col != P.nodes[2].col &&
                              P.nodes[1].col != P.nodes[2].col);
    }
    bool radix_applies(int64_t n) const {
        const DProg& P = L_.prog;
        if (no_radix_ || force_generic_ || stream_mode_ >= 0 || P.n_preds != 0 || !hc_cols_ok()) return false;
        if (occupied_ != 0 || stream_rows_ != 0 || radix_rows_ != 0) return false; // the first batch only: its groups must be the only ones
        if (n < radix_min_rows_ || n > ((int64_t)1 << 30)) return false;
        for (int i = 0; i < P.n_cols; i++)
            if (P.cols[i].width != 8) return false;
        // the planner's cardinality estimate says "many groups" (at most ~64 rows per key): with few, hot keys the partitions would overflow
        return plan_.group_hint > 0 && (int64_t)plan_.group_hint >= n / 64;
    }
    template <int NV>
    void radix_launch(const ColBind* cols, int64_t n, int fbits, uint32_t cap, const RxOut& out) {
        const DProg& P = L_.prog;
        const uint64_t* k = (const uint64_t*)cols[P.nodes[0].col].data;
        const uint64_t* v0 = (const uint64_t*)cols[P.nodes[1].col].data;
        const uint64_t* v1 = NV == 2 ? (const uint64_t*)cols[P.nodes[2].col].data : nullptr;
        static SmemOptIn opt;
        opt.set(k_rx_aggregate<NV>, (int)RxGeom<NV>::BYTES);
        k_rx_scatter<NV><<<ctx->n_sms * 8, 256, 0, stream>>>(k, v0, v1, n, P.phase_seed, fbits, cap, rx_cursor_.as<uint32_t>(), rx_a_.as<ulonglong2>(),
                                                            rx_b_.as<uint64_t>(), rx_ctr_.as<RxCounters>());
        k_rx_aggregate<NV><<<ctx->n_sms, RX_THREADS, RxGeom<NV>::BYTES, stream>>>(rx_a_.as<ulonglong2>(), rx_b_.as<uint64_t>(), rx_cursor_.as<uint32_t>(),
                                                                                   1 << fbits, cap, P.phase_seed, fbits, out, rx_ctr_.as<RxCounters>());
        launches += 2;
    }
    // -> false when a partition (a hot key) or the output overflowed: nothing is kept and the caller takes the table path
    bool radix_batch(const ColBind* cols, int64_t n) {
        const DProg& P = L_.prog;
        for (int i = 0; i < P.n_cols; i++)
            if (cols[i].nulls) return false; // (null maps without a set byte are stripped by the table path's launch_batch only)
        const int nv = shape_ == 4 ? 2 : 1;
        const int cap_max = nv == 1 ? RxGeom<1>::CAP : RxGeom<2>::CAP;
        // Rows of one key land in one partition: with c rows per key a partition's row count has variance mean * (c + 1), not mean.
        // The finest split whose mean + 6 sigma + slack still fits a partition buffer:
        const double c = std::max(1.0, (double)n / (double)plan_.group_hint);
        auto cap_for = [&](int fb) {
            const double m = (double)n / (double)((int64_t)1 << fb);
            return m + 6.0 * std::sqrt(m * (c + 1.0)) + 64.0;
        };
        int fbits = 8;
        while (cap_for(fbits) > (double)cap_max && fbits < 21) fbits++;
        if (cap_for(fbits) > (double)cap_max) return false;
        const int64_t F = (int64_t)1 << fbits;
        const uint32_t cap = std::min<uint32_t>(((uint32_t)cap_for(fbits) + 63u) & ~63u, (uint32_t)cap_max);
        rx_cursor_.ensure(ctx, (size_t)F * 4);
        rx_a_.ensure(ctx, (size_t)F * cap * 16);
        if (nv == 2) rx_b_.ensure(ctx, (size_t)F * cap * 8);
        rx_ctr_.ensure(ctx, sizeof(RxCounters));
        rx_hctr_.ensure(sizeof(RxCounters));
        SRB_CUDA(cudaMemsetAsync(rx_cursor_.p, 0, (size_t)F * 4, stream));
        SRB_CUDA(cudaMemsetAsync(rx_ctr_.p, 0, sizeof(RxCounters), stream));
        const uint64_t out_cap = std::min<uint64_t>((uint64_t)n, std::max<uint64_t>((uint64_t)plan_.group_hint * 2, (uint64_t)1 << 20));
        EmitArgs ea = emit_args(out_cap);
        RxOut out{ea.keys, (uint64_t*)ea.aggs[0].val, ea.aggs[0].null, (uint64_t*)ea.aggs[1].val, ea.aggs[1].null, 0ULL, out_cap};
        if (!ev0_) {
            ev0_ = ctx->take_event(true);
            ev1_ = ctx->take_event(true);
            if (!ev0_ || !ev1_) throw Error(SRB_INTERNAL_ERROR, "cudaEventCreate failed");
        }
        SRB_CUDA(cudaEventRecord(ev0_, stream));
        if (nv == 1)
            radix_launch<1>(cols, n, fbits, cap, out);
        else
            radix_launch<2>(cols, n, fbits, cap, out);
        SRB_CUDA(cudaEventRecord(ev1_, stream));
        SRB_CUDA(cudaGetLastError());
        RxCounters* hc = rx_hctr_.as<RxCounters>();
        SRB_CUDA(cudaMemcpyAsync(hc, rx_ctr_.p, sizeof(RxCounters), cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        float ms = 0;
        SRB_CUDA(cudaEventElapsedTime(&ms, ev0_, ev1_));
        kernel_ms_ += ms;
        timed_batches_++;
        if (hc->overflow) return false;
        radix_rows_ = (int64_t)hc->groups;
        stream_rows_ = radix_rows_; // the groups sit in front of whatever the table emits later, like pass-through rows
        selected_total_ += (uint64_t)n;
        radix_batches_++;
        return true;
    }
    void radix_ingest() {
        const DProg& P = L_.prog;
        const uint64_t need = (uint64_t)radix_rows_ * 2 + 1024;
        if (!table_ready_ || cap_ < need) grow_table(need);
        ensure_table();
        SRB_CUDA(cudaMemsetAsync(rx_ctr_.p, 0, sizeof(RxCounters), stream));
        RxCounters* dc = rx_ctr_.as<RxCounters>();
        k_rx_ingest<<<grid_for(radix_rows_), 256, 0, stream>>>(P, view(), ekeys_.as<ulonglong2>(), eval_[0].val.as<uint64_t>(), eval_[1].val.as<uint64_t>(),
                                                               radix_rows_, &dc->groups, &dc->overflow);
        launches++;
        RxCounters* hc = rx_hctr_.as<RxCounters>();
        SRB_CUDA(cudaMemcpyAsync(hc, rx_ctr_.p, sizeof(RxCounters), cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        if (hc->overflow) throw Error(SRB_INTERNAL_ERROR, "radix groups did not fit the group table");
        occupied_ += hc->groups;
        stream_rows_ = 0;
        radix_rows_ = 0;
    }

    // ---- L2-blocked high-cardinality path (srb_hcpart.cuh) ----------------------------------------------------------------
    // worth it when the table is far larger than L2 and the batch is large enough to amortise three extra launches
    bool hc_partition_applies(int64_t n) const {
        const DProg& P = L_.prog;
        const bool cols_ok = shape_ == 2 ? (P.n_cols == 2 && P.nodes[0].col != P.nodes[1].col)
                                         : (shape_ == 4 && P.n_cols == 3 && P.nodes[0].col != P.nodes[1].col && P.nodes[0].col != P.nodes[2].col &&
                                            P.nodes[1].col != P.nodes[2].col);
        return cols_ok && !force_generic_ && !no_partition_ && P.n_preds == 0 && (size_t)cap_ * (size_t)P.slot_bytes >= ((size_t)256 << 20) &&
               cap_ >= ((uint64_t)1 << 16) && n >= ((int64_t)1 << 22);
    }
    void hc_partition(const ColBind* cols, int64_t off, int64_t n) {
        const DProg& P = L_.prog;
        const int nv = shape_ == 4 ? 2 : 1; // phase 2 carries the partial sum and the partial count
        const int kc = P.nodes[0].col, vc = P.nodes[1].col, vc2 = nv == 2 ? P.nodes[2].col : -1;
        hc_k_.ensure(ctx, (size_t)n * 8);
        hc_v_.ensure(ctx, (size_t)n * 8);
        if (nv == 2) hc_v2_.ensure(ctx, (size_t)n * 8);
        if (!hc_ctr_.p) {
            hc_ctr_.ensure(ctx, 2 * HC_PARTS * sizeof(unsigned long long));
            SRB_CUDA(cudaMemsetAsync(hc_ctr_.p, 0, 2 * HC_PARTS * sizeof(unsigned long long), stream));
        }
        unsigned long long* hist = hc_ctr_.as<unsigned long long>();
        unsigned long long* cursor = hist + HC_PARTS;
        int log2cap = 0;
        while (((uint64_t)1 << log2cap) < cap_) log2cap++;
        const int shift = log2cap - HC_PBITS;
        const uint64_t* keys = (const uint64_t*)cols[kc].data + off;
        HcVals hv{};
        hv.src[0] = (const uint64_t*)cols[vc].data + off;
        hv.dst[0] = hc_v_.as<uint64_t>();
        if (nv == 2) {
            hv.src[1] = (const uint64_t*)cols[vc2].data + off;
            hv.dst[1] = hc_v2_.as<uint64_t>();
        }
        static SmemOptIn opt1, opt2; // per (kernel, device)
        opt1.set(k_hc_scatter<1>, (int)sizeof(HcSmem<1>));
        opt2.set(k_hc_scatter<2>, (int)sizeof(HcSmem<2>));
        k_hc_hist<<<ctx->n_sms * 2, 1024, 0, stream>>>(keys, n, P.phase_seed, cap_ - 1, shift, hist);
        k_hc_scan<<<1, HC_PARTS, 0, stream>>>(hist, cursor);
        if (nv == 1)
            k_hc_scatter<1><<<ctx->n_sms, HC_THREADS, sizeof(HcSmem<1>), stream>>>(keys, hv, n, P.phase_seed, cap_ - 1, shift, cursor, hc_k_.as<uint64_t>());
        else
            k_hc_scatter<2><<<ctx->n_sms, HC_THREADS, sizeof(HcSmem<2>), stream>>>(keys, hv, n, P.phase_seed, cap_ - 1, shift, cursor, hc_k_.as<uint64_t>());
        launches += 3;
        for (int i = 0; i < P.n_cols; i++) hc_cols_[i] = cols[i];
        hc_cols_[kc].data = hc_k_.p;
        hc_cols_[vc].data = hc_v_.p;
        if (nv == 2) hc_cols_[vc2].data = hc_v2_.p;
    }

    // ---- streaming pre-aggregation sink ------------------------------------------------------------------------
    // AggregateStreamingSinkOperator::push_chunk dispatches on TStreamingPreaggregationMode; one pushed range (at most AGG_BATCH
    // rows) plays the part of one chunk in the reference's AggrAutoState machine. What the reference decides from a dry-run lookup
    // of the chunk's keys (ADJUST: build_hash_map_with_selection, then one of three treatments of the whole chunk) is decided
    // here from the hit count of the selective pass itself — found rows update their groups, the others pass through — so the
    // decision costs no extra pass over the rows; the counters steer the following ranges exactly as the reference's do.
    // Thresholds that stand for CPU cache sizes in the reference (STREAMING_HT_MIN_REDUCTION: 256 KB / 2 MB; MaxHtSize 64 MB) are
    // restated for a B200: a table that fits the 126 MB L2 resolves its atomics there, a larger one pays HBM sectors per row.
    enum RowMode { ROWS_AGGREGATE = STREAM_ROWS_AGGREGATE, ROWS_STREAM = STREAM_ROWS_PASS, ROWS_SELECTIVE = STREAM_ROWS_SELECTIVE };

    void set_streaming_mode(int mode) override {
        if (mode < SRB_STREAMING_AUTO || mode > SRB_STREAMING_LIMITED_MEM) throw Error(SRB_INVALID_ARGUMENT, "unknown streaming pre-aggregation mode");
        if (rows_in > 0 || finishing) throw Error(SRB_INVALID_ARGUMENT, "the streaming mode is set before the first push_chunk");
        if (plan_.key_nodes.empty() || plan_.aggs.empty() || plan_.phase != SRB_AGG_UPDATE_SERIALIZE)
            throw Error(SRB_INVALID_ARGUMENT, "streaming pre-aggregation needs GROUP BY keys and the update + serialize phase");
        stream_mode_ = mode;
        if (getenv("SRB_STREAM_L2_BYTES")) policy_.l2_bytes = (size_t)atoll(getenv("SRB_STREAM_L2_BYTES")); // test hook: a small "L2"
    }
    void set_streaming_memory_limit(int64_t bytes) override {
        if (bytes < 0) throw Error(SRB_INVALID_ARGUMENT, "negative memory limit");
        stream_mem_limit_ = bytes;
    }
    void streaming_stats(int64_t* rows_passed_through, int64_t* groups_in_table, int32_t* auto_state) override {
        if (rows_passed_through) *rows_passed_through = stream_rows_;
        if (groups_in_table) *groups_in_table = (int64_t)occupied_;
        if (auto_state) *auto_state = (int32_t)policy_.state;
    }

    // AUTO learns from the hit counts of successive units: 2^24 rows each (a few per cent of an SF100 table), so that a poor
    // reduction is noticed after the first unit instead of after a whole 2^29-row batch. SRB_STREAM_UNIT: test hook.
    void stream_range(const ColBind* cols, int64_t n) {
        ensure_table();
        const int64_t unit = (stream_mode_ == SRB_STREAMING_AUTO || stream_mode_ == SRB_STREAMING_LIMITED_MEM) ? stream_unit_ : AGG_BATCH;
        for (int64_t off = 0; off < n; off += unit) stream_batch(cols, off, std::min<int64_t>(unit, n - off));
    }

    // rows of the range that found their group already present (aggregate: selected - newly inserted; selective: counted by the kernel)
    int64_t run_rows(const ColBind* cols, int64_t off, int64_t n, RowMode mode) {
        if (mode == ROWS_AGGREGATE) {
            const uint64_t sel0 = selected_total_, occ0 = occupied_;
            launch_batch(cols, off, n);
            check_batch(cols, off, n);
            return (int64_t)(selected_total_ - sel0) - (int64_t)(occupied_ - occ0);
        }
        const DProg& P = L_.prog;
        EmitArgs ea = emit_args((uint64_t)stream_rows_ + (uint64_t)n); // every row of the range may pass through
        bind_prog(cols);
        if (!ev0_) {
            ev0_ = ctx->take_event(true);
            ev1_ = ctx->take_event(true);
            if (!ev0_ || !ev1_) throw Error(SRB_INTERNAL_ERROR, "cudaEventCreate failed");
        }
        SRB_CUDA(cudaMemsetAsync(ctr_dev_.p, 0, sizeof(Counters), stream));
        SRB_CUDA(cudaEventRecord(ev0_, stream));
        k_stream_rows<<<grid_for(n), 256, 0, stream>>>(P, view(), off, n, mode == ROWS_SELECTIVE ? 1 : 0, ea, stream_rows_, ctr_dev_.as<Counters>());
        launches++;
        SRB_CUDA(cudaEventRecord(ev1_, stream));
        SRB_CUDA(cudaGetLastError());
        Counters* hc = ctr_host_.as<Counters>();
        SRB_CUDA(cudaMemcpyAsync(hc, ctr_dev_.p, sizeof(Counters), cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaMemsetAsync(ctr_dev_.p, 0, sizeof(Counters), stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        float ms = 0;
        SRB_CUDA(cudaEventElapsedTime(&ms, ev0_, ev1_));
        kernel_ms_ += ms;
        timed_batches_++;
        if (hc->key_too_long) throw Error(SRB_NOT_IMPLEMENTED, "serialized group key longer than its key area (16 bytes, or the declared 17-64)");
        selected_total_ += hc->rows_selected;
        stream_rows_ += (int64_t)hc->emitted;
        return (int64_t)hc->inserted; // the stream kernel reports its hits there
    }

    void stream_batch(const ColBind* cols, int64_t off, int64_t n) {
        const int64_t rows_before = stream_seen_;
        stream_seen_ += n;
        if (stream_mode_ == SRB_STREAMING_FORCE_STREAMING) {
            run_rows(cols, off, n, ROWS_STREAM);
            return;
        }
        if (stream_mode_ == SRB_STREAMING_FORCE_PREAGGREGATION) {
            run_rows(cols, off, n, ROWS_AGGREGATE);
            return;
        }
        const size_t slot_bytes = (size_t)L_.prog.slot_bytes;
        // LIMITED_MEM (_push_chunk_by_limited_memory): once the table has reached the limit every further row passes through
        if (stream_mode_ == SRB_STREAMING_LIMITED_MEM && stream_mem_limit_ > 0 && (int64_t)((size_t)cap_ * slot_bytes) >= stream_mem_limit_) {
            run_rows(cols, off, n, ROWS_STREAM);
            return;
        }
        // AUTO: srb_stream_policy.h
        const StreamRowMode mode = policy_.next(n, occupied_, cap_, (size_t)cap_ * slot_bytes, rows_before, stream_rows_);
        const uint64_t sel0 = selected_total_;
        const int64_t hits = run_rows(cols, off, n, (RowMode)mode);
        policy_.observe(hits, (int64_t)(selected_total_ - sel0), (size_t)cap_ * slot_bytes);
    }

    const ColBind* strip_clear_null_maps(const ColBind* cols, int64_t off, int64_t n) {
        const DProg& P = L_.prog;
        bool any_map = false;
        for (int i = 0; i < P.n_cols; i++) any_map = any_map || cols[i].nulls;
        if (!any_map) return cols;
        Counters* dc = ctr_dev_.as<Counters>();
        for (int i = 0; i < P.n_cols; i++)
            if (cols[i].nulls) {
                k_null_any<<<grid_for(n / 16 + 1), 256, 0, stream>>>(cols[i].nulls + off, n, &dc->any_null);
                launches++;
            }
        Counters* hc = ctr_host_.as<Counters>();
        SRB_CUDA(cudaMemcpyAsync(&hc->any_null, &dc->any_null, sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
        SRB_CUDA(cudaMemsetAsync(&dc->any_null, 0, sizeof(unsigned long long), stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        if (hc->any_null) {
            batch_has_nulls_ = true;
            return cols;
        }
        for (int i = 0; i < P.n_cols; i++) {
            nn_cols_[i] = cols[i];
            nn_cols_[i].nulls = nullptr;
        }
        batch_stripped_ = true;
        return nn_cols_;
    }

    void launch_batch(const ColBind* cols, int64_t off, int64_t n, const uint32_t* defer_in = nullptr) {
        if (radix_rows_ > 0) radix_ingest(); // a key may occur in the radix batch's groups and in this batch: both go through the table
        HostTrace tr;
        ensure_table();
        tr.mark("  ensure_table");
        if (!ev0_) {
            ev0_ = ctx->take_event(true);
            ev1_ = ctx->take_event(true);
            if (!ev0_ || !ev1_) throw Error(SRB_INTERNAL_ERROR, "cudaEventCreate failed");
        }
        SRB_CUDA(cudaEventRecord(ev0_, stream)); // device time of this batch's kernels, read in check_batch
        if (!defer_in) {
            // a recognised shape whose columns arrive with null maps: maps without a set byte are dropped for this batch, so that
            // the static kernel (and the L2-blocked permutation, which moves values but not null maps) can take it
            batch_has_nulls_ = batch_stripped_ = false;
            jit_kernel_ = nullptr;
            bool maps = false;
            for (int i = 0; i < L_.prog.n_cols; i++) maps = maps || cols[i].nulls;
            // a shape compiled at run time: start the compilation at the first large batch, use it once it is ready
            auto jit_variant = [&](int v) -> const JitLoaded* {
                if (jit_src_[v].empty() || jit_failed_[v]) return nullptr;
                if (L_.prog.n_keys) {
                    // a table-path shape sends every row to the table: with a few dozen groups the atomics pile up on the same addresses
                    // (SSB Q4.1's 35 groups: 2.1 ms for 9.5 M rows), where the interpreter's shared-memory mirror takes 0.1 ms. And a
                    // directory shape whose directory the table has outgrown (the planner's estimate was wrong) would do the same.
                    if (jit_gmax_[v] == 0 && generic_shared_cap() > 0) return nullptr;
                    if (jit_gmax_[v] > 0 && occupied_ > (uint64_t)jit_gmax_[v]) return nullptr;
                }
                if (!jit_[v] && n >= jit_min_rows_) jit_[v] = JitCache::get().acquire(jit_src_[v], jit_mode_ == 2);
                if (!jit_[v]) return nullptr;
                const JitLoaded* k = jit_[v]->get(ctx->cfg.device);
                if (!k && jit_[v]->failed()) {
                    jit_failed_[v] = true; // the interpreter serves the operator; the reason is kept for srb_jit_stats / SRB_TRACE
                    if (getenv("SRB_TRACE")) fprintf(stderr, "[srb] run-time compilation failed: %s\n", jit_[v]->error().c_str());
                }
                return k;
            };
            if (!maps) jit_kernel_ = jit_variant(0);
            if (maps && !force_generic_ && (shape_ >= 0 || !jit_src_[0].empty())) {
                cols = strip_clear_null_maps(cols, off, n); // maps without a set byte are dropped: the NULL-free kernel takes the batch
                jit_kernel_ = jit_variant(batch_has_nulls_ ? 1 : 0);
            }
            tr.mark("  strip null maps");
            batch_partitioned_ = !batch_has_nulls_ && hc_partition_applies(n);
            if (batch_partitioned_) hc_partition(cols, off, n);
            tr.mark("  hc_partition");
        } else if (!batch_partitioned_ && batch_stripped_) {
            cols = nn_cols_; // the replay sees the batch as its first launch did
        }
        if (batch_partitioned_) { // the batch (and a replay of its deferred rows) reads the permuted copy
            cols = hc_cols_;
            off = 0;
        }
        bind_prog(cols);
        if (L_.prog.n_keys) {
            size_t words = (size_t)((n + 31) / 32);
            for (int i = 0; i < 2; i++)
                if (defer_[i].bytes < words * 4) {
                    // ensure() rounds the size up and the pool hands back dirty memory: clear ALL of it, since a later, larger batch
                    // that still fits uses words past this batch's count without coming through here again
                    defer_[i].ensure(ctx, words * 4);
                    SRB_CUDA(cudaMemsetAsync(defer_[i].p, 0, defer_[i].bytes, stream));
                }
        }
        TableView tv = view();
        uint32_t* dout = L_.prog.n_keys ? defer_[defer_cur_].as<uint32_t>() : nullptr;
        // recognised shapes run the TMA-staged static kernel over the whole tiles; the generic kernel takes the rest
        int64_t covered = 0;
        Counters* dc = ctr_dev_.as<Counters>();
        if ((shape_ >= 0 || jit_kernel_) && !defer_in && !force_generic_) {
            covered = jit_kernel_ ? jit_launch(*jit_kernel_, L_.prog, tv, off, n, dout, &dc->rows_selected, &dc->deferred, &dc->inserted, ctx->n_sms, stream, !no_narrow_)
                                  : launch_static_shape(shape_, L_.prog, tv, off, n, dout, &dc->rows_selected, &dc->deferred, &dc->inserted, ctx->n_sms, stream,
                                                        !no_narrow_);
            if (covered) {
                launches++;
                static_launches_++;
                jit_launches_ += jit_kernel_ ? 1 : 0;
            }
        }
        if (covered < n) {
            const int lc_entries = defer_in ? 0 : lowcard_entries(n - covered);
            if (lc_entries) {
                const size_t smem = (size_t)lc_entries * L_.prog.n_aggs * 3 * LC_THREADS * 8;
                static SmemOptIn opt;
                opt.set(k_scan_agg_lowcard, 112 * 1024);
                const int64_t rest = n - covered;
                const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((rest + LC_THREADS - 1) / LC_THREADS, (int64_t)ctx->n_sms * 2));
                k_scan_agg_lowcard<<<blocks, LC_THREADS, smem, stream>>>(L_.prog, tv, off + covered, rest, covered, dout, dc, lc_entries);
            } else {
                const int sh_cap = generic_shared_cap();
                k_scan_agg_generic<<<grid_for(n - covered), 256, (size_t)sh_cap * L_.prog.n_aggs * 24, stream>>>(L_.prog, tv, off + covered, n - covered, covered,
                                                                                                                  defer_in, dout, dc, sh_cap);
            }
            launches++;
        }
        SRB_CUDA(cudaEventRecord(ev1_, stream));
        timed_pending_ = true;
        SRB_CUDA(cudaGetLastError());
        tr.mark("  kernels enqueued");
    }

    // waits for the batch, folds the counters, replays deferred rows after growing the table
    void check_batch(const ColBind* cols, int64_t off, int64_t n) {
        for (;;) {
            Counters* hc = ctr_host_.as<Counters>();
            SRB_CUDA(cudaMemcpyAsync(hc, ctr_dev_.p, sizeof(Counters), cudaMemcpyDeviceToHost, stream));
            SRB_CUDA(cudaMemsetAsync(ctr_dev_.p, 0, sizeof(Counters), stream));
            SRB_CUDA(cudaStreamSynchronize(stream));
            if (timed_pending_) {
                float ms = 0;
                SRB_CUDA(cudaEventElapsedTime(&ms, ev0_, ev1_));
                kernel_ms_ += ms;
                timed_batches_++;
                timed_pending_ = false;
            }
            selected_total_ += hc->rows_selected;
            if (hc->key_too_long) throw Error(SRB_NOT_IMPLEMENTED, "serialized group key longer than 16 bytes");
            occupied_ += hc->inserted;
            uint64_t deferred = hc->deferred;
            // deferred counts rows, not groups: bound one growth step so a burst of duplicates cannot over-allocate
            if (L_.prog.n_keys && (deferred || occupied_ * 2 > cap_))
                grow_table(std::min<uint64_t>(std::max<uint64_t>(cap_ * 2, (occupied_ + deferred) * 4), std::max<uint64_t>(cap_ * 8, occupied_ * 4)));
            if (!deferred) return;
            // replay: the bitmap just written becomes the input, the other one (cleared) the output
            int in = defer_cur_;
            defer_cur_ ^= 1;
            SRB_CUDA(cudaMemsetAsync(defer_[defer_cur_].p, 0, defer_[defer_cur_].bytes, stream));
            launch_batch(cols, off, n, defer_[in].as<uint32_t>());
            // the consumed bitmap is cleared for its next use as an output
            SRB_CUDA(cudaMemsetAsync(defer_[in].p, 0, defer_[in].bytes, stream));
        }
    }

    // ---- host chunk staging ------------------------------------------------------------------------------------
    void append_host(const srb_column* const* cols, int64_t n) {
        int64_t done = 0;
        while (done < n) {
            Stage& st = stage_[cur_stage_];
            if (st.rows == 0) begin_stage(st);
            int64_t take = std::min<int64_t>(n - done, staging_rows_ - st.rows);
            for (size_t i = 0; i < L_.cols.size(); i++) {
                const srb_column* c = cols[i];
                int w = L_.prog.cols[i].width;
                if (w) {
                    SRB_CUDA(cudaMemcpyAsync(st.data[i].as<uint8_t>() + (size_t)st.rows * w, (const uint8_t*)c->data + (size_t)done * w,
                                             (size_t)take * w, cudaMemcpyHostToDevice, copy_stream_));
                } else {
                    // binary: bytes appended, offsets rebased on the host (cheap, sequential) then copied
                    uint32_t b0 = c->offsets[done], b1 = c->offsets[done + take];
                    size_t nbytes = b1 - b0;
                    if ((size_t)st.bin_bytes[i] + nbytes + 16 > st.data[i].bytes) grow_bin(st, (int)i, (size_t)st.bin_bytes[i] + nbytes + 16);
                    if (nbytes)
                        SRB_CUDA(cudaMemcpyAsync(st.data[i].as<uint8_t>() + st.bin_bytes[i], (const uint8_t*)c->data + b0, nbytes,
                                                 cudaMemcpyHostToDevice, copy_stream_));
                    off_scratch_.resize((size_t)take + 1);
                    for (int64_t r = 0; r <= take; r++) off_scratch_[r] = c->offsets[done + r] - b0 + (uint32_t)st.bin_bytes[i];
                    SRB_CUDA(cudaMemcpyAsync(st.offsets[i].as<uint32_t>() + st.rows, off_scratch_.data(), (size_t)(take + 1) * 4,
                                             cudaMemcpyHostToDevice, copy_stream_));
                    SRB_CUDA(cudaStreamSynchronize(copy_stream_)); // off_scratch_ is reused
                    st.bin_bytes[i] += (int64_t)nbytes;
                }
                if (L_.cols[i].nullable) {
                    if (c->nulls)
                        SRB_CUDA(cudaMemcpyAsync(st.nulls[i].as<uint8_t>() + st.rows, c->nulls + done, (size_t)take, cudaMemcpyHostToDevice,
                                                 copy_stream_));
                    else
                        SRB_CUDA(cudaMemsetAsync(st.nulls[i].as<uint8_t>() + st.rows, 0, (size_t)take, copy_stream_));
                }
            }
            st.rows += take;
            done += take;
            if (st.rows >= staging_rows_) flush_stage();
        }
    }
    void begin_stage(Stage& st) {
        if (!copy_stream_ && !(copy_stream_ = ctx->take_stream())) throw Error(SRB_INTERNAL_ERROR, "cudaStreamCreate failed");
        if (!st.copied && !(st.copied = ctx->take_event(false))) throw Error(SRB_INTERNAL_ERROR, "cudaEventCreate failed");
        for (size_t i = 0; i < L_.cols.size(); i++) {
            int w = L_.prog.cols[i].width;
            if (w)
                st.data[i].ensure(ctx, (size_t)staging_rows_ * w);
            else {
                if (!st.data[i].p) st.data[i].ensure(ctx, (size_t)1 << 20);
                st.offsets[i].ensure(ctx, (size_t)(staging_rows_ + 1) * 4);
                st.bin_bytes[i] = 0;
            }
            if (L_.cols[i].nullable) st.nulls[i].ensure(ctx, (size_t)staging_rows_);
        }
    }
    void grow_bin(Stage& st, int i, size_t need) {
        SRB_CUDA(cudaStreamSynchronize(copy_stream_));
        size_t ncap = std::max(need, st.data[i].bytes * 2);
        DevBuf nb;
        nb.ensure(ctx, ncap);
        if (st.bin_bytes[i]) SRB_CUDA(cudaMemcpy(nb.p, st.data[i].p, (size_t)st.bin_bytes[i], cudaMemcpyDeviceToDevice));
        std::swap(st.data[i].p, nb.p);
        std::swap(st.data[i].bytes, nb.bytes);
        nb.ctx = ctx;
        st.data[i].ctx = ctx;
    }
    // launches the kernel over the current stage and switches to the other one; the previous stage's kernel is
    // checked first, so at most one batch is in flight while the host fills the next stage.
    void flush_stage() {
        Stage& st = stage_[cur_stage_];
        if (st.rows == 0) return;
        wait_pending();
        SRB_CUDA(cudaEventRecord(st.copied, copy_stream_));
        SRB_CUDA(cudaStreamWaitEvent(stream, st.copied, 0));
        ColBind* b = pending_cols_;
        for (size_t i = 0; i < L_.cols.size(); i++) {
            b[i] = L_.prog.cols[i];
            b[i].data = st.data[i].p;
            b[i].nulls = L_.cols[i].nullable ? st.nulls[i].as<uint8_t>() : nullptr;
            b[i].offsets = st.offsets[i].as<uint32_t>();
        }
        if (filter_only_) {
            filter_range(b, st.rows);
            st.rows = 0;
        } else if (stream_mode_ >= 0) { // the streaming sink decides per unit and waits for each unit's counters
            stream_range(b, st.rows);
            st.rows = 0;
        } else {
            launch_batch(b, 0, st.rows);
            pending_rows_ = st.rows;
            pending_ = true;
            pending_stage_ = cur_stage_;
        }
        cur_stage_ ^= 1;
    }
    void wait_pending() {
        if (!pending_) return;
        pending_ = false;
        check_batch(pending_cols_, 0, pending_rows_);
        stage_[pending_stage_].rows = 0;
    }

    // ---- filter-only -----------------------------------------------------------------------------------------------
    // One mask + count pass over the whole range, one exactly sized growth of the result columns, one compaction pass: the range
    // is read once (predicate columns by the mask kernel, output columns by the compaction) and the kept rows are written once.
    void filter_range(const ColBind* cols, int64_t n) {
        bind_prog(cols);
        const DProg& P = L_.prog;
        // predicates that lower to ranges over NULL-free, 16-byte aligned 4- / 8-byte columns take the vectorised mask kernel
        FRangeArgs fa{};
        bool fast = !no_fast_filter_;
        if (fast) {
            StaticArgs sa{};
            for (int c = 0; c < MAX_COLS; c++) {
                sa.col_off[c] = c;    // lower_ranges identifies a column by this field
                sa.null_off[c] = -1;  // and refuses columns with a staged null map (the null maps of this path are checked below)
            }
            fast = lower_ranges(P, sa);
            for (int i = 0; fast && i < P.n_preds; i++) fast = cols[P.preds[i].col].nulls == nullptr;
            for (int k = 0; fast && k < sa.n_ranges; k++) {
                const SRange& g = sa.rng[k];
                fa.r[k] = FRange{(const uint8_t*)cols[g.col_off].data, g.width, g.neg, g.lo, g.span};
                fast = ((uintptr_t)fa.r[k].data & 15) == 0;
            }
            fa.n = sa.n_ranges;
        }
        for (int64_t off = 0; off < n; off += BATCH) {
            int64_t m = std::min<int64_t>(BATCH, n - off);
            int64_t tiles = (m + FTILE - 1) / FTILE;
            mask_.ensure(ctx, (size_t)((m + 31) / 32) * 4 + 4);
            tile_counts_.ensure(ctx, (size_t)tiles * 4);
            tile_offsets_.ensure(ctx, (size_t)tiles * 8);
            if (fast) {
                FRangeArgs fb = fa;
                for (int k = 0; k < fb.n; k++) fb.r[k].data += (size_t)off * fb.r[k].width; // BATCH keeps the 16-byte alignment
                if (fb.n == 1 && fb.r[0].width == 4)
                    k_filter_mask_ranges<4><<<(int)tiles, 256, 0, stream>>>(fb, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>());
                else if (fb.n == 1)
                    k_filter_mask_ranges<8><<<(int)tiles, 256, 0, stream>>>(fb, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>());
                else
                    k_filter_mask_ranges<0><<<(int)tiles, 256, 0, stream>>>(fb, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>());
            } else {
                k_filter_mask<<<(int)tiles, 256, 0, stream>>>(L_.prog, off, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>());
            }
            // opt-in (SRB_RF_FUSED=1, not yet measured): every runtime filter of the scan in one pass over the tiles
            bool fused = rf_fused_ && rfs_.size() >= 2 && rfs_.size() <= (size_t)RF_FUSED_MAX;
            for (const auto& r : rfs_) fused = fused && P.cols[r.col].width == P.cols[rfs_[0].col].width && !r.rf->is_bitset;
            if (fused) {
                RfMultiArgs ma{};
                ma.n = (int)rfs_.size();
                const int w = P.cols[rfs_[0].col].width;
                for (int i = 0; i < ma.n; i++) {
                    const int ci = rfs_[i].col;
                    ma.f[i] = RfArgs{rfs_[i].rf->view(), (const uint8_t*)cols[ci].data + (size_t)off * w, cols[ci].nulls ? cols[ci].nulls + off : nullptr};
                }
                switch (w) {
                case 1: k_filter_rf_multi<1><<<(int)tiles, 256, 0, stream>>>(ma, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>()); break;
                case 2: k_filter_rf_multi<2><<<(int)tiles, 256, 0, stream>>>(ma, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>()); break;
                case 4: k_filter_rf_multi<4><<<(int)tiles, 256, 0, stream>>>(ma, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>()); break;
                default: k_filter_rf_multi<8><<<(int)tiles, 256, 0, stream>>>(ma, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>());
                }
                launches++;
            }
            for (const auto& r : rfs_) { // runtime filters narrow the mask before anything is copied
                if (fused) break;
                const int ci = r.col;
                const int w = P.cols[ci].width;
                RfArgs ra{r.rf->view(), (const uint8_t*)cols[ci].data + (size_t)off * w, cols[ci].nulls ? cols[ci].nulls + off : nullptr};
                switch (w) {
                case 1: k_filter_rf<1><<<(int)tiles, 256, 0, stream>>>(ra, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>()); break;
                case 2: k_filter_rf<2><<<(int)tiles, 256, 0, stream>>>(ra, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>()); break;
                case 4: k_filter_rf<4><<<(int)tiles, 256, 0, stream>>>(ra, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>()); break;
                default: k_filter_rf<8><<<(int)tiles, 256, 0, stream>>>(ra, m, mask_.as<uint32_t>(), tile_counts_.as<uint32_t>());
                }
                launches++;
            }
            k_scan_tiles<<<1, 1024, 0, stream>>>(tile_counts_.as<uint32_t>(), tile_offsets_.as<uint64_t>(), tiles, ctr_dev_.as<Counters>());
            launches += 2;
            Counters* hc = ctr_host_.as<Counters>();
            SRB_CUDA(cudaMemcpyAsync(hc, ctr_dev_.p, sizeof(Counters), cudaMemcpyDeviceToHost, stream));
            SRB_CUDA(cudaStreamSynchronize(stream));
            int64_t sel = (int64_t)hc->rows_selected;
            // grow the device result columns, then compact into them
            CompactArgs ca{};
            ca.n_cols = (int)out_col_idx_.size();
            fout_.resize(out_col_idx_.size());
            for (size_t k = 0; k < out_col_idx_.size(); k++) {
                int ci = out_col_idx_[k];
                int w = L_.prog.cols[ci].width;
                grow_out(fout_[k].data, (size_t)out_rows_ * w, (size_t)(out_rows_ + sel) * w);
                if (L_.cols[ci].nullable) grow_out(fout_[k].nulls, (size_t)out_rows_, (size_t)(out_rows_ + sel));
                ca.cols[k].src = (const uint8_t*)cols[ci].data + (size_t)off * w;
                ca.cols[k].dst = fout_[k].data.as<uint8_t>() + (size_t)out_rows_ * w;
                // a later device chunk may come without a null map for a slot that had one before: its rows are all non-NULL
                ca.cols[k].src_null = (L_.cols[ci].nullable && cols[ci].nulls) ? cols[ci].nulls + off : nullptr;
                if (L_.cols[ci].nullable && !cols[ci].nulls && sel > 0)
                    SRB_CUDA(cudaMemsetAsync(fout_[k].nulls.as<uint8_t>() + out_rows_, 0, (size_t)sel, stream));
                ca.cols[k].dst_null = L_.cols[ci].nullable ? fout_[k].nulls.as<uint8_t>() + out_rows_ : nullptr;
                ca.cols[k].width = w;
            }
            if (sel > 0 && ca.n_cols > 0) {
                k_compact<<<(int)tiles, 256, 0, stream>>>(ca, m, mask_.as<uint32_t>(), tile_offsets_.as<uint64_t>());
                launches++;
            }
            SRB_CUDA(cudaGetLastError());
            out_rows_ += sel;
            selected_total_ += (uint64_t)sel;
        }
        SRB_CUDA(cudaStreamSynchronize(stream)); // the caller may reuse the input stage / read the result columns right away
    }
    struct FOut {
        DevBuf data, nulls;
    };
    struct RfBind {
        const srb_runtime_filter* rf;
        int slot, col;
    };
    std::vector<RfBind> rfs_;
    void grow_out(DevBuf& b, size_t used, size_t need) {
        if (need <= b.bytes) return;
        size_t ncap = std::max(need, b.bytes * 2);
        DevBuf nb;
        nb.ensure(ctx, ncap);
        if (used) SRB_CUDA(cudaMemcpyAsync(nb.p, b.p, used, cudaMemcpyDeviceToDevice, stream));
        SRB_CUDA(cudaStreamSynchronize(stream));
        std::swap(b.p, nb.p);
        std::swap(b.bytes, nb.bytes);
        nb.ctx = ctx;
        b.ctx = ctx;
    }

    // ---- output ------------------------------------------------------------------------------------------------------
    // output arrays for `rows` result rows; the first stream_rows_ rows (pass-through rows of a streaming sink) are preserved
    // bytes of one packed key record in the result's key buffer
    size_t key_rec_bytes() const { return L_.prog.key_words > 2 ? (size_t)L_.prog.key_words * 8 : 16; }
    EmitArgs emit_args(uint64_t rows) {
        const DProg& P = L_.prog;
        const size_t cap = (size_t)std::max<uint64_t>(1, rows), keep = (size_t)stream_rows_;
        EmitArgs ea{};
        grow_out(ekeys_, keep * key_rec_bytes(), cap * key_rec_bytes());
        ea.keys = ekeys_.as<ulonglong2>();
        ea.key_rec_words = (int32_t)(key_rec_bytes() / 8);
        eval_.resize(P.n_aggs);
        for (int a = 0; a < P.n_aggs; a++) {
            const AggOutMeta& m = L_.agg_out[a];
            EmitAgg& e = ea.aggs[a];
            grow_out(eval_[a].val, keep * m.out_width, cap * m.out_width);
            e.val = eval_[a].val.p;
            e.out_width = m.out_width;
            if (m.nullable) {
                grow_out(eval_[a].null, keep, cap);
                e.null = eval_[a].null.as<uint8_t>();
            }
            if (m.blob) {
                grow_out(eval_[a].cnt, keep * 8, cap * 8);
                e.cnt = eval_[a].cnt.as<int64_t>();
            }
            e.scale = m.arg_scale;
            e.decode = m.fn == SRB_AGG_MAX ? 1 : (m.fn == SRB_AGG_MIN ? 2 : 0);
            if (m.fn == SRB_AGG_COUNT)
                e.mode = 0;
            else if (m.fn != SRB_AGG_AVG || m.blob)
                e.mode = 1;
            else if (is_decimal(m.arg_type))
                e.mode = 2;
            else if (m.arg_type == SRB_TYPE_DATE)
                e.mode = 3;
            else if (m.arg_type == SRB_TYPE_LARGEINT)
                e.mode = 4;
            else
                e.mode = 5;
        }
        return ea;
    }

    void produce_output() { produce_output(L_.prog.n_keys ? occupied_ : 1, true); }
    // max_rows: an upper bound of the table's groups (exact when `exact`: a mismatch is an internal error)
    void produce_output(uint64_t max_rows, bool exact) {
        if (filter_only_) return; // rows already sit in fout_
        const DProg& P = L_.prog;
        HostTrace tr;
        EmitArgs ea = emit_args((uint64_t)stream_rows_ + max_rows);
        tr.mark("emit_args (output arrays)");
        if (!table_ready_) { // every batch went through the radix path (or none arrived): the result rows are in place already
            if (P.n_keys == 0) {
                ensure_table();
            } else {
                out_rows_ = stream_rows_;
                return;
            }
        }
        SRB_CUDA(cudaMemsetAsync(ctr_dev_.p, 0, sizeof(Counters), stream));
        int blocks = (int)std::min<uint64_t>((cap_ + 255) / 256, (uint64_t)ctx->n_sms * 16);
        k_emit<<<blocks, 256, 0, stream>>>(P, slots_.as<uint8_t>(), cap_, ea, ctr_dev_.as<Counters>(), stream_rows_);
        launches++;
        SRB_CUDA(cudaGetLastError());
        Counters* hc = ctr_host_.as<Counters>();
        SRB_CUDA(cudaMemcpyAsync(hc, ctr_dev_.p, sizeof(Counters), cudaMemcpyDeviceToHost, stream));
        prefetch_small_result((int64_t)stream_rows_ + (int64_t)max_rows); // a handful of rows reach the host with this same wait
        SRB_CUDA(cudaStreamSynchronize(stream));
        if (exact ? hc->emitted != max_rows : hc->emitted > max_rows) throw Error(SRB_INTERNAL_ERROR, fmt("emit produced %llu rows, table holds %llu", (unsigned long long)hc->emitted, (unsigned long long)max_rows));
        tr.mark("k_emit + wait");
        out_rows_ = stream_rows_ + (int64_t)hc->emitted; // pass-through rows first, then the table's groups
    }

    // the device arrays a result row is made of: (base pointer, bytes per row), in the order materialize() consumes them
    struct RowPiece {
        const uint8_t* base;
        size_t unit;
    };
    std::vector<RowPiece> row_pieces() const {
        const DProg& P = L_.prog;
        std::vector<RowPiece> v;
        if (P.n_keys) v.push_back({ekeys_.as<uint8_t>(), key_rec_bytes()});
        for (int a = 0; a < P.n_aggs; a++) {
            const AggOutMeta& m = L_.agg_out[a];
            v.push_back({eval_[a].val.as<uint8_t>(), (size_t)m.out_width});
            if (m.nullable) v.push_back({eval_[a].null.as<uint8_t>(), 1});
            if (m.blob) v.push_back({(const uint8_t*)eval_[a].cnt.as<int64_t>(), 8});
        }
        return v;
    }
    // Results of at most 64 rows (Q1's four, Q6's one) are packed and copied to the host right behind the emit kernel, so that
    // pull_chunk finds them there: one host wait per set_finishing + pull instead of two.
    void prefetch_small_result(int64_t rows) {
        prefetch_rows_ = 0;
        if (rows <= 0 || rows > 64) return;
        const std::vector<RowPiece> rp = row_pieces();
        if ((int)rp.size() > PackArgs::MAX) return;
        PackArgs pa{};
        pa.n = (int)rp.size();
        size_t off = 0;
        prefetch_off_.assign(rp.size(), 0);
        for (int i = 0; i < pa.n; i++) {
            pa.src[i] = rp[(size_t)i].base;
            pa.bytes[i] = (uint32_t)(rp[(size_t)i].unit * (size_t)rows);
            pa.dst_off[i] = (uint32_t)off;
            prefetch_off_[(size_t)i] = off;
            off += (pa.bytes[i] + 15) & ~(size_t)15;
        }
        if (!pa.n) return;
        pack_dev_.ensure(ctx, off);
        pack_host_.ensure(off);
        k_pack_result<<<std::min(64, pa.n * 4), 256, 0, stream>>>(pa, pack_dev_.as<uint8_t>());
        launches++;
        SRB_CUDA(cudaMemcpyAsync(pack_host_.p, pack_dev_.p, off, cudaMemcpyDeviceToHost, stream));
        prefetch_rows_ = rows;
    }

    // copies rows [pos, pos+n) of the device result into host columns
    void materialize(int64_t pos, int64_t n) {
        hcols_.clear();
        hslots_.clear();
        if (filter_only_) {
            for (size_t k = 0; k < out_col_idx_.size(); k++) {
                const ColMeta& cm = L_.cols[out_col_idx_[k]];
                HostCol h;
                h.type = cm.type;
                h.precision = cm.precision;
                h.scale = cm.scale;
                int w = type_width(cm.type);
                h.data.resize((size_t)n * w);
                if (n) SRB_CUDA(cudaMemcpyAsync(h.data.data(), fout_[k].data.as<uint8_t>() + (size_t)pos * w, (size_t)n * w, cudaMemcpyDeviceToHost, stream));
                if (cm.nullable) {
                    h.has_nulls = true;
                    h.nulls.resize((size_t)n);
                    if (n) SRB_CUDA(cudaMemcpyAsync(h.nulls.data(), fout_[k].nulls.as<uint8_t>() + pos, (size_t)n, cudaMemcpyDeviceToHost, stream));
                }
                hcols_.push_back(std::move(h));
                hslots_.push_back(cm.slot);
            }
            SRB_CUDA(cudaStreamSynchronize(stream));
            return;
        }
        const DProg& P = L_.prog;
        // keys: packed records (16 bytes; a wide key's words) -> one column per key
        const size_t krec = key_rec_bytes();
        std::vector<uint8_t> packed;
        if (P.n_keys) packed.resize((size_t)n * krec);
        std::vector<std::vector<uint8_t>> vals(P.n_aggs), nulls(P.n_aggs);
        std::vector<std::vector<int64_t>> cnts(P.n_aggs);
        // every piece of the result rows [pos, pos+n): (device source, bytes, host destination)
        struct Piece {
            const uint8_t* src;
            size_t bytes;
            void* dst;
        };
        std::vector<Piece> pieces;
        if (P.n_keys && n) pieces.push_back({ekeys_.as<uint8_t>() + (size_t)pos * krec, (size_t)n * krec, packed.data()});
        for (int a = 0; a < P.n_aggs; a++) {
            const AggOutMeta& m = L_.agg_out[a];
            vals[a].resize((size_t)n * m.out_width);
            if (n) pieces.push_back({eval_[a].val.as<uint8_t>() + (size_t)pos * m.out_width, (size_t)n * m.out_width, vals[a].data()});
            if (m.nullable) {
                nulls[a].resize((size_t)n);
                if (n) pieces.push_back({eval_[a].null.as<uint8_t>() + pos, (size_t)n, nulls[a].data()});
            }
            if (m.blob) {
                cnts[a].resize((size_t)n);
                if (n) pieces.push_back({(const uint8_t*)(eval_[a].cnt.as<int64_t>() + pos), (size_t)n * 8, cnts[a].data()});
            }
        }
        size_t total = 0;
        for (auto& pc : pieces) total += (pc.bytes + 15) & ~(size_t)15;
        if (n > 0 && prefetch_rows_ > 0 && pos + n <= prefetch_rows_) { // already on the host (prefetch_small_result): same piece order
            const std::vector<RowPiece> rp = row_pieces();
            for (size_t i = 0; i < pieces.size() && i < rp.size(); i++)
                memcpy(pieces[i].dst, pack_host_.as<uint8_t>() + prefetch_off_[i] + (size_t)pos * rp[i].unit, pieces[i].bytes);
        } else if (total <= ((size_t)1 << 20) && (int)pieces.size() <= PackArgs::MAX) { // (<= 1 MB: the pinned buffer comes from the pool)
            // small result: one pack kernel + ONE pinned copy instead of a pageable copy per column
            PackArgs pa{};
            pa.n = (int)pieces.size();
            size_t off = 0;
            for (int i = 0; i < pa.n; i++) {
                pa.src[i] = pieces[i].src;
                pa.bytes[i] = (uint32_t)pieces[i].bytes;
                pa.dst_off[i] = (uint32_t)off;
                off += (pieces[i].bytes + 15) & ~(size_t)15;
            }
            if (pa.n) {
                prefetch_rows_ = 0; // the buffers are reused
                pack_dev_.ensure(ctx, total);
                pack_host_.ensure(total);
                k_pack_result<<<std::min(64, pa.n * 4), 256, 0, stream>>>(pa, pack_dev_.as<uint8_t>());
                launches++;
                SRB_CUDA(cudaMemcpyAsync(pack_host_.p, pack_dev_.p, total, cudaMemcpyDeviceToHost, stream));
            }
            SRB_CUDA(cudaStreamSynchronize(stream));
            for (int i = 0; i < pa.n; i++) memcpy(pieces[i].dst, pack_host_.as<uint8_t>() + pa.dst_off[i], pieces[i].bytes);
        } else {
            for (auto& pc : pieces) SRB_CUDA(cudaMemcpyAsync(pc.dst, pc.src, pc.bytes, cudaMemcpyDeviceToHost, stream));
            SRB_CUDA(cudaStreamSynchronize(stream));
        }
        for (int k = 0; k < P.n_keys; k++) {
            const ColMeta& cm = L_.key_out[k];
            HostCol h;
            h.type = cm.type;
            h.precision = cm.precision;
            h.scale = cm.scale;
            h.has_nulls = cm.nullable;
            if (is_binary(cm.type)) h.offsets.push_back(0);
            hcols_.push_back(std::move(h));
            hslots_.push_back(k);
        }
        for (int64_t r = 0; r < n; r++) {
            const uint8_t* kb = packed.data() + (size_t)r * krec;
            int off = 0;
            for (int k = 0; k < P.n_keys; k++) {
                HostCol& h = hcols_[k];
                const DKey& kd = P.keys[k];
                if (kd.nullable) h.nulls.push_back(kb[off++]);
                if (kd.width == 0) {
                    uint32_t len;
                    memcpy(&len, kb + off, 4);
                    off += 4;
                    h.data.insert(h.data.end(), kb + off, kb + off + len);
                    off += (int)len;
                    h.offsets.push_back((uint32_t)h.data.size());
                } else {
                    h.data.insert(h.data.end(), kb + off, kb + off + kd.width);
                    off += kd.width;
                }
            }
        }
        for (int a = 0; a < P.n_aggs; a++) {
            const AggOutMeta& m = L_.agg_out[a];
            HostCol h;
            h.type = m.out_type;
            h.precision = m.out_precision;
            h.scale = m.out_scale;
            h.has_nulls = m.nullable;
            h.nulls = std::move(nulls[a]);
            if (m.blob) { // {sum, count}: 24 bytes (int128 sum) or 16 bytes (double sum); empty when NULL (avg.h:168-181)
                h.offsets.push_back(0);
                int sw = m.out_width;
                for (int64_t r = 0; r < n; r++) {
                    if (!h.nulls[r]) {
                        h.data.insert(h.data.end(), vals[a].data() + (size_t)r * sw, vals[a].data() + (size_t)(r + 1) * sw);
                        const uint8_t* cp = (const uint8_t*)&cnts[a][r];
                        h.data.insert(h.data.end(), cp, cp + 8);
                    }
                    h.offsets.push_back((uint32_t)h.data.size());
                }
            } else {
                h.data = std::move(vals[a]);
            }
            hcols_.push_back(std::move(h));
            hslots_.push_back(P.n_keys + a);
        }
    }

    void fill_chunk(srb_chunk* out, int64_t n) {
        cview_.resize(hcols_.size());
        for (size_t i = 0; i < hcols_.size(); i++) {
            HostCol& h = hcols_[i];
            srb_column& c = cview_[i];
            c.type = h.type;
            c.precision = h.precision;
            c.scale = h.scale;
            c.mem_kind = SRB_MEM_HOST;
            if (h.data.empty()) h.data.reserve(16);
            c.data = h.data.data();
            c.nulls = h.has_nulls ? h.nulls.data() : nullptr;
            if (h.has_nulls && h.nulls.empty()) {
                h.nulls.reserve(16);
                c.nulls = h.nulls.data();
            }
            c.offsets = is_binary(h.type) ? h.offsets.data() : nullptr;
            c.n_bytes = (int64_t)h.data.size();
        }
        out->n_rows = n;
        out->n_cols = (int32_t)hcols_.size();
        out->slot_ids = hslots_.data();
        out->cols = cview_.data();
    }

    PlanCopy plan_;
    Lowered L_;
    bool lowered_ = false, filter_only_ = false;
    int shape_ = -1;
    // the static kernel compiled at run time for this plan (srb_jit.h); SRB_JIT = off | async (default) | sync
    static constexpr int SHAPE_JIT = 100; // what profile() reports as the shape once such a kernel has run
    int jit_mode_ = jit_mode_from_env();
    int64_t jit_min_rows_ = getenv("SRB_JIT_MIN_ROWS") ? atoll(getenv("SRB_JIT_MIN_ROWS")) : ((int64_t)4 << 20);
    std::string jit_src_[2];
    std::shared_ptr<JitEntry> jit_[2];
    const JitLoaded* jit_kernel_ = nullptr; // the variant this batch runs, nullptr: a built-in shape or the interpreter
    bool jit_failed_[2] = {false, false};
    int jit_gmax_[2] = {0, 0}; // directory size of the variant's shape (0: table path)
    int64_t jit_launches_ = 0;
    static int jit_mode_from_env() {
        const char* m = getenv("SRB_JIT");
        if (!m || !*m || !strcmp(m, "async") || !strcmp(m, "1")) return 1;
        if (!strcmp(m, "sync")) return 2;
        return (!strcmp(m, "off") || !strcmp(m, "0")) ? 0 : 1;
    }
    bool force_generic_ = getenv("SRB_FORCE_GENERIC") != nullptr; // test hook: compare the two kernel families
    bool no_fast_filter_ = getenv("SRB_NO_FAST_FILTER") != nullptr; // test hook: the per-row mask kernel for every filter-only plan
    bool no_narrow_ = getenv("SRB_NO_NARROW") != nullptr;         // test / measurement hook: the static kernels' general tile body only
    bool rf_fused_ = getenv("SRB_RF_FUSED") != nullptr;           // opt-in: all runtime filters of a scan in one kernel (k_filter_rf_multi)
    // k_scan_agg_lowcard measured SLOWER than the shared mirror on a B200 (Q1 with VARCHAR keys: 92 vs 49 ms per 134 M rows — its
    // 96 KB of private accumulators leave 8 warps per SM for an interpreter that lives on latency hiding): opt-in, SRB_LOWCARD=1
    bool no_lowcard_ = getenv("SRB_LOWCARD") == nullptr || getenv("SRB_NO_LOWCARD") != nullptr;
    bool no_shared_agg_ = getenv("SRB_NO_SHARED_AGG") != nullptr; // test / measurement hook: the interpreter updates the table per row
    bool no_partition_ = getenv("SRB_NO_PARTITION") != nullptr;   // test / measurement hook: direct table updates for every batch
    // The radix-partitioned aggregation is correct (tests/test_gpu_radix_agg.py) but measured SLOWER than the table path on a B200
    // (61 ms vs 26 ms for 600 M rows / 100 M keys), so it is opt-in: SRB_RADIX=1.
    bool no_radix_ = getenv("SRB_RADIX") == nullptr || getenv("SRB_NO_RADIX") != nullptr;
    int64_t radix_min_rows_ = getenv("SRB_RADIX_MIN_ROWS") ? atoll(getenv("SRB_RADIX_MIN_ROWS")) : ((int64_t)1 << 24); // (test hook: smaller batches)
    DevBuf page_err_, merge_send_, merge_recv_;
    int64_t prefetch_rows_ = 0;        // result rows [0, prefetch_rows_) sit in pack_host_ (prefetch_small_result)
    std::vector<size_t> prefetch_off_; // where each row piece starts there
    bool pages_pushed_ = false;
    bool table_ready_ = false;                                     // the group table is allocated and cleared (init_table defers large ones)
    int64_t radix_rows_ = 0, radix_batches_ = 0;                   // groups the radix path left at the front of the output arrays
    DevBuf rx_cursor_, rx_a_, rx_b_, rx_ctr_;
    PinnedBuf rx_hctr_;
    DevBuf hc_k_, hc_v_, hc_v2_, hc_ctr_;                                 // permuted (key, value) copy of the current batch + its 2 x 256 counters
    ColBind hc_cols_[MAX_COLS], nn_cols_[MAX_COLS];
    bool batch_partitioned_ = false, batch_has_nulls_ = false, batch_stripped_ = false;
    DevBuf pack_dev_;    // small results: all pulled columns packed contiguously ...
    PinnedBuf pack_host_; // ... and their pinned landing buffer
    int64_t staging_rows_, chunk_size_;
    cudaStream_t copy_stream_ = nullptr;
    cudaEvent_t kernel_done_ = nullptr, ev0_ = nullptr, ev1_ = nullptr;
    bool timed_pending_ = false;
    double kernel_ms_ = 0;
    int64_t timed_batches_ = 0, static_launches_ = 0;
    // table
    DevBuf slots_, defer_[2];
    int defer_cur_ = 0;
    uint64_t cap_ = 0, occupied_ = 0, selected_total_ = 0;
    DevBuf ctr_dev_;
    PinnedBuf ctr_host_;
    // staging
    Stage stage_[2];
    int cur_stage_ = 0, pending_stage_ = 0;
    bool pending_ = false;
    int64_t pending_rows_ = 0;
    ColBind pending_cols_[MAX_COLS];
    std::vector<uint32_t> off_scratch_;
    // filter-only
    DevBuf mask_, tile_counts_, tile_offsets_;
    std::vector<int> out_col_idx_;
    std::vector<FOut> fout_;
    // output
    struct EVal {
        DevBuf val, null, cnt;
    };
    DevBuf ekeys_;
    std::vector<FOut> dkeys_;
    std::vector<EVal> eval_;
    int64_t out_rows_ = 0, out_pos_ = 0;
    // streaming pre-aggregation sink
    int stream_mode_ = -1;    // -1: blocking sink; otherwise TStreamingPreaggregationMode
    int64_t stream_mem_limit_ = 0; // LIMITED_MEM: bytes of group table from which on rows pass through (0: never)
    int64_t stream_rows_ = 0; // pass-through rows at the front of the output arrays
    int64_t stream_seen_ = 0; // input rows the sink has been handed
    int64_t stream_unit_ = getenv("SRB_STREAM_UNIT") ? std::max<int64_t>(1, atoll(getenv("SRB_STREAM_UNIT"))) : ((int64_t)1 << 24);
    StreamPolicy policy_;
    bool pulled_empty_ = false;
    std::vector<HostCol> hcols_;
    std::vector<int32_t> hslots_;
    std::vector<srb_column> cview_;
};

void ScanAggOp::device_result_columns(int64_t n) {
    const DProg& P = L_.prog;
    UnpackArgs ua{};
    ua.n_keys = P.n_keys;
    dkeys_.resize(P.n_keys);
    for (int k = 0; k < P.n_keys; k++) {
        const ColMeta& cm = L_.key_out[k];
        if (P.keys[k].width == 0) throw Error(SRB_NOT_IMPLEMENTED, "device pull of binary group keys");
        dkeys_[k].data.ensure(ctx, (size_t)std::max<int64_t>(n, 1) * P.keys[k].width);
        if (cm.nullable) dkeys_[k].nulls.ensure(ctx, (size_t)std::max<int64_t>(n, 1));
        ua.keys[k] = UnpackKey{dkeys_[k].data.as<uint8_t>(), cm.nullable ? dkeys_[k].nulls.as<uint8_t>() : nullptr, P.keys[k].width,
                               cm.nullable ? 1 : 0};
        srb_column c{};
        c.type = cm.type;
        c.precision = cm.precision;
        c.scale = cm.scale;
        c.mem_kind = SRB_MEM_DEVICE;
        c.data = dkeys_[k].data.p;
        c.nulls = cm.nullable ? dkeys_[k].nulls.as<uint8_t>() : nullptr;
        cview_.push_back(c);
        hslots_.push_back(k);
    }
    if (P.n_keys && n > 0) {
        k_unpack_keys<<<grid_for(n), 256, 0, stream>>>(ua, ekeys_.as<uint64_t>() + (size_t)out_pos_ * (key_rec_bytes() / 8), (int)(key_rec_bytes() / 8), n);
        launches++;
        SRB_CUDA(cudaGetLastError());
    }
    for (int a = 0; a < P.n_aggs; a++) {
        const AggOutMeta& m = L_.agg_out[a];
        if (m.blob) throw Error(SRB_NOT_IMPLEMENTED, "device pull of AVG intermediate blobs");
        srb_column c{};
        c.type = m.out_type;
        c.precision = m.out_precision;
        c.scale = m.out_scale;
        c.mem_kind = SRB_MEM_DEVICE;
        c.data = eval_[a].val.as<uint8_t>() + (size_t)out_pos_ * m.out_width;
        c.nulls = m.nullable ? eval_[a].null.as<uint8_t>() + out_pos_ : nullptr;
        cview_.push_back(c);
        hslots_.push_back(P.n_keys + a);
    }
    SRB_CUDA(cudaStreamSynchronize(stream));
}

void jit_wait_all() { JitCache::get().wait_all(); }
void jit_stats(int64_t out[4], std::string* last_error) {
    JitCache::get().stats(out);
    if (last_error) *last_error = JitCache::get().last_error();
}
// lowers the plan against the first chunk's column types, generates its shape and compiles it (on the calling thread; no GPU needed)
void jit_check_plan(const srb_scan_agg_plan* plan, const srb_chunk* first, std::string* description) {
    if (!plan || !first) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    PlanCopy pc(plan);
    if (pc.aggs.empty()) throw Error(SRB_NOT_IMPLEMENTED, "filter-only plans have no static kernel");
    Lowered L = lower_plan(pc, first);
    JitShape js;
    if (match_static_shape(L.prog) >= 0) {
        if (description) *description = std::string("built-in shape: ") + static_shape_name(match_static_shape(L.prog));
        return;
    }
    bool nullable[MAX_COLS] = {};
    bool any = false;
    for (size_t i = 0; i < L.cols.size() && i < (size_t)MAX_COLS; i++) any = (nullable[i] = L.cols[i].nullable) || any;
    if (!jit_shape_of(L.prog, (int)std::min<int64_t>(pc.group_hint, 1 << 30), any ? nullable : nullptr, js))
        throw Error(SRB_NOT_IMPLEMENTED, "the plan is outside what the static kernel covers: the interpreter runs it");
    std::shared_ptr<JitEntry> e = JitCache::get().acquire(jit_source(js), true);
    if (e->failed()) throw Error(SRB_INTERNAL_ERROR, e->error());
    if (description)
        *description = fmt("compiled shape: %d nodes, %d aggregates, %d keys, directory %d, %d threads x %d rows, cubin %zu bytes, %.0f ms%s", js.n_nodes,
                           js.n_aggs, js.n_keys, js.gmax, js.threads, js.rows, e->cubin.size(), e->compile_ms, e->from_disk ? " (disk cache)" : "");
}

srb_op* make_scan_agg_op(srb_ctx* ctx, const srb_scan_agg_plan* plan) {
    if (!plan) throw Error(SRB_INVALID_ARGUMENT, "null plan");
    return new ScanAggOp(ctx, plan);
}

} // namespace srb
