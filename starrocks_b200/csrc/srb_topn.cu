// srb_topn.cu — ORDER BY ... [OFFSET] LIMIT on the device (the last operator of TPC-H Q3's plan,
// ORDER BY revenue DESC, o_orderdate LIMIT 10).
//
// Reference: ChunksSorterTopn (be/src/exec/chunks_sorter_topn.cpp: update() buffers chunks and filters them against the current
// top rows, done() leaves rows [offset, offset + limit) of the sorted order), SortDesc (be/src/exec/sorting/sorting.h:86-101:
// direction per key; NULLs first or last in the OUTPUT whatever the direction), column-wise three-way compare
// (be/src/exec/sorting/compare_column.cpp), rows that tie on every key keep their arrival order (ChunksSorterTest.stable_sort).
//
// Design. Every row gets a 32-byte order-preserving composite key: per sort column one NULL-rank byte and the value's bytes in
// big-endian, sign-flipped (and complemented for DESC) form, so that comparing composite keys as unsigned big-endian strings IS the
// reference's multi-column compare; the row's arrival sequence number breaks ties (stable, total order, deterministic).
//   k_topn_encode: key columns -> entries {key[4], seq, loc}                               (one pass over the key columns only)
//   k_topn_reduce: tiles of 2048 entries are bitonic-sorted in shared memory and only the best L = offset + limit of each tile
//                  survive; repeated (>= 2x fewer entries per pass) until one tile is left, which is then fully sorted
//   k_topn_gather: the payload columns of the <= L survivors are copied into operator-owned state, so the pushed chunk can be
//                  released when push_chunk returns; the state's rows re-enter the next push's tournament as L ready-made entries.
// The payload columns are read once, for L rows; the key columns once per pushed row. L <= 1024.
#include <algorithm>
#include <cstring>
#include <vector>

#include "srb_host.h"
#include "srb_lower.h"

namespace srb {

constexpr int TOPN_MAX_ROWS = 1024, TOPN_TILE = 2048, TOPN_THREADS = 1024, TOPN_MAX_KEYS = 8, TOPN_KEY_BYTES = 32, TOPN_MAX_COLS = 32;
constexpr uint64_t TOPN_STATE_FLAG = 1ULL << 63;

struct TEntry {
    uint64_t k[4];
    uint64_t seq; // arrival number of the row: the stable tie-break
    uint64_t loc; // row inside the pushed chunk, or TOPN_STATE_FLAG | row inside the operator's state
};

struct TKeyCol {
    const uint8_t* data;
    const uint8_t* nulls;
    int32_t width;      // 1, 2, 4, 8, 16
    int32_t is_double;  // IEEE-754 order-preserving encoding instead of the sign flip
    int32_t desc;       // complement the value bytes
    int32_t null_first; // rank byte: NULLs before (1) or after (0) the values
    int32_t byte_off;   // position of this column's rank byte inside the composite key
};
struct TEncodeArgs {
    int32_t n_keys;
    TKeyCol cols[TOPN_MAX_KEYS];
};

SRB_D void key_put_byte(uint64_t (&k)[4], int pos, uint32_t b) { k[pos >> 3] |= (uint64_t)(b & 0xFFu) << (56 - 8 * (pos & 7)); }

// Rows [row0, row0 + n) of the pushed chunk -> entries. thr == nullptr: entry j of the range is written to out[j]. thr != nullptr (the
// worst of `keep` already sorted candidates): only rows that sort strictly before *thr can still reach the result (ChunksSorterTopn filters
// every new chunk against its current top rows the same way, chunks_sorter_topn.cpp); they are appended at out[atomicAdd(counter)].
__global__ void __launch_bounds__(256) k_topn_encode(const __grid_constant__ TEncodeArgs a, int64_t row0, int64_t n, uint64_t seq_base, TEntry* __restrict__ out,
                                                      const TEntry* __restrict__ thr, unsigned long long* __restrict__ counter) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    uint64_t t[5] = {0, 0, 0, 0, 0};
    if (thr) {
        t[0] = thr->k[0];
        t[1] = thr->k[1];
        t[2] = thr->k[2];
        t[3] = thr->k[3];
        t[4] = thr->seq;
    }
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride) {
        const int64_t i = row0 + j;
        uint64_t k[4] = {0, 0, 0, 0};
        for (int c = 0; c < a.n_keys; c++) {
            const TKeyCol& kc = a.cols[c];
            const bool isnull = kc.nulls != nullptr && kc.nulls[i] != 0;
            key_put_byte(k, kc.byte_off, isnull ? (kc.null_first ? 0u : 1u) : (kc.null_first ? 1u : 0u));
            if (isnull) continue; // value bytes stay zero: every NULL of a column compares equal
            // the value as an unsigned 128-bit number whose order is the column's order
            uint64_t lo = 0, hi = 0;
            const uint8_t* p = kc.data + (size_t)i * kc.width;
            switch (kc.width) {
            case 1: lo = (uint8_t)(*(const int8_t*)p) ^ 0x80u; break;
            case 2: lo = (uint16_t)(*(const int16_t*)p) ^ 0x8000u; break;
            case 4: lo = (uint32_t)(*(const int32_t*)p) ^ 0x80000000u; break;
            case 8:
                lo = *(const uint64_t*)p;
                if (kc.is_double) {
                    if (lo == (1ULL << 63)) lo = 0; // -0.0 compares equal to +0.0
                    lo = (lo >> 63) ? ~lo : (lo | (1ULL << 63)); // negative doubles: all bits flipped; others: sign bit set
                }
                else
                    lo ^= 1ULL << 63;
                break;
            default:
                lo = ((const uint64_t*)p)[0];
                hi = ((const uint64_t*)p)[1] ^ (1ULL << 63);
            }
            if (kc.desc) {
                lo = ~lo;
                hi = ~hi;
            }
            // big-endian bytes, most significant first
            for (int b = 0; b < kc.width; b++) {
                const int byte = kc.width - 1 - b; // significance
                const uint32_t v = byte < 8 ? (uint32_t)(lo >> (8 * byte)) : (uint32_t)(hi >> (8 * (byte - 8)));
                key_put_byte(k, kc.byte_off + 1 + b, v);
            }
        }
        TEntry e;
        e.k[0] = k[0];
        e.k[1] = k[1];
        e.k[2] = k[2];
        e.k[3] = k[3];
        e.seq = seq_base + (uint64_t)j;
        e.loc = (uint64_t)i;
        if (!thr) {
            out[j] = e;
            continue;
        }
        bool before = false; // e < *thr in the order of entry_less
        const uint64_t w[5] = {k[0], k[1], k[2], k[3], e.seq};
#pragma unroll
        for (int f = 0; f < 5; f++) {
            if (w[f] != t[f]) {
                before = w[f] < t[f];
                break;
            }
        }
        if (before) out[atomicAdd(counter, 1ULL)] = e;
    }
}

// shared memory of the tournament: structure of arrays, one 8-byte word per entry and field (conflict-free by construction)
struct TSmem {
    uint64_t w[6][TOPN_TILE]; // k0..k3, seq, loc
};

SRB_D bool entry_less(const TSmem& s, int a, int b) {
#pragma unroll
    for (int f = 0; f < 5; f++) { // loc is not part of the order
        const uint64_t x = s.w[f][a], y = s.w[f][b];
        if (x != y) return x < y;
    }
    return false;
}

__global__ void __launch_bounds__(TOPN_THREADS, 1) k_topn_reduce(const TEntry* __restrict__ in, int64_t n, int keep, TEntry* __restrict__ out) {
    extern __shared__ __align__(16) uint8_t raw[];
    TSmem& s = *reinterpret_cast<TSmem*>(raw);
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * TOPN_TILE;
    const int rows = (int)(n - base < TOPN_TILE ? n - base : TOPN_TILE);
    for (int i = tid; i < TOPN_TILE; i += TOPN_THREADS) {
        if (i < rows) {
            const TEntry e = in[base + i];
            s.w[0][i] = e.k[0];
            s.w[1][i] = e.k[1];
            s.w[2][i] = e.k[2];
            s.w[3][i] = e.k[3];
            s.w[4][i] = e.seq;
            s.w[5][i] = e.loc;
        } else { // padding sorts behind every real entry
#pragma unroll
            for (int f = 0; f < 6; f++) s.w[f][i] = ~0ULL;
        }
    }
    __syncthreads();
    for (int k = 2; k <= TOPN_TILE; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int i = 2 * j * (tid / j) + (tid % j), p = i + j; // TOPN_TILE / 2 == TOPN_THREADS: one pair per thread
            const bool ascending = (i & k) == 0;
            const bool swap = ascending ? entry_less(s, p, i) : entry_less(s, i, p);
            if (swap) {
#pragma unroll
                for (int f = 0; f < 6; f++) {
                    const uint64_t t = s.w[f][i];
                    s.w[f][i] = s.w[f][p];
                    s.w[f][p] = t;
                }
            }
            __syncthreads();
        }
    }
    const int m = rows < keep ? rows : keep; // only the last tile can hold fewer than `keep` rows: the output stays dense
    for (int i = tid; i < m; i += TOPN_THREADS) {
        TEntry e;
        e.k[0] = s.w[0][i];
        e.k[1] = s.w[1][i];
        e.k[2] = s.w[2][i];
        e.k[3] = s.w[3][i];
        e.seq = s.w[4][i];
        e.loc = s.w[5][i];
        out[(int64_t)blockIdx.x * keep + i] = e;
    }
}

struct TGatherCol {
    const uint8_t* src_new;
    const uint8_t* src_state;
    uint8_t* dst;
    const uint8_t* null_new; // may be nullptr: the pushed column has no null map
    const uint8_t* null_state;
    uint8_t* null_dst;       // nullptr: the column is not nullable
    int32_t width;
};
struct TGatherArgs {
    int32_t n_cols;
    TGatherCol cols[TOPN_MAX_COLS];
};

// one thread per surviving row: every column's element moves from the pushed chunk or the old state into the new state
__global__ void __launch_bounds__(256) k_topn_gather(const __grid_constant__ TGatherArgs a, TEntry* __restrict__ entries, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t loc = entries[i].loc;
    const bool from_state = (loc & TOPN_STATE_FLAG) != 0;
    const size_t row = (size_t)(loc & ~TOPN_STATE_FLAG);
    for (int c = 0; c < a.n_cols; c++) {
        const TGatherCol& g = a.cols[c];
        const uint8_t* src = (from_state ? g.src_state : g.src_new) + row * g.width;
        uint8_t* dst = g.dst + (size_t)i * g.width;
        switch (g.width) {
        case 1: *dst = *src; break;
        case 2: *(uint16_t*)dst = *(const uint16_t*)src; break;
        case 4: *(uint32_t*)dst = *(const uint32_t*)src; break;
        case 8: *(uint64_t*)dst = *(const uint64_t*)src; break;
        default: *(ulonglong2*)dst = *(const ulonglong2*)src;
        }
        if (g.null_dst) {
            const uint8_t* ns = from_state ? g.null_state : g.null_new;
            g.null_dst[i] = ns ? ns[row] : 0;
        }
    }
    entries[i].loc = TOPN_STATE_FLAG | (uint64_t)i;
}

class TopNOp : public srb_op {
public:
    TopNOp(srb_ctx* c, const srb_topn_plan* p) : srb_op(c) {
        if (!p) throw Error(SRB_INVALID_ARGUMENT, "null plan");
        if (p->n_keys < 1 || p->n_keys > TOPN_MAX_KEYS) throw Error(SRB_NOT_IMPLEMENTED, fmt("ORDER BY with %d keys", p->n_keys));
        if (p->offset < 0 || p->limit < 0) throw Error(SRB_INVALID_ARGUMENT, "negative offset or limit");
        if (p->offset + p->limit > TOPN_MAX_ROWS)
            throw Error(SRB_NOT_IMPLEMENTED, fmt("offset + limit above %d rows (a full sort, not a TopN)", TOPN_MAX_ROWS));
        keys_.assign(p->key_slots, p->key_slots + p->n_keys);
        asc_.assign(p->is_asc, p->is_asc + p->n_keys);
        nulls_first_.assign(p->nulls_first, p->nulls_first + p->n_keys);
        offset_ = p->offset;
        keep_ = (int)(p->offset + p->limit);
        cudaFuncSetAttribute(k_topn_reduce, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TSmem));
    }
    ~TopNOp() override { cudaStreamSynchronize(stream); }

    void push_chunk(const srb_chunk* ch) override {
        if (finishing) throw Error(SRB_INTERNAL_ERROR, "push_chunk after set_finishing");
        if (cancelled) throw Error(SRB_CANCELLED, "operator cancelled");
        if (!ch || ch->n_rows < 0) throw Error(SRB_INVALID_ARGUMENT, "bad chunk");
        if (ch->n_rows == 0) return;
        if (ch->n_rows >= ((int64_t)1 << 32)) throw Error(SRB_INVALID_ARGUMENT, "chunk exceeds Column::MAX_CAPACITY_LIMIT rows");
        if (!described_) describe(ch);
        const int64_t n = ch->n_rows;
        rows_in += n;
        if (keep_ == 0) { // LIMIT 0
            seq_ += (uint64_t)n;
            return;
        }
        // the columns of this chunk on the device (device columns in place, host columns staged)
        const int nc = (int)cols_.size();
        std::vector<const uint8_t*> data(nc), nulls(nc, nullptr);
        for (int i = 0; i < nc; i++) {
            const srb_column* c = find_slot(ch, cols_[i].slot);
            if (!c) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d not in chunk", cols_[i].slot));
            if (c->type != cols_[i].type) throw Error(SRB_INVALID_ARGUMENT, fmt("slot %d changed type", cols_[i].slot));
            if (c->nulls && !cols_[i].nullable) throw Error(SRB_NOT_IMPLEMENTED, fmt("slot %d: null map appeared after the first chunk had none", cols_[i].slot));
            if (!c->data) throw Error(SRB_INVALID_ARGUMENT, "null data pointer");
            const int w = cols_[i].width;
            if (c->mem_kind == SRB_MEM_DEVICE) {
                data[i] = (const uint8_t*)c->data;
                nulls[i] = c->nulls;
            } else {
                stage_[i].data.ensure(ctx, (size_t)n * w);
                SRB_CUDA(cudaMemcpyAsync(stage_[i].data.p, c->data, (size_t)n * w, cudaMemcpyHostToDevice, stream));
                data[i] = stage_[i].data.as<uint8_t>();
                if (c->nulls) {
                    stage_[i].nulls.ensure(ctx, (size_t)n);
                    SRB_CUDA(cudaMemcpyAsync(stage_[i].nulls.p, c->nulls, (size_t)n, cudaMemcpyHostToDevice, stream));
                    nulls[i] = stage_[i].nulls.as<uint8_t>();
                }
            }
        }
        // entries: the state's survivors first (already encoded, loc = state row), then the pushed rows
        const int64_t total = state_rows_ + n;
        ent_[0].ensure(ctx, (size_t)total * sizeof(TEntry));
        ent_[1].ensure(ctx, (size_t)total * sizeof(TEntry)); // both sized for the worst case up front: ensure() does not keep contents
        TEntry* e0 = ent_[0].as<TEntry>();
        if (state_rows_) SRB_CUDA(cudaMemcpyAsync(e0, state_ent_.p, (size_t)state_rows_ * sizeof(TEntry), cudaMemcpyDeviceToDevice, stream));
        TEncodeArgs ea{};
        ea.n_keys = (int)keys_.size();
        for (size_t k = 0; k < keys_.size(); k++) {
            const ColDesc& cd = cols_[key_col_[k]];
            ea.cols[k] = TKeyCol{data[key_col_[k]], nulls[key_col_[k]], cd.width, cd.type == SRB_TYPE_DOUBLE ? 1 : 0, asc_[k] ? 0 : 1,
                                 nulls_first_[k] ? 1 : 0, key_off_[k]};
        }
        auto grid_for = [&](int64_t rows) { return (int)std::max<int64_t>(1, std::min<int64_t>((rows + 255) / 256, (int64_t)ctx->n_sms * 8)); };
        // tournament: every pass keeps the best keep_ entries of each 2048-entry tile, until one (fully sorted) tile is left
        int cur = 0;
        auto tournament = [&](int64_t cnt) {
            for (;;) {
                const int64_t tiles = (cnt + TOPN_TILE - 1) / TOPN_TILE;
                const int64_t last_rows = cnt - (tiles - 1) * TOPN_TILE;
                const int64_t out_cnt = (tiles - 1) * keep_ + std::min<int64_t>(keep_, last_rows);
                k_topn_reduce<<<(int)tiles, TOPN_THREADS, sizeof(TSmem), stream>>>(ent_[cur].as<TEntry>(), cnt, keep_, ent_[cur ^ 1].as<TEntry>());
                launches++;
                cur ^= 1;
                cnt = out_cnt;
                if (tiles == 1) return cnt;
            }
        };
        // A large chunk is not sorted whole: the state and a leading sample give keep_ sorted candidates, whose worst entry is a bound no
        // later result row can exceed; the rest of the chunk is encoded through that bound and only the rows before it (about
        // keep_ * rows / sample of them unless the chunk arrives sorted the wrong way round) join the tournament.
        const int64_t sample = (int64_t)ctx->n_sms * TOPN_TILE;
        int64_t cnt;
        if (n <= 2 * sample || getenv("SRB_TOPN_NO_BOUND")) {
            k_topn_encode<<<grid_for(n), 256, 0, stream>>>(ea, 0, n, seq_, e0 + state_rows_, nullptr, nullptr);
            launches++;
            cnt = tournament(total);
        } else {
            k_topn_encode<<<grid_for(sample), 256, 0, stream>>>(ea, 0, sample, seq_, e0 + state_rows_, nullptr, nullptr);
            launches++;
            cnt = tournament(state_rows_ + sample); // == keep_: the sample alone holds more rows than that
            counter_.ensure(ctx, 8);
            hcounter_.ensure(8);
            unsigned long long* hc = hcounter_.as<unsigned long long>();
            *hc = (unsigned long long)cnt;
            SRB_CUDA(cudaMemcpyAsync(counter_.p, hc, 8, cudaMemcpyHostToDevice, stream));
            TEntry* cand = ent_[cur].as<TEntry>();
            k_topn_encode<<<grid_for(n - sample), 256, 0, stream>>>(ea, sample, n - sample, seq_ + (uint64_t)sample, cand, cand + (cnt - 1),
                                                                    counter_.as<unsigned long long>());
            launches++;
            SRB_CUDA(cudaMemcpyAsync(hc, counter_.p, 8, cudaMemcpyDeviceToHost, stream));
            SRB_CUDA(cudaStreamSynchronize(stream));
            cnt = tournament((int64_t)*hc);
        }
        seq_ += (uint64_t)n;
        // payload of the survivors into the other state buffer
        const int ns = state_cur_ ^ 1;
        TGatherArgs ga{};
        ga.n_cols = nc;
        for (int i = 0; i < nc; i++) {
            const int w = cols_[i].width;
            state_[ns][i].data.ensure(ctx, (size_t)keep_ * w);
            if (cols_[i].nullable) state_[ns][i].nulls.ensure(ctx, (size_t)keep_);
            ga.cols[i] = TGatherCol{data[i], state_[state_cur_][i].data.as<uint8_t>(), state_[ns][i].data.as<uint8_t>(), nulls[i],
                                    cols_[i].nullable ? state_[state_cur_][i].nulls.as<uint8_t>() : nullptr,
                                    cols_[i].nullable ? state_[ns][i].nulls.as<uint8_t>() : nullptr, w};
        }
        k_topn_gather<<<(int)((cnt + 255) / 256), 256, 0, stream>>>(ga, ent_[cur].as<TEntry>(), (int)cnt);
        launches++;
        state_ent_.ensure(ctx, (size_t)keep_ * sizeof(TEntry));
        SRB_CUDA(cudaMemcpyAsync(state_ent_.p, ent_[cur].p, (size_t)cnt * sizeof(TEntry), cudaMemcpyDeviceToDevice, stream));
        SRB_CUDA(cudaGetLastError());
        SRB_CUDA(cudaStreamSynchronize(stream)); // the caller may release the chunk (and the staging buffers are reused)
        state_cur_ = ns;
        state_rows_ = cnt;
    }

    void set_finishing() override {
        if (finishing) return;
        finishing = true;
        out_rows_ = std::max<int64_t>(0, state_rows_ - offset_);
        rows_selected = out_rows_;
    }
    bool has_output() override { return finishing && !pulled_ && out_rows_ > 0; }
    bool need_input() override { return !finishing; }
    bool is_finished() override { return finishing && (pulled_ || out_rows_ == 0); }

    // the whole result (at most 1024 rows) in one chunk: sorted rows [offset, offset + limit)
    srb_status pull_chunk(srb_chunk* out, bool) override {
        if (!finishing) throw Error(SRB_INTERNAL_ERROR, "pull_chunk before set_finishing");
        if (pulled_ || (out_rows_ == 0 && !described_)) {
            pulled_ = true;
            return SRB_END_OF_FILE;
        }
        const int nc = (int)cols_.size();
        host_.assign(nc, {});
        view_.assign(nc, srb_column{});
        slots_.resize(nc);
        for (int i = 0; i < nc; i++) {
            const int w = cols_[i].width;
            host_[i].data.resize((size_t)std::max<int64_t>(1, out_rows_) * w);
            if (out_rows_)
                SRB_CUDA(cudaMemcpyAsync(host_[i].data.data(), state_[state_cur_][i].data.as<uint8_t>() + (size_t)offset_ * w, (size_t)out_rows_ * w,
                                         cudaMemcpyDeviceToHost, stream));
            if (cols_[i].nullable) {
                host_[i].nulls.resize((size_t)std::max<int64_t>(1, out_rows_));
                if (out_rows_)
                    SRB_CUDA(cudaMemcpyAsync(host_[i].nulls.data(), state_[state_cur_][i].nulls.as<uint8_t>() + offset_, (size_t)out_rows_,
                                             cudaMemcpyDeviceToHost, stream));
            }
        }
        SRB_CUDA(cudaStreamSynchronize(stream));
        for (int i = 0; i < nc; i++) {
            srb_column& c = view_[i];
            c.type = cols_[i].type;
            c.precision = cols_[i].precision;
            c.scale = cols_[i].scale;
            c.mem_kind = SRB_MEM_HOST;
            c.data = host_[i].data.data();
            c.nulls = cols_[i].nullable ? host_[i].nulls.data() : nullptr;
            c.offsets = nullptr;
            c.n_bytes = (int64_t)out_rows_ * cols_[i].width;
            slots_[i] = cols_[i].slot;
        }
        out->n_rows = out_rows_;
        out->n_cols = nc;
        out->slot_ids = slots_.data();
        out->cols = view_.data();
        pulled_ = true;
        return SRB_OK;
    }

private:
    struct ColDesc {
        int slot, type, precision, scale, width;
        bool nullable;
    };
    struct DevCol {
        DevBuf data, nulls;
    };
    struct HostCol {
        std::vector<uint8_t> data, nulls;
    };

    void describe(const srb_chunk* ch) {
        if (ch->n_cols > TOPN_MAX_COLS) throw Error(SRB_NOT_IMPLEMENTED, fmt("more than %d columns through a TopN", TOPN_MAX_COLS));
        for (int i = 0; i < ch->n_cols; i++) {
            const srb_column& c = ch->cols[i];
            if (is_binary(c.type)) throw Error(SRB_NOT_IMPLEMENTED, "VARCHAR / VARBINARY columns through the device TopN");
            cols_.push_back(ColDesc{ch->slot_ids[i], c.type, c.precision, c.scale, type_width(c.type), c.nulls != nullptr});
        }
        int off = 0;
        for (size_t k = 0; k < keys_.size(); k++) {
            int ci = -1;
            for (size_t i = 0; i < cols_.size(); i++)
                if (cols_[i].slot == keys_[k]) ci = (int)i;
            if (ci < 0) throw Error(SRB_INVALID_ARGUMENT, fmt("sort key slot %d not in chunk", keys_[k]));
            key_col_.push_back(ci);
            key_off_.push_back(off);
            off += 1 + cols_[ci].width;
        }
        if (off > TOPN_KEY_BYTES) throw Error(SRB_NOT_IMPLEMENTED, fmt("sort keys of %d bytes (the composite key holds %d)", off, TOPN_KEY_BYTES));
        stage_.resize(cols_.size());
        state_[0].resize(cols_.size());
        state_[1].resize(cols_.size());
        described_ = true;
    }

    std::vector<int32_t> keys_;
    std::vector<uint8_t> asc_, nulls_first_;
    int64_t offset_ = 0;
    int keep_ = 0;
    bool described_ = false, pulled_ = false;
    std::vector<ColDesc> cols_;
    std::vector<int> key_col_, key_off_;
    std::vector<DevCol> stage_, state_[2];
    int state_cur_ = 0;
    int64_t state_rows_ = 0, out_rows_ = 0;
    uint64_t seq_ = 0;
    DevBuf ent_[2], state_ent_, counter_;
    PinnedBuf hcounter_;
    std::vector<HostCol> host_;
    std::vector<srb_column> view_;
    std::vector<int32_t> slots_;
};

srb_op* make_topn_op(srb_ctx* ctx, const srb_topn_plan* plan) { return new TopNOp(ctx, plan); }

} // namespace srb
