// srb_exchange.cu — hash partitioning (K9) and the NCCL exchange that stands where the brpc shuffle is in the reference.
//
// Replaces:
//   ExchangeSinkOperator::push_chunk, HASH_PARTITIONED branch    be/src/exec/pipeline/exchange/exchange_sink_operator.cpp:569-633
//     per-row hash = FNV-1 chained over the partition columns    be/src/column/fixed_length_column_base.cpp:223-227,
//       (NULL: h ^= 0x9e3779b9 + (h<<6) + (h>>2))                be/src/column/nullable_column.cpp:346-369, util/hash_util.hpp:195-214
//     channel = ReduceOp(hash, n) = (hash * n) >> 32             be/src/exec/pipeline/exchange/shuffler.h:70-85
//     counting sort of row ids per channel + append_selective    exchange_sink_operator.cpp:595-630
//   UNPARTITIONED (broadcast) branch                              exchange_sink_operator.cpp:508-547
//   the transport: PTransmitChunkParams over brpc                 be/src/runtime/data_stream_{sender,recvr}.cpp
// by: one counting pass + one scatter pass on the device that leave every destination's rows contiguous, then a grouped
// ncclSend/ncclRecv all-to-all (counts exchanged first with ncclAllGather), or ncclBroadcast / ncclAllGather for replicated
// build sides. NCCL is bound at run time (dlopen libnccl.so.2 — the copy torch already loaded when there is one), so a
// single-GPU process never needs it.
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <cstring>
#include <string>
#include <vector>

#include "srb_dev.cuh"
#include "srb_host.h"

namespace srb {

constexpr int PMAX_COLS = 16, PMAX_KEYS = 4, PMAX_PARTS = 64;

struct PCol {
    const uint8_t* src;
    uint8_t* dst;
    const uint8_t* src_null;
    uint8_t* dst_null;
    int32_t width;
    int32_t _pad;
};
struct PArgs {
    int32_t n_cols, n_keys, n_parts, _pad;
    int32_t key_col[PMAX_KEYS];
    PCol cols[PMAX_COLS];
    unsigned long long base[PMAX_PARTS]; // first output row of each partition (scatter pass)
};

SRB_D uint32_t part_of(const PArgs& a, int64_t r) {
    uint32_t h = 0x811C9DC5u; // HashUtil::FNV_SEED
    for (int k = 0; k < a.n_keys; k++) {
        const PCol& c = a.cols[a.key_col[k]];
        if (c.src_null && c.src_null[r]) {
            h = h ^ (0x9e3779b9u + (h << 6) + (h >> 2));
        } else {
            // FNV-1 over the value's bytes, low byte first; 4- and 8-byte values are fetched with one load
            if (c.width == 4) {
                uint32_t x = ((const uint32_t*)c.src)[r];
#pragma unroll
                for (int b = 0; b < 4; b++, x >>= 8) h = ((x & 0xffu) ^ h) * 0x01000193u;
            } else if (c.width == 8) {
                uint64_t x = ((const uint64_t*)c.src)[r];
#pragma unroll
                for (int b = 0; b < 8; b++, x >>= 8) h = ((uint32_t)(x & 0xffu) ^ h) * 0x01000193u;
            } else {
                const uint8_t* p = c.src + (size_t)r * c.width;
                for (int b = 0; b < c.width; b++) h = (p[b] ^ h) * 0x01000193u;
            }
        }
    }
    return reduce_op(h, (uint32_t)a.n_parts);
}

// pass 1: destination per row (1 byte) + per-partition totals
__global__ void __launch_bounds__(256) k_part_count(const __grid_constant__ PArgs a, int64_t n, uint8_t* __restrict__ dest,
                                                     unsigned long long* __restrict__ totals) {
    __shared__ unsigned int hist[PMAX_PARTS];
    for (int i = threadIdx.x; i < PMAX_PARTS; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        uint32_t d = part_of(a, r);
        dest[r] = (uint8_t)d;
        unsigned peers = __match_any_sync(__activemask(), d);
        if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&hist[d], __popc(peers));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < a.n_parts; i += blockDim.x)
        if (hist[i]) atomicAdd(&totals[i], (unsigned long long)hist[i]);
}

// pass 2: each block reserves a contiguous run per partition for its tile, then scatters every column
constexpr int PTILE = 2048;
__global__ void __launch_bounds__(256) k_part_scatter(const __grid_constant__ PArgs a, int64_t n, const uint8_t* __restrict__ dest,
                                                       unsigned long long* __restrict__ cursors) {
    __shared__ unsigned int hist[PMAX_PARTS];
    __shared__ unsigned long long run_base[PMAX_PARTS];
    const int64_t n_tiles = (n + PTILE - 1) / PTILE;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int i = threadIdx.x; i < PMAX_PARTS; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        uint32_t d[PTILE / 256], rank[PTILE / 256];
#pragma unroll
        for (int j = 0; j < PTILE / 256; j++) {
            int64_t r = tile * PTILE + j * 256 + threadIdx.x;
            d[j] = 0xffffffffu;
            rank[j] = 0;
            bool on = r < n;
            if (on) d[j] = dest[r];
            unsigned act = __ballot_sync(0xffffffffu, on);
            if (on) {
                unsigned peers = __match_any_sync(act, d[j]);
                int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
                unsigned int b = 0;
                if (lane == leader) b = atomicAdd(&hist[d[j]], __popc(peers));
                b = __shfl_sync(peers, b, leader);
                rank[j] = b + __popc(peers & ((1u << lane) - 1));
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < a.n_parts; i += blockDim.x)
            run_base[i] = hist[i] ? a.base[i] + atomicAdd(&cursors[i], (unsigned long long)hist[i]) : 0;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < PTILE / 256; j++) {
            if (d[j] == 0xffffffffu) continue;
            const int64_t r = tile * PTILE + j * 256 + threadIdx.x;
            const unsigned long long pos = run_base[d[j]] + rank[j];
            for (int c = 0; c < a.n_cols; c++) {
                const PCol& pc = a.cols[c];
                switch (pc.width) {
                case 1: pc.dst[pos] = pc.src[r]; break;
                case 2: ((uint16_t*)pc.dst)[pos] = ((const uint16_t*)pc.src)[r]; break;
                case 4: ((uint32_t*)pc.dst)[pos] = ((const uint32_t*)pc.src)[r]; break;
                case 8: ((uint64_t*)pc.dst)[pos] = ((const uint64_t*)pc.src)[r]; break;
                default: ((ulonglong2*)pc.dst)[pos] = ((const ulonglong2*)pc.src)[r];
                }
                if (pc.src_null) pc.dst_null[pos] = pc.src_null[r];
            }
        }
        __syncthreads();
    }
}


// ---- shuffle fused with the exchange: scatter straight into the destination ranks' receive buffers ------------------------------
// The receive buffers of every rank are mapped into this process (cudaIpc, NVLink peer memory). The scatter pass of the partitioner
// then IS the all-to-all: a tile reserves a run per destination (shared-memory histogram + one atomic per destination and tile) and
// every row is stored at  peer[d].column[c] + (first_row_of_this_source_in_d + run + rank) * width  — remote stores over NVLink for
// d != me, local stores for d == me. No staging copy of the grouped rows, no ncclSend/ncclRecv; the W x W count matrix travels in
// one ncclAllGather and a second tiny collective after the kernel tells every rank that all writes have landed.
struct XDst {
    uint8_t* data[PMAX_COLS];
    uint8_t* nulls[PMAX_COLS];
};
template <class T>
SRB_D void scatter_col(const T* __restrict__ src, int64_t row0, const uint32_t (&d)[PTILE / 256], const unsigned long long (&pos)[PTILE / 256],
                       const XDst* __restrict__ dst, int col, bool nulls) {
    constexpr int R = PTILE / 256, U = sizeof(T) > 8 ? 4 : 8;
#pragma unroll
    for (int j0 = 0; j0 < R; j0 += U) {
        T v[U] = {};
#pragma unroll
        for (int u = 0; u < U; u++)
            if (d[j0 + u] != 0xffffffffu) v[u] = src[row0 + (j0 + u) * 256];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (d[j0 + u] != 0xffffffffu) ((T*)(nulls ? dst[d[j0 + u]].nulls[col] : dst[d[j0 + u]].data[col]))[pos[j0 + u]] = v[u];
    }
}

__global__ void __launch_bounds__(256, 2) k_part_scatter_peers(const __grid_constant__ PArgs a, int64_t n, const uint8_t* __restrict__ dest,
                                                                unsigned long long* __restrict__ cursors, const XDst* __restrict__ dst) {
    __shared__ unsigned int hist[PMAX_PARTS];
    __shared__ unsigned long long run_base[PMAX_PARTS];
    constexpr int R = PTILE / 256;
    const int64_t n_tiles = (n + PTILE - 1) / PTILE;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int i = threadIdx.x; i < PMAX_PARTS; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        uint32_t d[R], rank[R];
        const int64_t row0 = tile * PTILE + threadIdx.x;
#pragma unroll
        for (int j = 0; j < R; j++) {
            const int64_t r = row0 + j * 256;
            d[j] = 0xffffffffu;
            rank[j] = 0;
            const bool on = r < n;
            if (on) d[j] = dest[r];
            const unsigned act = __ballot_sync(0xffffffffu, on);
            if (on) {
                const unsigned peers = __match_any_sync(act, d[j]);
                const int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
                unsigned int b = 0;
                if (lane == leader) b = atomicAdd(&hist[d[j]], __popc(peers));
                b = __shfl_sync(peers, b, leader);
                rank[j] = b + __popc(peers & ((1u << lane) - 1));
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < a.n_parts; i += blockDim.x)
            run_base[i] = hist[i] ? a.base[i] + atomicAdd(&cursors[i], (unsigned long long)hist[i]) : 0;
        __syncthreads();
        unsigned long long pos[R];
#pragma unroll
        for (int j = 0; j < R; j++) pos[j] = d[j] != 0xffffffffu ? run_base[d[j]] + rank[j] : 0;
        for (int c = 0; c < a.n_cols; c++) {
            const PCol& pc = a.cols[c];
            switch (pc.width) {
            case 1: scatter_col((const uint8_t*)pc.src, row0, d, pos, dst, c, false); break;
            case 2: scatter_col((const uint16_t*)pc.src, row0, d, pos, dst, c, false); break;
            case 4: scatter_col((const uint32_t*)pc.src, row0, d, pos, dst, c, false); break;
            case 8: scatter_col((const uint64_t*)pc.src, row0, d, pos, dst, c, false); break;
            default: scatter_col((const ulonglong2*)pc.src, row0, d, pos, dst, c, false);
            }
            if (pc.src_null) scatter_col(pc.src_null, row0, d, pos, dst, c, true);
        }
        __syncthreads();
    }
}

} // namespace srb

using namespace srb;

struct srb_partitioner {
    srb_ctx* ctx;
    int n_parts;
    cudaStream_t stream = nullptr;
    DevBuf dest, totals, out_data[PMAX_COLS], out_null[PMAX_COLS];
    PinnedBuf htotals;
    std::vector<srb_column> cview;
    std::vector<int32_t> slots;
    int64_t launches = 0;
};

// ---- NCCL, bound at run time ---------------------------------------------------------------------------------------------------
namespace {
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId_ {
    char internal[128];
};
enum { ncclInt8_ = 0, ncclUint8_ = 1, ncclInt64_ = 4, ncclUint64_ = 5 };
struct Nccl {
    void* h = nullptr;
    int (*GetUniqueId)(ncclUniqueId_*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId_, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*CommAbort)(ncclComm_t) = nullptr;
    int (*CommGetAsyncError)(ncclComm_t, int*) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
Nccl g_nccl;
Nccl& nccl() {
    if (g_nccl.h) return g_nccl;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) throw Error(SRB_INTERNAL_ERROR, std::string("cannot load libnccl.so.2: ") + dlerror());
    auto sym = [&](const char* n) {
        void* p = dlsym(h, n);
        if (!p) throw Error(SRB_INTERNAL_ERROR, std::string("libnccl lacks ") + n);
        return p;
    };
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))sym("ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))sym("ncclCommInitRank");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))sym("ncclCommDestroy");
    g_nccl.CommAbort = (decltype(g_nccl.CommAbort))sym("ncclCommAbort");
    g_nccl.CommGetAsyncError = (decltype(g_nccl.CommGetAsyncError))sym("ncclCommGetAsyncError");
    g_nccl.GroupStart = (decltype(g_nccl.GroupStart))sym("ncclGroupStart");
    g_nccl.GroupEnd = (decltype(g_nccl.GroupEnd))sym("ncclGroupEnd");
    g_nccl.Send = (decltype(g_nccl.Send))sym("ncclSend");
    g_nccl.Recv = (decltype(g_nccl.Recv))sym("ncclRecv");
    g_nccl.AllGather = (decltype(g_nccl.AllGather))sym("ncclAllGather");
    g_nccl.Broadcast = (decltype(g_nccl.Broadcast))sym("ncclBroadcast");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))sym("ncclGetErrorString");
    g_nccl.h = h;
    return g_nccl;
}
#define SRB_NCCL(expr)                                                                                              \
    do {                                                                                                            \
        int _r = (expr);                                                                                            \
        if (_r != 0) throw Error(SRB_INTERNAL_ERROR, fmt("%s failed: %s", #expr, nccl().GetErrorString(_r)));        \
    } while (0)
} // namespace

struct srb_comm {
    srb_ctx* ctx;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    cudaStream_t stream = nullptr;
    DevBuf counts_dev, recv_data[PMAX_COLS], recv_null[PMAX_COLS];
    PinnedBuf counts_host;
    std::vector<srb_column> cview;
    std::vector<int32_t> slots;
    // host-chunk gather (a few intermediate rows): wire buffers and the concatenated result
    DevBuf wire_dev;
    PinnedBuf wire_host;
    struct HostOut {
        std::vector<uint8_t> data, nulls;
        std::vector<uint32_t> offsets;
    };
    std::vector<HostOut> hout;
    // fused shuffle (srb_comm_shuffle): this rank's receive arena (plain cudaMalloc: exportable with cudaIpc) and the mapped
    // arenas of the peers
    struct Peer {
        void* base = nullptr;
        size_t cap = 0;
        uint64_t gen = 0;
    };
    void* arena = nullptr;
    size_t arena_cap = 0;
    uint64_t arena_gen = 0;
    std::vector<Peer> peers;
    DevBuf x_dest, x_totals, x_dst_dev, x_hs_dev;
    PinnedBuf x_dst_host, x_hs_host;
    int64_t x_launches = 0;
    cudaEvent_t ev_in = nullptr, ev_out = nullptr; // ordering between an operator's stream and the communicator's (comm_all_gather_device)
    std::atomic<int> aborted{0}; // srb_comm_abort: the NCCL communicator is gone, every later call fails with SRB_CANCELLED
};

static inline void comm_alive(const srb_comm* c) {
    if (c->aborted.load(std::memory_order_acquire)) throw srb::Error(SRB_CANCELLED, "communicator aborted (srb_comm_abort)");
}

// the C ABI's error string lives in srb_capi.cu; exchange entry points report through the same convention
namespace srb {
void set_last_error(const std::string& s);
}
#define X_BEGIN try {
#define X_END                                 \
    }                                         \
    catch (const srb::Error& e) {             \
        srb::set_last_error(e.what());        \
        return e.code;                        \
    }                                         \
    catch (const std::exception& e) {         \
        srb::set_last_error(e.what());        \
        return SRB_INTERNAL_ERROR;            \
    }

// Host chunks (any column kind, binary included): the few intermediate rows of a low-cardinality aggregation. Each rank
// serialises its chunk as [n_rows][per column: data bytes][payload: data | nulls | offsets], the blobs travel in ONE
// fixed-slot ncclAllGather when every rank's fits SMALL_SLOT (a second, exactly sized one otherwise), and the columns are
// concatenated in rank order into buffers owned by the communicator (valid until its next call).
static constexpr size_t SMALL_SLOT = 16 << 10;

static size_t host_col_data_bytes(const srb_chunk* ch, const srb_column& sc) {
    const int w = type_width(sc.type);
    const size_t n = (size_t)ch->n_rows;
    return w ? n * (size_t)w : (n ? (size_t)(sc.offsets[n] - sc.offsets[0]) : 0);
}
static size_t host_chunk_wire_bytes(const srb_chunk* ch) {
    size_t b = 8 + (size_t)ch->n_cols * 8;
    const size_t n = (size_t)ch->n_rows;
    for (int col = 0; col < ch->n_cols; col++) {
        const srb_column& sc = ch->cols[col];
        b += host_col_data_bytes(ch, sc) + (sc.nulls ? n : 0);
        if (type_width(sc.type) == 0) b += 4 + (n + 1) * 4; // offsets, 4-byte aligned inside the blob
    }
    return (b + 15) & ~(size_t)15;
}
static void host_chunk_write(const srb_chunk* ch, uint8_t* dst) {
    const size_t n = (size_t)ch->n_rows;
    int64_t* hdr = (int64_t*)dst;
    hdr[0] = (int64_t)n;
    uint8_t* p = dst + 8 + (size_t)ch->n_cols * 8;
    for (int col = 0; col < ch->n_cols; col++) {
        const srb_column& sc = ch->cols[col];
        const bool binary = type_width(sc.type) == 0;
        const uint32_t base = (binary && n) ? sc.offsets[0] : 0;
        const size_t data = host_col_data_bytes(ch, sc);
        hdr[1 + col] = (int64_t)data | (sc.nulls ? (int64_t)1 << 62 : 0); // bit 62: a null map follows the data
        if (data) memcpy(p, (const uint8_t*)sc.data + base, data);
        p += data;
        if (sc.nulls) {
            if (n) memcpy(p, sc.nulls, n);
            p += n;
        }
        if (binary) { // offsets rebased to 0
            uint32_t* o = (uint32_t*)(dst + (((size_t)(p - dst) + 3) & ~(size_t)3));
            o[0] = 0;
            for (size_t r = 1; r <= n; r++) o[r] = sc.offsets[r] - base;
            p = (uint8_t*)(o + n + 1);
        }
    }
}
static void gather_host_chunk(srb_comm* c, const srb_chunk* mine, srb_chunk* out) {
    const int W = c->world;
    for (int col = 0; col < mine->n_cols; col++) {
        const srb_column& sc = mine->cols[col];
        if (type_width(sc.type) == 0 && mine->n_rows && !sc.offsets) throw Error(SRB_INVALID_ARGUMENT, "binary column without offsets");
        if (mine->n_rows && !sc.data && type_width(sc.type) != 0) throw Error(SRB_INVALID_ARGUMENT, "null data pointer");
    }
    const size_t need = host_chunk_wire_bytes(mine);
    size_t slot = SMALL_SLOT;
    auto exchange = [&](size_t slot_bytes, bool with_payload) {
        c->wire_host.ensure(slot_bytes * (size_t)(W + 1));
        c->wire_dev.ensure(c->ctx, slot_bytes * (size_t)(W + 1));
        uint8_t* h = c->wire_host.as<uint8_t>();
        ((int64_t*)h)[0] = (int64_t)need; // first word of the slot: this rank's blob size; second: its column count
        ((int64_t*)h)[1] = (int64_t)mine->n_cols;
        if (with_payload) host_chunk_write(mine, h + 16);
        uint8_t* d = c->wire_dev.as<uint8_t>();
        SRB_CUDA(cudaMemcpyAsync(d, h, with_payload ? 16 + need : 16, cudaMemcpyHostToDevice, c->stream));
        SRB_NCCL(nccl().AllGather(d, d + slot_bytes, slot_bytes, ncclUint8_, c->comm, c->stream));
        SRB_CUDA(cudaMemcpyAsync(h + slot_bytes, d + slot_bytes, slot_bytes * (size_t)W, cudaMemcpyDeviceToHost, c->stream));
        SRB_CUDA(cudaStreamSynchronize(c->stream));
    };
    exchange(slot, 16 + need <= SMALL_SLOT);
    size_t max_need = 0;
    for (int r = 0; r < W; r++) {
        const int64_t* head = (const int64_t*)(c->wire_host.as<uint8_t>() + slot * (size_t)(r + 1));
        max_need = std::max(max_need, (size_t)head[0]);
        if (head[1] != (int64_t)mine->n_cols) throw Error(SRB_INVALID_ARGUMENT, fmt("all-gather: rank %d sent %d columns, this rank %d", r, (int)head[1], mine->n_cols));
    }
    if (16 + max_need > SMALL_SLOT) { // some rank did not fit: every rank sees that and repeats with an exactly sized slot
        slot = (16 + max_need + 255) & ~(size_t)255;
        exchange(slot, true);
    }
    // concatenate in rank order; a column has a null map in the result when any rank sent one
    c->hout.assign((size_t)mine->n_cols, srb_comm::HostOut{});
    std::vector<char> any_nulls((size_t)mine->n_cols, 0);
    for (int r = 0; r < W; r++) {
        const int64_t* hdr = (const int64_t*)(c->wire_host.as<uint8_t>() + slot * (size_t)(r + 1) + 16);
        for (int col = 0; col < mine->n_cols; col++) any_nulls[(size_t)col] |= (char)((hdr[1 + col] >> 62) & 1);
    }
    int64_t total = 0;
    for (int r = 0; r < W; r++) {
        const uint8_t* blob = c->wire_host.as<uint8_t>() + slot * (size_t)(r + 1) + 16;
        const int64_t* hdr = (const int64_t*)blob;
        const size_t n = (size_t)hdr[0];
        const uint8_t* p = blob + 8 + (size_t)mine->n_cols * 8;
        for (int col = 0; col < mine->n_cols; col++) {
            srb_comm::HostOut& o = c->hout[(size_t)col];
            const size_t data = (size_t)(hdr[1 + col] & (((int64_t)1 << 62) - 1));
            const uint32_t byte_base = (uint32_t)o.data.size();
            o.data.insert(o.data.end(), p, p + data);
            p += data;
            if ((hdr[1 + col] >> 62) & 1) {
                o.nulls.insert(o.nulls.end(), p, p + n);
                p += n;
            } else if (any_nulls[(size_t)col]) {
                o.nulls.insert(o.nulls.end(), n, (uint8_t)0);
            }
            if (type_width(mine->cols[col].type) == 0) {
                const uint32_t* off = (const uint32_t*)(blob + (((size_t)(p - blob) + 3) & ~(size_t)3));
                if (o.offsets.empty()) o.offsets.push_back(0);
                for (size_t i = 1; i <= n; i++) o.offsets.push_back(byte_base + off[i]);
                p = (const uint8_t*)(off + n + 1);
            }
        }
        total += (int64_t)n;
    }
    c->cview.assign(mine->cols, mine->cols + mine->n_cols);
    c->slots.assign(mine->slot_ids, mine->slot_ids + mine->n_cols);
    for (int col = 0; col < mine->n_cols; col++) {
        srb_comm::HostOut& o = c->hout[(size_t)col];
        const bool binary = type_width(mine->cols[col].type) == 0;
        if (binary && o.offsets.empty()) o.offsets.push_back(0);
        const size_t used = o.data.size();
        o.data.resize(used + 16); // BinaryColumn readers may touch 16 bytes past the end (column/bytes.h:27)
        c->cview[col].data = o.data.data();
        c->cview[col].nulls = any_nulls[(size_t)col] ? o.nulls.data() : nullptr;
        c->cview[col].offsets = binary ? o.offsets.data() : nullptr;
        c->cview[col].n_bytes = binary ? (int64_t)used : 0;
        c->cview[col].mem_kind = SRB_MEM_HOST;
    }
    out->n_rows = total;
    out->n_cols = mine->n_cols;
    out->slot_ids = c->slots.data();
    out->cols = c->cview.data();
}

extern "C" {

srb_status srb_partition_create(srb_ctx* ctx, int32_t n_partitions, int32_t reserved, srb_partitioner** out) {
    X_BEGIN
    (void)reserved;
    if (!ctx || !out) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    if (n_partitions < 1 || n_partitions > PMAX_PARTS) throw Error(SRB_INVALID_ARGUMENT, fmt("1..%d partitions supported", PMAX_PARTS));
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    auto* p = new srb_partitioner();
    p->ctx = ctx;
    p->n_parts = n_partitions;
    SRB_CUDA(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    p->totals.ensure(ctx, 2 * PMAX_PARTS * 8);
    p->htotals.ensure(PMAX_PARTS * 8);
    *out = p;
    return SRB_OK;
    X_END
}

void srb_partition_destroy(srb_partitioner* p) {
    if (!p) return;
    cudaSetDevice(p->ctx->cfg.device);
    if (p->stream) {
        cudaStreamSynchronize(p->stream);
        cudaStreamDestroy(p->stream);
    }
    delete p;
}

// in: device chunk of fixed-width columns; out: same columns with rows grouped by destination (library-owned device
// buffers, valid until the next run/destroy); counts[d] = rows for destination d.
srb_status srb_partition_run(srb_partitioner* p, const srb_chunk* in, int32_t n_key_slots, const int32_t* key_slots, srb_chunk* out,
                             int64_t* counts) {
    X_BEGIN
    if (!p || !in || !out || !counts || !key_slots) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    if (n_key_slots < 1 || n_key_slots > PMAX_KEYS) throw Error(SRB_INVALID_ARGUMENT, "bad number of partition columns");
    if (in->n_cols < 1 || in->n_cols > PMAX_COLS) throw Error(SRB_NOT_IMPLEMENTED, "too many columns");
    SRB_CUDA(cudaSetDevice(p->ctx->cfg.device));
    const int64_t n = in->n_rows;
    PArgs a{};
    a.n_cols = in->n_cols;
    a.n_keys = n_key_slots;
    a.n_parts = p->n_parts;
    p->cview.assign(in->cols, in->cols + in->n_cols);
    p->slots.assign(in->slot_ids, in->slot_ids + in->n_cols);
    for (int c = 0; c < in->n_cols; c++) {
        const srb_column& col = in->cols[c];
        int w = type_width(col.type);
        if (w == 0) throw Error(SRB_NOT_IMPLEMENTED, "partitioning of binary columns");
        if (col.mem_kind != SRB_MEM_DEVICE) throw Error(SRB_INVALID_ARGUMENT, "partitioner takes device-resident chunks");
        p->out_data[c].ensure(p->ctx, (size_t)std::max<int64_t>(n, 1) * w);
        if (col.nulls) p->out_null[c].ensure(p->ctx, (size_t)std::max<int64_t>(n, 1));
        a.cols[c] = PCol{(const uint8_t*)col.data, p->out_data[c].as<uint8_t>(), col.nulls, col.nulls ? p->out_null[c].as<uint8_t>() : nullptr, w, 0};
        p->cview[c].data = p->out_data[c].p;
        p->cview[c].nulls = col.nulls ? p->out_null[c].as<uint8_t>() : nullptr;
    }
    for (int k = 0; k < n_key_slots; k++) {
        int idx = -1;
        for (int c = 0; c < in->n_cols; c++)
            if (in->slot_ids[c] == key_slots[k]) idx = c;
        if (idx < 0) throw Error(SRB_INVALID_ARGUMENT, fmt("partition slot %d not in chunk", key_slots[k]));
        a.key_col[k] = idx;
    }
    unsigned long long* totals = p->totals.as<unsigned long long>();
    unsigned long long* cursors = totals + PMAX_PARTS;
    SRB_CUDA(cudaMemsetAsync(totals, 0, 2 * PMAX_PARTS * 8, p->stream));
    unsigned long long* ht = p->htotals.as<unsigned long long>();
    memset(ht, 0, PMAX_PARTS * 8);
    if (n > 0) {
        p->dest.ensure(p->ctx, (size_t)n);
        int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)p->ctx->n_sms * 8));
        k_part_count<<<grid, 256, 0, p->stream>>>(a, n, p->dest.as<uint8_t>(), totals);
        SRB_CUDA(cudaMemcpyAsync(ht, totals, PMAX_PARTS * 8, cudaMemcpyDeviceToHost, p->stream));
        SRB_CUDA(cudaStreamSynchronize(p->stream));
        unsigned long long acc = 0;
        for (int d = 0; d < p->n_parts; d++) {
            a.base[d] = acc;
            acc += ht[d];
        }
        int grid2 = (int)std::max<int64_t>(1, std::min<int64_t>((n + PTILE - 1) / PTILE, (int64_t)p->ctx->n_sms * 4));
        k_part_scatter<<<grid2, 256, 0, p->stream>>>(a, n, p->dest.as<uint8_t>(), cursors);
        SRB_CUDA(cudaGetLastError());
        SRB_CUDA(cudaStreamSynchronize(p->stream));
        p->launches += 2;
    }
    for (int d = 0; d < p->n_parts; d++) counts[d] = (int64_t)ht[d];
    out->n_rows = n;
    out->n_cols = in->n_cols;
    out->slot_ids = p->slots.data();
    out->cols = p->cview.data();
    return SRB_OK;
    X_END
}

// ---- communicator ------------------------------------------------------------------------------------------------------------
srb_status srb_comm_unique_id(void* out128) {
    X_BEGIN
    if (!out128) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    ncclUniqueId_ id;
    SRB_NCCL(nccl().GetUniqueId(&id));
    memcpy(out128, &id, 128);
    return SRB_OK;
    X_END
}

srb_status srb_comm_create(srb_ctx* ctx, const void* id128, srb_comm** out) {
    X_BEGIN
    if (!ctx || !id128 || !out) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    if (ctx->cfg.world_size > PMAX_PARTS) throw Error(SRB_NOT_IMPLEMENTED, "world size");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    auto* c = new srb_comm();
    c->ctx = ctx;
    c->rank = ctx->cfg.rank;
    c->world = ctx->cfg.world_size;
    ncclUniqueId_ id;
    memcpy(&id, id128, 128);
    SRB_NCCL(nccl().CommInitRank(&c->comm, c->world, id, c->rank));
    SRB_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    c->counts_dev.ensure(ctx, (size_t)(c->world + c->world * c->world) * 8);
    c->counts_host.ensure((size_t)(c->world + c->world * c->world) * 8);
    *out = c;
    return SRB_OK;
    X_END
}

void srb_comm_destroy(srb_comm* c) {
    if (!c) return;
    cudaSetDevice(c->ctx->cfg.device);
    if (c->stream) {
        cudaStreamSynchronize(c->stream);
        cudaStreamDestroy(c->stream);
    }
    for (int d = 0; d < (int)c->peers.size(); d++)
        if (d != c->rank && c->peers[d].base) cudaIpcCloseMemHandle(c->peers[d].base);
    if (c->arena) cudaFree(c->arena);
    if (c->ev_in) cudaEventDestroy(c->ev_in);
    if (c->ev_out) cudaEventDestroy(c->ev_out);
    if (c->comm && !c->aborted.load()) nccl().CommDestroy(c->comm);
    delete c;
}

void* srb_comm_stream(srb_comm* c) { return c ? (void*)c->stream : nullptr; }

// Cancellation that reaches in-flight collectives (PipelineDriver cancels a fragment's drivers, pipeline_driver.cpp:316; the brpc exchange
// of the reference fails its pending RPCs): ncclCommAbort makes every collective queued or running on this communicator return, on this
// rank, so a peer that died or a cancelled query cannot leave the stream blocked. May be called from another thread than the one inside a
// collective. Afterwards every call on the communicator fails with SRB_CANCELLED; only srb_comm_destroy is allowed.
srb_status srb_comm_abort(srb_comm* c) {
    X_BEGIN
    if (!c) throw Error(SRB_INVALID_ARGUMENT, "null communicator");
    if (c->aborted.exchange(1)) return SRB_OK;
    if (c->comm) SRB_NCCL(nccl().CommAbort(c->comm));
    return SRB_OK;
    X_END
}
// the communicator's asynchronous error state (ncclCommGetAsyncError): SRB_OK, SRB_CANCELLED after an abort, SRB_INTERNAL_ERROR otherwise
srb_status srb_comm_check(srb_comm* c) {
    X_BEGIN
    if (!c) throw Error(SRB_INVALID_ARGUMENT, "null communicator");
    comm_alive(c);
    int err = 0;
    SRB_NCCL(nccl().CommGetAsyncError(c->comm, &err));
    if (err) throw Error(SRB_INTERNAL_ERROR, fmt("NCCL asynchronous error: %s", nccl().GetErrorString(err)));
    return SRB_OK;
    X_END
}
} // extern "C" (reopened below)

namespace srb {
// Fixed-size all-gather of device bytes ordered against a caller's stream, no host synchronisation: `send` (bytes_per_rank) of every rank
// lands in recv[rank * bytes_per_rank ...] on every rank. Work queued on `user_stream` before the call happens before the collective,
// work queued after it sees the gathered bytes. (The gather exchange of a two-phase low-cardinality aggregation: see
// ScanAggOp::set_finishing_merged.)
void comm_all_gather_device(srb_comm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t user_stream) {
    if (!c) throw Error(SRB_INVALID_ARGUMENT, "null communicator");
    comm_alive(c);
    if (!c->ev_in) {
        SRB_CUDA(cudaEventCreateWithFlags(&c->ev_in, cudaEventDisableTiming));
        SRB_CUDA(cudaEventCreateWithFlags(&c->ev_out, cudaEventDisableTiming));
    }
    SRB_CUDA(cudaEventRecord(c->ev_in, user_stream));
    SRB_CUDA(cudaStreamWaitEvent(c->stream, c->ev_in, 0));
    SRB_NCCL(nccl().AllGather(send, recv, bytes_per_rank, ncclUint8_, c->comm, c->stream));
    SRB_CUDA(cudaEventRecord(c->ev_out, c->stream));
    SRB_CUDA(cudaStreamWaitEvent(user_stream, c->ev_out, 0));
    c->x_launches++;
}
int comm_world(const srb_comm* c) { return c->world; }
int comm_rank(const srb_comm* c) { return c->rank; }
} // namespace srb

extern "C" {

// send: device chunk whose rows are grouped by destination rank (srb_partition_run output), send_counts[world].
// recv: library-owned device chunk holding the rows received from rank 0, 1, ... in that order; recv_counts[world].
srb_status srb_comm_all_to_all(srb_comm* c, const srb_chunk* send, const int64_t* send_counts, srb_chunk* recv, int64_t* recv_counts) {
    X_BEGIN
    if (!c || !send || !send_counts || !recv || !recv_counts) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    if (send->n_cols > PMAX_COLS) throw Error(SRB_NOT_IMPLEMENTED, "too many columns");
    comm_alive(c);
    SRB_CUDA(cudaSetDevice(c->ctx->cfg.device));
    const int W = c->world;
    // counts: every rank learns the whole W x W matrix
    int64_t* hc = c->counts_host.as<int64_t>();
    int64_t* dc = c->counts_dev.as<int64_t>();
    for (int d = 0; d < W; d++) hc[d] = send_counts[d];
    SRB_CUDA(cudaMemcpyAsync(dc, hc, (size_t)W * 8, cudaMemcpyHostToDevice, c->stream));
    SRB_NCCL(nccl().AllGather(dc, dc + W, (size_t)W, ncclInt64_, c->comm, c->stream));
    SRB_CUDA(cudaMemcpyAsync(hc + W, dc + W, (size_t)W * W * 8, cudaMemcpyDeviceToHost, c->stream));
    SRB_CUDA(cudaStreamSynchronize(c->stream));
    int64_t total = 0;
    std::vector<int64_t> roff(W), soff(W);
    int64_t sacc = 0;
    for (int s = 0; s < W; s++) {
        recv_counts[s] = hc[W + (size_t)s * W + c->rank]; // what rank s sends to me
        roff[s] = total;
        total += recv_counts[s];
        soff[s] = sacc;
        sacc += send_counts[s];
    }
    if (sacc != send->n_rows) throw Error(SRB_INVALID_ARGUMENT, "send_counts do not add up to the chunk's rows");
    c->cview.assign(send->cols, send->cols + send->n_cols);
    c->slots.assign(send->slot_ids, send->slot_ids + send->n_cols);
    SRB_NCCL(nccl().GroupStart());
    for (int col = 0; col < send->n_cols; col++) {
        const srb_column& sc = send->cols[col];
        const int w = type_width(sc.type);
        if (w == 0 || sc.mem_kind != SRB_MEM_DEVICE) throw Error(SRB_INVALID_ARGUMENT, "all-to-all takes device-resident fixed-width columns");
        c->recv_data[col].ensure(c->ctx, (size_t)std::max<int64_t>(total, 1) * w);
        if (sc.nulls) c->recv_null[col].ensure(c->ctx, (size_t)std::max<int64_t>(total, 1));
        for (int peer = 0; peer < W; peer++) {
            if (send_counts[peer])
                SRB_NCCL(nccl().Send((const uint8_t*)sc.data + (size_t)soff[peer] * w, (size_t)send_counts[peer] * w, ncclUint8_, peer, c->comm, c->stream));
            if (recv_counts[peer])
                SRB_NCCL(nccl().Recv(c->recv_data[col].as<uint8_t>() + (size_t)roff[peer] * w, (size_t)recv_counts[peer] * w, ncclUint8_, peer, c->comm,
                                     c->stream));
            if (sc.nulls) {
                if (send_counts[peer]) SRB_NCCL(nccl().Send(sc.nulls + soff[peer], (size_t)send_counts[peer], ncclUint8_, peer, c->comm, c->stream));
                if (recv_counts[peer])
                    SRB_NCCL(nccl().Recv(c->recv_null[col].as<uint8_t>() + roff[peer], (size_t)recv_counts[peer], ncclUint8_, peer, c->comm, c->stream));
            }
        }
        c->cview[col].data = c->recv_data[col].p;
        c->cview[col].nulls = sc.nulls ? c->recv_null[col].as<uint8_t>() : nullptr;
    }
    SRB_NCCL(nccl().GroupEnd());
    SRB_CUDA(cudaStreamSynchronize(c->stream));
    recv->n_rows = total;
    recv->n_cols = send->n_cols;
    recv->slot_ids = c->slots.data();
    recv->cols = c->cview.data();
    return SRB_OK;
    X_END
}

// HASH_PARTITIONED exchange in one step (ExchangeSinkOperator::push_chunk -> transmit -> ExchangeSourceOperator, exchange_sink_operator.cpp:
// 569-633): count pass, one ncclAllGather of the W x W count matrix, then the scatter kernel writes every row straight into its
// destination rank's receive arena over peer memory (k_part_scatter_peers). Output as srb_comm_all_to_all.
namespace {
struct XHandshake {
    cudaIpcMemHandle_t handle;
    uint64_t cap, gen;
};
inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }
} // namespace

srb_status srb_comm_shuffle(srb_comm* c, const srb_chunk* in, int32_t n_key_slots, const int32_t* key_slots, srb_chunk* recv, int64_t* recv_counts) {
    X_BEGIN
    if (!c || !in || !key_slots || !recv || !recv_counts) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    if (n_key_slots < 1 || n_key_slots > PMAX_KEYS) throw Error(SRB_INVALID_ARGUMENT, "bad number of partition columns");
    if (in->n_cols < 1 || in->n_cols > PMAX_COLS) throw Error(SRB_NOT_IMPLEMENTED, "too many columns");
    comm_alive(c);
    SRB_CUDA(cudaSetDevice(c->ctx->cfg.device));
    const int W = c->world, me = c->rank;
    const int64_t n = in->n_rows;
    if (c->peers.empty()) c->peers.resize(W);
    PArgs a{};
    a.n_cols = in->n_cols;
    a.n_keys = n_key_slots;
    a.n_parts = W;
    uint64_t nullmask = 0;
    int width[PMAX_COLS];
    for (int col = 0; col < in->n_cols; col++) {
        const srb_column& sc = in->cols[col];
        width[col] = type_width(sc.type);
        if (width[col] == 0) throw Error(SRB_NOT_IMPLEMENTED, "partitioning of binary columns");
        if (n > 0 && sc.mem_kind != SRB_MEM_DEVICE) throw Error(SRB_INVALID_ARGUMENT, "the shuffle takes device-resident chunks");
        a.cols[col] = PCol{(const uint8_t*)sc.data, nullptr, sc.nulls, nullptr, width[col], 0};
        if (sc.nulls) nullmask |= 1ull << col;
    }
    for (int k = 0; k < n_key_slots; k++) {
        int idx = -1;
        for (int col = 0; col < in->n_cols; col++)
            if (in->slot_ids[col] == key_slots[k]) idx = col;
        if (idx < 0) throw Error(SRB_INVALID_ARGUMENT, fmt("partition slot %d not in chunk", key_slots[k]));
        a.key_col[k] = idx;
    }
    // pass 1: destination per row + totals per destination
    c->x_totals.ensure(c->ctx, 2 * PMAX_PARTS * 8);
    unsigned long long* totals = c->x_totals.as<unsigned long long>();
    unsigned long long* cursors = totals + PMAX_PARTS;
    SRB_CUDA(cudaMemsetAsync(totals, 0, 2 * PMAX_PARTS * 8, c->stream));
    c->x_dest.ensure(c->ctx, (size_t)std::max<int64_t>(n, 1));
    if (n > 0) {
        int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)c->ctx->n_sms * 8));
        k_part_count<<<grid, 256, 0, c->stream>>>(a, n, c->x_dest.as<uint8_t>(), totals);
        c->x_launches++;
    }
    // counts (+ this rank's null-map mask in one extra word): every rank learns the whole matrix
    const int RW = W + 1; // words per rank
    c->counts_dev.ensure(c->ctx, (size_t)(RW + RW * W) * 8);
    c->counts_host.ensure((size_t)(RW + RW * W) * 8);
    int64_t* hc = c->counts_host.as<int64_t>();
    int64_t* dc = c->counts_dev.as<int64_t>();
    hc[W] = (int64_t)nullmask;
    SRB_CUDA(cudaMemcpyAsync(dc, totals, (size_t)W * 8, cudaMemcpyDeviceToDevice, c->stream));
    SRB_CUDA(cudaMemcpyAsync(dc + W, hc + W, 8, cudaMemcpyHostToDevice, c->stream));
    SRB_NCCL(nccl().AllGather(dc, dc + RW, (size_t)RW, ncclInt64_, c->comm, c->stream));
    SRB_CUDA(cudaMemcpyAsync(hc + RW, dc + RW, (size_t)RW * W * 8, cudaMemcpyDeviceToHost, c->stream));
    SRB_CUDA(cudaStreamSynchronize(c->stream));
    auto M = [&](int s, int d) { return hc[RW + (size_t)s * RW + d]; };
    std::vector<int64_t> T(W, 0), off(W, 0);
    for (int d = 0; d < W; d++)
        for (int s = 0; s < W; s++) {
            if (s < me) off[d] += M(s, d);
            T[d] += M(s, d);
        }
    for (int s = 0; s < W; s++) {
        recv_counts[s] = M(s, me);
        if ((uint64_t)hc[RW + (size_t)s * RW + W] != nullmask) throw Error(SRB_NOT_IMPLEMENTED, "null maps differ between the ranks of a shuffle");
    }
    // receive layout of rank d: columns back to back (256-byte aligned), then the null maps
    auto layout = [&](int d, size_t* data_off, size_t* null_off) {
        size_t pos = 0;
        for (int col = 0; col < in->n_cols; col++) {
            data_off[col] = pos;
            pos += align256((size_t)T[d] * width[col]);
        }
        for (int col = 0; col < in->n_cols; col++) {
            null_off[col] = pos;
            if (nullmask >> col & 1) pos += align256((size_t)T[d]);
        }
        return std::max<size_t>(pos, 256);
    };
    size_t doff[PMAX_COLS], noff[PMAX_COLS];
    bool handshake = false;
    for (int d = 0; d < W; d++) handshake = handshake || layout(d, doff, noff) > (d == me ? c->arena_cap : c->peers[d].cap);
    void* old_arena = nullptr;
    if (handshake) { // some rank grows its arena: everyone republishes (the decision is the same on every rank)
        const size_t need = layout(me, doff, noff);
        if (need > c->arena_cap) {
            old_arena = c->arena;
            size_t cap = std::max(need, c->arena_cap + c->arena_cap / 2);
            cap = (cap + ((size_t)2 << 20) - 1) & ~(((size_t)2 << 20) - 1);
            SRB_CUDA(cudaMalloc(&c->arena, cap));
            c->arena_cap = cap;
            c->arena_gen++;
        }
        c->x_hs_host.ensure(sizeof(XHandshake) * (size_t)(W + 1));
        c->x_hs_dev.ensure(c->ctx, sizeof(XHandshake) * (size_t)(W + 1));
        XHandshake* hs = c->x_hs_host.as<XHandshake>();
        memset(&hs[0], 0, sizeof(XHandshake));
        if (W > 1) SRB_CUDA(cudaIpcGetMemHandle(&hs[0].handle, c->arena));
        hs[0].cap = c->arena_cap;
        hs[0].gen = c->arena_gen;
        XHandshake* hd = c->x_hs_dev.as<XHandshake>();
        SRB_CUDA(cudaMemcpyAsync(hd, hs, sizeof(XHandshake), cudaMemcpyHostToDevice, c->stream));
        SRB_NCCL(nccl().AllGather(hd, hd + 1, sizeof(XHandshake), ncclUint8_, c->comm, c->stream));
        SRB_CUDA(cudaMemcpyAsync(hs + 1, hd + 1, sizeof(XHandshake) * (size_t)W, cudaMemcpyDeviceToHost, c->stream));
        SRB_CUDA(cudaStreamSynchronize(c->stream));
        for (int d = 0; d < W; d++) {
            srb_comm::Peer& p = c->peers[d];
            const XHandshake& r = hs[1 + d];
            if (d == me) {
                p.base = c->arena;
            } else if (r.gen != p.gen || !p.base) {
                if (p.base) SRB_CUDA(cudaIpcCloseMemHandle(p.base));
                p.base = nullptr;
                cudaError_t e = cudaIpcOpenMemHandle(&p.base, r.handle, cudaIpcMemLazyEnablePeerAccess);
                if (e != cudaSuccess) {
                    cudaGetLastError();
                    throw Error(SRB_INTERNAL_ERROR, fmt("cannot map the receive arena of rank %d (%s): the fused shuffle needs NVLink/PCIe peer access "
                                                        "between the GPUs; use srb_partition_run + srb_comm_all_to_all instead", d, cudaGetErrorString(e)));
                }
            }
            p.cap = r.cap;
            p.gen = r.gen;
        }
    }
    c->peers[me].base = c->arena;
    // destination table + first row of this source inside every destination
    c->x_dst_host.ensure(sizeof(XDst) * (size_t)W);
    c->x_dst_dev.ensure(c->ctx, sizeof(XDst) * (size_t)W);
    XDst* xh = c->x_dst_host.as<XDst>();
    for (int d = 0; d < W; d++) {
        layout(d, doff, noff);
        uint8_t* base = (uint8_t*)c->peers[d].base;
        for (int col = 0; col < PMAX_COLS; col++) {
            xh[d].data[col] = col < in->n_cols ? base + doff[col] : nullptr;
            xh[d].nulls[col] = col < in->n_cols && (nullmask >> col & 1) ? base + noff[col] : nullptr;
        }
        a.base[d] = (unsigned long long)off[d];
    }
    if (n > 0) {
        SRB_CUDA(cudaMemcpyAsync(c->x_dst_dev.p, xh, sizeof(XDst) * (size_t)W, cudaMemcpyHostToDevice, c->stream));
        int grid2 = (int)std::max<int64_t>(1, std::min<int64_t>((n + PTILE - 1) / PTILE, (int64_t)c->ctx->n_sms * 2));
        k_part_scatter_peers<<<grid2, 256, 0, c->stream>>>(a, n, c->x_dest.as<uint8_t>(), cursors, c->x_dst_dev.as<XDst>());
        SRB_CUDA(cudaGetLastError());
        c->x_launches++;
    }
    // every rank's scatter has finished (its stores are visible) once this collective, queued behind the kernel, completes everywhere
    SRB_NCCL(nccl().AllGather(dc, dc + RW, 1, ncclInt64_, c->comm, c->stream));
    SRB_CUDA(cudaStreamSynchronize(c->stream));
    if (old_arena) SRB_CUDA(cudaFree(old_arena)); // every peer re-mapped the new arena before it reached the collective above
    layout(me, doff, noff);
    c->cview.assign(in->cols, in->cols + in->n_cols);
    c->slots.assign(in->slot_ids, in->slot_ids + in->n_cols);
    for (int col = 0; col < in->n_cols; col++) {
        c->cview[col].mem_kind = SRB_MEM_DEVICE;
        c->cview[col].data = (uint8_t*)c->arena + doff[col];
        c->cview[col].nulls = (nullmask >> col & 1) ? (uint8_t*)c->arena + noff[col] : nullptr;
        c->cview[col].offsets = nullptr;
    }
    recv->n_rows = T[me];
    recv->n_cols = in->n_cols;
    recv->slot_ids = c->slots.data();
    recv->cols = c->cview.data();
    return SRB_OK;
    X_END
}

// Global runtime filter (RuntimeFilterMerger / SimdBlockFilter::merge, exprs/runtime_filter.cpp): the partial filters of all ranks,
// built with the same expected_rows, are all-gathered and ORed bucket by bucket; min / max / has_null / inserted combine on the host.
namespace srb {
__global__ void __launch_bounds__(256) k_rf_or(uint32_t* __restrict__ dir, const uint32_t* __restrict__ all, int64_t words, int world) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t v = 0;
        for (int r = 0; r < world; r++) v |= all[(size_t)r * words + i];
        dir[i] = v;
    }
}
} // namespace srb

srb_status srb_comm_merge_runtime_filter(srb_comm* c, srb_runtime_filter* rf) {
    X_BEGIN
    if (!c || !rf) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    comm_alive(c);
    // "RuntimeBitsetFilter only support for broadcast join, so merge and concat are not supported" (exprs/runtime_filter.h:1686-1688)
    if (rf->is_bitset && c->world > 1) throw Error(SRB_INVALID_ARGUMENT, "a bitset runtime filter cannot be merged: build partial filters without SRB_RF_MAYBE_BITSET");
    SRB_CUDA(cudaSetDevice(c->ctx->cfg.device));
    const int W = c->world;
    c->counts_dev.ensure(c->ctx, (size_t)(5 + 5 * W) * 8);
    c->counts_host.ensure((size_t)(5 + 5 * W) * 8);
    int64_t* hc = c->counts_host.as<int64_t>();
    int64_t* dc = c->counts_dev.as<int64_t>();
    hc[0] = rf->log_num_buckets;
    hc[1] = rf->min;
    hc[2] = rf->max;
    hc[3] = rf->has_null ? 1 : 0;
    hc[4] = rf->inserted;
    hc[0] |= (int64_t)rf->mix << 32; // filters hashed differently cannot be merged either
    SRB_CUDA(cudaMemcpyAsync(dc, hc, 5 * 8, cudaMemcpyHostToDevice, c->stream));
    SRB_NCCL(nccl().AllGather(dc, dc + 5, 5, ncclInt64_, c->comm, c->stream));
    SRB_CUDA(cudaMemcpyAsync(hc + 5, dc + 5, (size_t)5 * W * 8, cudaMemcpyDeviceToHost, c->stream));
    SRB_CUDA(cudaStreamSynchronize(c->stream));
    int64_t mn = INT64_MAX, mx = INT64_MIN, ins = 0;
    bool has_null = false;
    for (int r = 0; r < W; r++) {
        const int64_t* m = hc + 5 + (size_t)r * 5;
        if (m[0] != hc[0])
            throw Error(SRB_INVALID_ARGUMENT, "runtime filters of different sizes / hash modes cannot be merged: build them with the same expected_rows and hash_mix");
        mn = std::min(mn, m[1]);
        mx = std::max(mx, m[2]);
        has_null = has_null || m[3] != 0;
        ins += m[4];
    }
    if (W > 1) {
        const size_t bytes = rf->dir_bytes();
        DevBuf all;
        all.ensure(c->ctx, bytes * (size_t)W);
        SRB_NCCL(nccl().AllGather(rf->dir.p, all.p, bytes, ncclUint8_, c->comm, c->stream));
        const int64_t words = (int64_t)(bytes / 4);
        int grid = (int)std::max<int64_t>(1, std::min<int64_t>((words + 255) / 256, (int64_t)c->ctx->n_sms * 8));
        k_rf_or<<<grid, 256, 0, c->stream>>>((uint32_t*)rf->dir.p, all.as<uint32_t>(), words, W);
        SRB_CUDA(cudaGetLastError());
        SRB_CUDA(cudaStreamSynchronize(c->stream));
        c->x_launches++;
    }
    rf->min = mn;
    rf->max = mx;
    rf->has_null = has_null;
    rf->inserted = ins;
    return SRB_OK;
    X_END
}

srb_status srb_comm_counters(srb_comm* c, int64_t* kernel_launches) {
    if (!c || !kernel_launches) return SRB_INVALID_ARGUMENT;
    *kernel_launches = c->x_launches;
    return SRB_OK;
}

// every rank contributes `mine` (device, fixed width, same schema); out holds all ranks' rows, rank 0 first. Used to replicate
// a filtered build side (broadcast join) and to gather small partial aggregates.
srb_status srb_comm_all_gather(srb_comm* c, const srb_chunk* mine, srb_chunk* out) {
    X_BEGIN
    if (!c || !mine || !out) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    comm_alive(c);
    SRB_CUDA(cudaSetDevice(c->ctx->cfg.device));
    const int W = c->world;
    bool all_host = mine->n_cols > 0;
    for (int col = 0; col < mine->n_cols; col++) all_host = all_host && mine->cols[col].mem_kind == SRB_MEM_HOST;
    if (all_host) {
        gather_host_chunk(c, mine, out);
        return SRB_OK;
    }
    int64_t* hc = c->counts_host.as<int64_t>();
    int64_t* dc = c->counts_dev.as<int64_t>();
    hc[0] = mine->n_rows;
    SRB_CUDA(cudaMemcpyAsync(dc, hc, 8, cudaMemcpyHostToDevice, c->stream));
    SRB_NCCL(nccl().AllGather(dc, dc + W, 1, ncclInt64_, c->comm, c->stream));
    SRB_CUDA(cudaMemcpyAsync(hc + W, dc + W, (size_t)W * 8, cudaMemcpyDeviceToHost, c->stream));
    SRB_CUDA(cudaStreamSynchronize(c->stream));
    std::vector<int64_t> off(W);
    int64_t total = 0;
    for (int r = 0; r < W; r++) {
        off[r] = total;
        total += hc[W + r];
    }
    c->cview.assign(mine->cols, mine->cols + mine->n_cols);
    c->slots.assign(mine->slot_ids, mine->slot_ids + mine->n_cols);
    SRB_NCCL(nccl().GroupStart());
    for (int col = 0; col < mine->n_cols; col++) {
        const srb_column& sc = mine->cols[col];
        const int w = type_width(sc.type);
        if (w == 0 || sc.mem_kind != SRB_MEM_DEVICE) throw Error(SRB_INVALID_ARGUMENT, "all-gather takes device-resident fixed-width columns");
        c->recv_data[col].ensure(c->ctx, (size_t)std::max<int64_t>(total, 1) * w);
        if (sc.nulls) c->recv_null[col].ensure(c->ctx, (size_t)std::max<int64_t>(total, 1));
        for (int r = 0; r < W; r++) { // one broadcast per root: variable-size all-gather
            if (hc[W + r] == 0) continue;
            const void* src = r == c->rank ? sc.data : nullptr;
            SRB_NCCL(nccl().Broadcast(src ? src : c->recv_data[col].as<uint8_t>() + (size_t)off[r] * w, c->recv_data[col].as<uint8_t>() + (size_t)off[r] * w,
                                      (size_t)hc[W + r] * w, ncclUint8_, r, c->comm, c->stream));
            if (sc.nulls) {
                const void* ns = r == c->rank ? sc.nulls : nullptr;
                SRB_NCCL(nccl().Broadcast(ns ? ns : c->recv_null[col].as<uint8_t>() + off[r], c->recv_null[col].as<uint8_t>() + off[r], (size_t)hc[W + r],
                                          ncclUint8_, r, c->comm, c->stream));
            }
        }
        c->cview[col].data = c->recv_data[col].p;
        c->cview[col].nulls = sc.nulls ? c->recv_null[col].as<uint8_t>() : nullptr;
    }
    SRB_NCCL(nccl().GroupEnd());
    SRB_CUDA(cudaStreamSynchronize(c->stream));
    out->n_rows = total;
    out->n_cols = mine->n_cols;
    out->slot_ids = c->slots.data();
    out->cols = c->cview.data();
    return SRB_OK;
    X_END
}

} // extern "C"

// ---- srb_partition_run_noalloc: the partitioner for callers that cannot allocate ------------------------------------------------
// No counterpart in the reference: the same routing as srb_partition_run, but the pass-1
// scratch (one destination byte per row, the per-destination totals, cursors and first rows) lives in a file-scope __device__ array of
// fixed size, and the first output row of every destination is derived on the device, so the call allocates nothing, never waits for
// the device and never copies to the host. Every call on a device shares the one array: calls that overlap on the GPU race.
namespace srb {
constexpr int64_t NA_MAX_ROWS = SRB_PARTITION_NOALLOC_MAX_ROWS;
constexpr size_t NA_HEAD = 2048; // totals[64] | cursors[64] | base[64] (u64 each), then the destination bytes
static __device__ __align__(256) uint8_t g_na_scratch[NA_HEAD + NA_MAX_ROWS];

// totals -> first output row per destination (exclusive scan) and the caller's counts
__global__ void k_part_bases_na(const unsigned long long* __restrict__ totals, unsigned long long* __restrict__ base, int64_t* __restrict__ counts,
                                int n_parts) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        unsigned long long acc = 0;
        for (int d = 0; d < n_parts; d++) {
            base[d] = acc;
            counts[d] = (int64_t)totals[d];
            acc += totals[d];
        }
    }
}

// k_part_scatter with the destinations' first rows read from device memory instead of the launch arguments
__global__ void __launch_bounds__(256) k_part_scatter_na(const __grid_constant__ PArgs a, int64_t n, const uint8_t* __restrict__ dest,
                                                          unsigned long long* __restrict__ cursors, const unsigned long long* __restrict__ base) {
    __shared__ unsigned int hist[PMAX_PARTS];
    __shared__ unsigned long long run_base[PMAX_PARTS];
    constexpr int R = PTILE / 256;
    const int64_t n_tiles = (n + PTILE - 1) / PTILE;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int i = threadIdx.x; i < PMAX_PARTS; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        uint32_t d[R], rank[R];
        const int64_t row0 = tile * PTILE + threadIdx.x;
#pragma unroll
        for (int j = 0; j < R; j++) {
            const int64_t r = row0 + j * 256;
            d[j] = 0xffffffffu;
            rank[j] = 0;
            const bool on = r < n;
            if (on) d[j] = dest[r];
            const unsigned act = __ballot_sync(0xffffffffu, on);
            if (on) {
                const unsigned peers = __match_any_sync(act, d[j]);
                const int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
                unsigned int b = 0;
                if (lane == leader) b = atomicAdd(&hist[d[j]], __popc(peers));
                b = __shfl_sync(peers, b, leader);
                rank[j] = b + __popc(peers & ((1u << lane) - 1));
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < a.n_parts; i += blockDim.x)
            run_base[i] = hist[i] ? base[i] + atomicAdd(&cursors[i], (unsigned long long)hist[i]) : 0;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < R; j++) {
            if (d[j] == 0xffffffffu) continue;
            const int64_t r = row0 + j * 256;
            const unsigned long long pos = run_base[d[j]] + rank[j];
            for (int c = 0; c < a.n_cols; c++) {
                const PCol& pc = a.cols[c];
                switch (pc.width) {
                case 1: pc.dst[pos] = pc.src[r]; break;
                case 2: ((uint16_t*)pc.dst)[pos] = ((const uint16_t*)pc.src)[r]; break;
                case 4: ((uint32_t*)pc.dst)[pos] = ((const uint32_t*)pc.src)[r]; break;
                case 8: ((uint64_t*)pc.dst)[pos] = ((const uint64_t*)pc.src)[r]; break;
                default: ((ulonglong2*)pc.dst)[pos] = ((const ulonglong2*)pc.src)[r];
                }
                if (pc.src_null) pc.dst_null[pos] = pc.src_null[r];
            }
        }
        __syncthreads();
    }
}

// The array's name is a host-side stand-in in host code: its device address is looked up once per device, with that device current.
static std::mutex g_na_mu;
static uint8_t* g_na_ptr[64] = {};
static int g_na_sms[64] = {};
static uint8_t* na_scratch(int device, int* n_sms) {
    if (device < 0 || device >= 64) throw Error(SRB_INVALID_ARGUMENT, "device ordinal");
    std::lock_guard<std::mutex> g(g_na_mu);
    if (!g_na_ptr[device]) {
        void* p = nullptr;
        SRB_CUDA(cudaGetSymbolAddress(&p, g_na_scratch));
        int sms = 0;
        SRB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        g_na_sms[device] = sms;
        g_na_ptr[device] = (uint8_t*)p;
    }
    *n_sms = g_na_sms[device];
    return g_na_ptr[device];
}
} // namespace srb

extern "C" srb_status srb_partition_run_noalloc(int32_t device, void* cuda_stream, int32_t n_partitions, const srb_chunk* in, int32_t n_key_slots,
                                                const int32_t* key_slots, const srb_chunk* out, int64_t* counts_dev) {
    X_BEGIN
    if (!in || !out || !counts_dev || !key_slots) throw Error(SRB_INVALID_ARGUMENT, "null argument");
    if (n_partitions < 1 || n_partitions > PMAX_PARTS) throw Error(SRB_INVALID_ARGUMENT, fmt("n_partitions must be 1..%d", PMAX_PARTS));
    if (n_key_slots < 1 || n_key_slots > PMAX_KEYS) throw Error(SRB_INVALID_ARGUMENT, "bad number of partition columns");
    if (in->n_cols < 1 || in->n_cols > PMAX_COLS) throw Error(SRB_NOT_IMPLEMENTED, "too many columns");
    if (in->n_rows < 0) throw Error(SRB_INVALID_ARGUMENT, "bad chunk");
    if (in->n_rows > NA_MAX_ROWS)
        throw Error(SRB_INVALID_ARGUMENT, fmt("srb_partition_run_noalloc takes at most %lld rows per call (SRB_PARTITION_NOALLOC_MAX_ROWS), got %lld: "
                                              "split the chunk or use srb_partition_run",
                                              (long long)NA_MAX_ROWS, (long long)in->n_rows));
    if (out->n_cols != in->n_cols || out->n_rows < in->n_rows) throw Error(SRB_INVALID_ARGUMENT, "output chunk: same columns and room for every input row");
    SRB_CUDA(cudaSetDevice(device));
    int n_sms = 0;
    uint8_t* scratch = na_scratch(device, &n_sms);
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    const int64_t n = in->n_rows;
    PArgs a{};
    a.n_cols = in->n_cols;
    a.n_keys = n_key_slots;
    a.n_parts = n_partitions;
    for (int c = 0; c < in->n_cols; c++) {
        const srb_column& col = in->cols[c];
        const srb_column& oc = out->cols[c];
        const int w = type_width(col.type);
        if (w == 0) throw Error(SRB_NOT_IMPLEMENTED, "partitioning of binary columns");
        if (col.mem_kind != SRB_MEM_DEVICE || oc.mem_kind != SRB_MEM_DEVICE) throw Error(SRB_INVALID_ARGUMENT, "partitioner takes device-resident chunks");
        if (oc.type != col.type || out->slot_ids[c] != in->slot_ids[c]) throw Error(SRB_INVALID_ARGUMENT, "output chunk: columns differ from the input's");
        if (n > 0 && (!col.data || !oc.data || (col.nulls && !oc.nulls))) throw Error(SRB_INVALID_ARGUMENT, "null column buffer");
        a.cols[c] = PCol{(const uint8_t*)col.data, (uint8_t*)oc.data, col.nulls, col.nulls ? (uint8_t*)oc.nulls : nullptr, w, 0};
    }
    for (int k = 0; k < n_key_slots; k++) {
        int idx = -1;
        for (int c = 0; c < in->n_cols; c++)
            if (in->slot_ids[c] == key_slots[k]) idx = c;
        if (idx < 0) throw Error(SRB_INVALID_ARGUMENT, fmt("partition slot %d not in chunk", key_slots[k]));
        a.key_col[k] = idx;
    }
    unsigned long long* totals = (unsigned long long*)scratch;
    unsigned long long* cursors = totals + PMAX_PARTS;
    unsigned long long* base = cursors + PMAX_PARTS;
    uint8_t* dest = scratch + NA_HEAD;
    // the array keeps the last call's contents: totals and cursors must start at zero (base and dest are fully rewritten)
    SRB_CUDA(cudaMemsetAsync(totals, 0, 2 * PMAX_PARTS * 8, stream));
    if (n > 0) {
        const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n + 255) / 256, (int64_t)n_sms * 8));
        k_part_count<<<grid, 256, 0, stream>>>(a, n, dest, totals);
    }
    k_part_bases_na<<<1, 32, 0, stream>>>(totals, base, counts_dev, n_partitions);
    if (n > 0) {
        const int grid2 = (int)std::max<int64_t>(1, std::min<int64_t>((n + PTILE - 1) / PTILE, (int64_t)n_sms * 4));
        k_part_scatter_na<<<grid2, 256, 0, stream>>>(a, n, dest, cursors, base);
    }
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
    X_END
}
