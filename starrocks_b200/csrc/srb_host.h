// srb_host.h — host-side plumbing of libsrb.so: context, operator base class, error convention, device buffers.
//
// The operator contract mirrored here is starrocks::pipeline::Operator (be/src/exec/pipeline/operator.h:43-344):
// push_chunk / pull_chunk / set_finishing / has_output / need_input / is_finished, one thread at a time per
// operator instance (pipeline_driver.cpp:246). Status codes are TStatusCode values (include/srb/srb_c_api.h).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "srb_dev.cuh"

#include "../../include/srb/srb_c_api.h"

namespace srb {

struct Error : public std::runtime_error {
    srb_status code;
    Error(srb_status c, const std::string& m) : std::runtime_error(m), code(c) {}
};

inline std::string fmt(const char* f, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, f);
    vsnprintf(buf, sizeof(buf), f, ap);
    va_end(ap);
    return buf;
}

#define SRB_CUDA(expr)                                                                                            \
    do {                                                                                                          \
        cudaError_t _e = (expr);                                                                                  \
        if (_e != cudaSuccess)                                                                                    \
            throw srb::Error(_e == cudaErrorMemoryAllocation ? SRB_MEM_LIMIT_EXCEEDED : SRB_INTERNAL_ERROR,       \
                             srb::fmt("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__)); \
    } while (0)

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: one flag per (kernel, device), set with that device
// current. Racing first calls both set it (idempotent); the return code is checked.
struct SmemOptIn {
    std::atomic<uint8_t> done[64] = {};
    template <class F>
    void set(F* kernel, int bytes) {
        int d = 0;
        SRB_CUDA(cudaGetDevice(&d));
        if (d >= 0 && d < 64 && done[d].load(std::memory_order_acquire)) return;
        SRB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        if (d >= 0 && d < 64) done[d].store(1, std::memory_order_release);
    }
};

inline int type_width(int t) {
    switch (t) {
    case SRB_TYPE_TINYINT:
    case SRB_TYPE_BOOLEAN: return 1;
    case SRB_TYPE_SMALLINT: return 2;
    case SRB_TYPE_INT:
    case SRB_TYPE_DATE:
    case SRB_TYPE_DECIMAL32: return 4;
    case SRB_TYPE_BIGINT:
    case SRB_TYPE_DECIMAL64:
    case SRB_TYPE_DOUBLE: return 8;
    case SRB_TYPE_LARGEINT:
    case SRB_TYPE_DECIMAL128: return 16;
    default: return 0;
    }
}
inline bool is_decimal(int t) { return t == SRB_TYPE_DECIMAL32 || t == SRB_TYPE_DECIMAL64 || t == SRB_TYPE_DECIMAL128; }
inline bool is_binary(int t) { return t == SRB_TYPE_VARCHAR || t == SRB_TYPE_VARBINARY; }

} // namespace srb

struct srb_ctx {
    srb_config cfg{};
    int n_sms = 0;
    int64_t l2_bytes = 0;
    int sm_arch = 0;
    int64_t hbm_bytes = 0;
    int64_t mem_in_use = 0;
    // freed device buffers kept for reuse: operators are created per query fragment, and cudaMalloc / cudaFree (which
    // synchronises the device) would otherwise dominate short pipelines
    std::mutex pool_mu;
    std::multimap<size_t, void*> pool;
    int64_t pool_bytes = 0;
    // how often (and for how long) the driver's allocator was reached: what a steady loop of operators should not do (srb_ctx_alloc_stats)
    std::atomic<int64_t> n_malloc{0}, n_free{0}, malloc_us{0}, free_us{0};
    cudaError_t timed_malloc(void** p, size_t n) {
        const auto t0 = std::chrono::steady_clock::now();
        const cudaError_t e = cudaMalloc(p, n);
        malloc_us += std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count();
        n_malloc++;
        return e;
    }
    void timed_free(void* p) {
        const auto t0 = std::chrono::steady_clock::now();
        cudaFree(p);
        free_us += std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count();
        n_free++;
    }
    static size_t round_size(size_t n) {
        if (n <= 256) return 256;
        if (n >= ((size_t)2 << 20)) return (n + ((size_t)2 << 20) - 1) & ~(((size_t)2 << 20) - 1);
        size_t r = 256;
        while (r < n) r <<= 1;
        return r;
    }
    void* pool_take(size_t rounded) {
        std::lock_guard<std::mutex> g(pool_mu);
        auto it = pool.find(rounded);
        if (it == pool.end()) return nullptr;
        void* p = it->second;
        pool.erase(it);
        pool_bytes -= (int64_t)rounded;
        return p;
    }
    // The buffers an operator returns at close are the ones the next, identical operator asks for: when the pool is full, older
    // entries of OTHER sizes make room (freed here) before the newcomer is turned away. Refusing the newcomer instead made every
    // step of a steady loop pay cudaFree + cudaMalloc of gigabytes once the pool had filled with stale sizes.
    bool pool_give(size_t rounded, void* p) {
        std::vector<void*> evict;
        {
            std::lock_guard<std::mutex> g(pool_mu);
            const int64_t limit = std::max<int64_t>((int64_t)48 << 30, hbm_bytes / 2);
            if ((int64_t)rounded > limit) return false;
            for (auto it = pool.begin(); it != pool.end() && pool_bytes + (int64_t)rounded > limit;) {
                if (it->first == rounded) {
                    ++it;
                    continue;
                }
                evict.push_back(it->second);
                pool_bytes -= (int64_t)it->first;
                it = pool.erase(it);
            }
            if (pool_bytes + (int64_t)rounded > limit) {
                for (void* q : evict) timed_free(q);
                return false;
            }
            pool.emplace(rounded, p);
            pool_bytes += (int64_t)rounded;
        }
        for (void* q : evict) timed_free(q);
        return true;
    }
    void pool_clear() {
        std::lock_guard<std::mutex> g(pool_mu);
        for (auto& kv : pool) timed_free(kv.second);
        pool.clear();
        pool_bytes = 0;
    }
    // idle streams and events of closed operators, reused by the next operator (a returned stream has been synchronised)
    std::vector<cudaStream_t> idle_streams;
    std::vector<cudaEvent_t> idle_events, idle_events_notiming;
    cudaStream_t take_stream() {
        {
            std::lock_guard<std::mutex> g(pool_mu);
            if (!idle_streams.empty()) {
                cudaStream_t s = idle_streams.back();
                idle_streams.pop_back();
                return s;
            }
        }
        cudaStream_t s = nullptr;
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        return s;
    }
    void give_stream(cudaStream_t s) {
        if (!s) return;
        std::lock_guard<std::mutex> g(pool_mu);
        if (idle_streams.size() < 64)
            idle_streams.push_back(s);
        else
            cudaStreamDestroy(s);
    }
    cudaEvent_t take_event(bool timing) {
        {
            std::lock_guard<std::mutex> g(pool_mu);
            auto& v = timing ? idle_events : idle_events_notiming;
            if (!v.empty()) {
                cudaEvent_t e = v.back();
                v.pop_back();
                return e;
            }
        }
        cudaEvent_t e = nullptr;
        if (cudaEventCreateWithFlags(&e, timing ? cudaEventDefault : cudaEventDisableTiming) != cudaSuccess) return nullptr;
        return e;
    }
    void give_event(cudaEvent_t e, bool timing) {
        if (!e) return;
        std::lock_guard<std::mutex> g(pool_mu);
        auto& v = timing ? idle_events : idle_events_notiming;
        if (v.size() < 256)
            v.push_back(e);
        else
            cudaEventDestroy(e);
    }
    ~srb_ctx() {
        pool_clear();
        for (auto s : idle_streams) cudaStreamDestroy(s);
        for (auto e : idle_events) cudaEventDestroy(e);
        for (auto e : idle_events_notiming) cudaEventDestroy(e);
    }
};

namespace srb {

// RAII device / pinned buffers that charge the ctx's mem_limit
struct DevBuf {
    srb_ctx* ctx = nullptr;
    void* p = nullptr;
    size_t bytes = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : ctx(o.ctx), p(o.p), bytes(o.bytes) {
        o.p = nullptr;
        o.bytes = 0;
    }
    ~DevBuf() { release(); }
    void release() {
        if (p) {
            if (!ctx)
                cudaFree(p);
            else if (!ctx->pool_give(bytes, p))
                ctx->timed_free(p);
            if (ctx) ctx->mem_in_use -= (int64_t)bytes;
        }
        p = nullptr;
        bytes = 0;
    }
    // (re)allocate when the current capacity is too small; contents are NOT preserved
    void ensure(srb_ctx* c, size_t n) {
        if (n <= bytes && p) return;
        release();
        ctx = c;
        n = srb_ctx::round_size(n);
        if (c->cfg.mem_limit > 0 && c->mem_in_use + (int64_t)n > c->cfg.mem_limit)
            throw Error(SRB_MEM_LIMIT_EXCEEDED, fmt("device memory limit exceeded: %lld + %zu > %lld", (long long)c->mem_in_use, n,
                                                    (long long)c->cfg.mem_limit));
        p = c->pool_take(n);
        if (!p) {
            cudaError_t e = c->timed_malloc(&p, n);
            if (e == cudaErrorMemoryAllocation) { // give the cached buffers back to the driver and retry once
                cudaGetLastError();
                c->pool_clear();
                e = c->timed_malloc(&p, n);
            }
            if (e != cudaSuccess) {
                p = nullptr;
                SRB_CUDA(e);
            }
        }
        bytes = n;
        c->mem_in_use += (int64_t)n;
    }
    template <typename T>
    T* as() const {
        return (T*)p;
    }
};

// Small pinned buffers (counters, totals) are recycled process-wide: cudaMallocHost / cudaFreeHost cost about a
// millisecond each, which is as long as a whole kernel of a short pipeline. Large ones go straight to the driver.
struct PinnedPool {
    static constexpr size_t MAX_POOLED = (size_t)1 << 20, MAX_TOTAL = (size_t)64 << 20;
    std::mutex mu;
    std::multimap<size_t, void*> free_;
    size_t total = 0;
    static PinnedPool& get() {
        static PinnedPool* g = new PinnedPool(); // never destroyed: buffers may outlive static destruction order
        return *g;
    }
    static size_t round_size(size_t n) {
        size_t r = 256;
        while (r < n) r <<= 1;
        return r;
    }
    void* take(size_t rounded) {
        std::lock_guard<std::mutex> g(mu);
        auto it = free_.find(rounded);
        if (it == free_.end()) return nullptr;
        void* p = it->second;
        free_.erase(it);
        total -= rounded;
        return p;
    }
    bool give(size_t rounded, void* p) {
        std::lock_guard<std::mutex> g(mu);
        if (rounded > MAX_POOLED || total + rounded > MAX_TOTAL) return false;
        free_.emplace(rounded, p);
        total += rounded;
        return true;
    }
};

struct PinnedBuf {
    void* p = nullptr;
    size_t bytes = 0;
    PinnedBuf() = default;
    PinnedBuf(const PinnedBuf&) = delete;
    PinnedBuf& operator=(const PinnedBuf&) = delete;
    ~PinnedBuf() { release(); }
    void release() {
        if (p && !PinnedPool::get().give(bytes, p)) cudaFreeHost(p);
        p = nullptr;
        bytes = 0;
    }
    void ensure(size_t n) {
        if (n <= bytes && p) return;
        release();
        if (n == 0) n = 16;
        if (n <= PinnedPool::MAX_POOLED) {
            n = PinnedPool::round_size(n);
            p = PinnedPool::get().take(n);
        }
        if (!p) SRB_CUDA(cudaMallocHost(&p, n));
        bytes = n;
    }
    template <typename T>
    T* as() const {
        return (T*)p;
    }
};

} // namespace srb

// Operator base: the C ABI forwards to these virtuals.
struct srb_op {
    srb_ctx* ctx;
    cudaStream_t stream = nullptr;
    bool finishing = false, finished = false, cancelled = false;
    int64_t rows_in = 0, rows_selected = 0, launches = 0;
    explicit srb_op(srb_ctx* c) : ctx(c) {
        stream = c->take_stream();
        if (!stream) throw srb::Error(SRB_INTERNAL_ERROR, "cudaStreamCreate failed");
    }
    virtual ~srb_op() {
        if (stream) {
            cudaStreamSynchronize(stream);
            ctx->give_stream(stream);
        }
    }
    virtual void push_chunk(const srb_chunk* ch) = 0;
    virtual void set_finishing() = 0;
    virtual bool has_output() = 0;
    virtual bool need_input() = 0;
    virtual bool is_finished() = 0;
    virtual srb_status pull_chunk(srb_chunk* out, bool all) = 0;
    virtual srb_status pull_device(srb_chunk*) { throw srb::Error(SRB_NOT_IMPLEMENTED, "device pull not supported by this operator"); }
    virtual void refresh_counters() {}
    virtual void profile(double* kernel_ms, int64_t* batches, int64_t* static_launches, int32_t* shape) {
        *kernel_ms = 0;
        *batches = 0;
        *static_launches = 0;
        *shape = -1;
    }
    // join ops override the probe side
    virtual void probe_push_chunk(const srb_chunk*) { throw srb::Error(SRB_INVALID_ARGUMENT, "not a join operator"); }
    virtual void probe_set_finishing() { throw srb::Error(SRB_INVALID_ARGUMENT, "not a join operator"); }
    virtual bool probe_has_output() { return false; }
    virtual bool probe_need_input() { return false; }
    virtual bool probe_is_finished() { return true; }
    virtual srb_status probe_pull_chunk(srb_chunk*, bool) { throw srb::Error(SRB_INVALID_ARGUMENT, "not a join operator"); }
    virtual srb_status probe_pull_device(srb_chunk*) { throw srb::Error(SRB_INVALID_ARGUMENT, "not a join operator"); }
    // runtime filters: join ops build them from their build side, scan-filter ops apply them
    virtual struct srb_runtime_filter* build_runtime_filter(int, int64_t, int) { throw srb::Error(SRB_INVALID_ARGUMENT, "not a join operator"); }
    virtual void add_runtime_filter(const struct srb_runtime_filter*, int) { throw srb::Error(SRB_NOT_IMPLEMENTED, "this operator does not take runtime filters"); }
    // set_finishing + gather exchange + merge phase in one (aggregate operators with small tables)
    virtual void set_finishing_merged(struct srb_comm*) { throw srb::Error(SRB_NOT_IMPLEMENTED, "this operator has no cross-rank merge"); }
    // encoded segment pages (scan operators): decoded on the device in front of the scan
    virtual void push_pages(int32_t, const int32_t*, const srb_page_column*, int64_t) { throw srb::Error(SRB_NOT_IMPLEMENTED, "this operator does not take encoded pages"); }
    // streaming pre-aggregation (aggregate operators in the update+serialize phase only)
    virtual void set_streaming_mode(int) { throw srb::Error(SRB_INVALID_ARGUMENT, "not an aggregate operator"); }
    virtual void set_streaming_memory_limit(int64_t) { throw srb::Error(SRB_INVALID_ARGUMENT, "not an aggregate operator"); }
    virtual void streaming_stats(int64_t*, int64_t*, int32_t*) { throw srb::Error(SRB_INVALID_ARGUMENT, "not an aggregate operator"); }
};

// a runtime filter built from a join's build side (HashJoinBuildOperator -> RuntimeBloomFilter): device-resident SimdBlockFilter
// directory + min / max / has_null kept on the host
struct srb_runtime_filter {
    srb_ctx* ctx = nullptr;
    srb::DevBuf dir;
    int log_num_buckets = 1;
    uint32_t directory_mask = 1;
    int key_type = 0, key_width = 0;
    int64_t min = INT64_MAX, max = INT64_MIN, inserted = 0;
    bool has_null = false;
    int mix = 0; // RfView::mix
    // RuntimeBitsetFilter instead of the bloom filter (chosen by create_join_runtime_filter's rule, runtime_filter_bank.cpp:160-205): one bit
    // per value of [min, max], padded to whole 32-bit words
    srb::DevBuf bits;
    bool is_bitset = false;
    size_t bits_bytes() const { return is_bitset ? (size_t)((((uint64_t)max - (uint64_t)min + 1) + 31) / 32) * 4 : 0; }
    srb::RfView view() const {
        return srb::RfView{(uint32_t*)dir.p, directory_mask, log_num_buckets, min, max, has_null ? 1 : 0, mix, is_bitset ? (const uint32_t*)bits.p : nullptr};
    }
    size_t dir_bytes() const { return (size_t)32 << log_num_buckets; }
};

namespace srb {
srb_op* make_scan_agg_op(srb_ctx* ctx, const srb_scan_agg_plan* plan);
srb_op* make_join_op(srb_ctx* ctx, const srb_join_plan* plan);
srb_op* make_topn_op(srb_ctx* ctx, const srb_topn_plan* plan);
// run-time compiled shapes (srb_jit.h; defined in srb_scan_agg.cu, the one translation unit that carries the embedded kernel headers)
void jit_wait_all();
void jit_stats(int64_t out[4], std::string* last_error);
void jit_check_plan(const srb_scan_agg_plan* plan, const srb_chunk* first, std::string* description);
} // namespace srb

struct srb_comm;
namespace srb {
void comm_all_gather_device(srb_comm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t user_stream);
int comm_world(const srb_comm* c);
int comm_rank(const srb_comm* c);
} // namespace srb
