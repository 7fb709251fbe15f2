// srb_capi.cu — the extern "C" surface of libsrb.so (include/srb/srb_c_api.h). No C++ type or exception crosses it.
#include <cstring>
#include <string>

#include "srb_host.h"

static thread_local std::string g_last_error;
namespace srb {
void set_last_error(const std::string& s) { g_last_error = s; }
} // namespace srb

#define SRB_API_BEGIN try {
#define SRB_API_END                                   \
    }                                                 \
    catch (const srb::Error& e) {                     \
        g_last_error = e.what();                      \
        return e.code;                                \
    }                                                 \
    catch (const std::bad_alloc&) {                   \
        g_last_error = "host allocation failed";      \
        return SRB_MEM_LIMIT_EXCEEDED;                \
    }                                                 \
    catch (const std::exception& e) {                 \
        g_last_error = e.what();                      \
        return SRB_INTERNAL_ERROR;                    \
    }

extern "C" {

const char* srb_last_error(void) { return g_last_error.c_str(); }

srb_status srb_ctx_create(const srb_config* cfg, srb_ctx** out) {
    SRB_API_BEGIN
    if (!cfg || !out) throw srb::Error(SRB_INVALID_ARGUMENT, "null argument");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        throw srb::Error(SRB_INTERNAL_ERROR, std::string("no CUDA device: libsrb has no CPU path (") + cudaGetErrorString(e) + ")");
    if (cfg->device < 0 || cfg->device >= n) throw srb::Error(SRB_INVALID_ARGUMENT, "bad device ordinal");
    SRB_CUDA(cudaSetDevice(cfg->device));
    cudaDeviceProp p;
    SRB_CUDA(cudaGetDeviceProperties(&p, cfg->device));
    if (p.major < 10) throw srb::Error(SRB_INTERNAL_ERROR, srb::fmt("device is sm_%d%d; libsrb is built for sm_100a only", p.major, p.minor));
    // the group table is probed with random 16..32-byte accesses: ask L2 not to over-fetch neighbouring sectors
    cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, 32);
    auto* c = new srb_ctx();
    c->cfg = *cfg;
    if (c->cfg.world_size <= 0) c->cfg.world_size = 1;
    c->n_sms = p.multiProcessorCount;
    c->l2_bytes = p.l2CacheSize;
    c->sm_arch = p.major * 10 + p.minor;
    c->hbm_bytes = (int64_t)p.totalGlobalMem;
    *out = c;
    return SRB_OK;
    SRB_API_END
}

void srb_ctx_destroy(srb_ctx* ctx) { delete ctx; }

srb_status srb_ctx_device_info(srb_ctx* ctx, int32_t* n_sms, int64_t* hbm_bytes, int32_t* sm_arch) {
    SRB_API_BEGIN
    if (!ctx) throw srb::Error(SRB_INVALID_ARGUMENT, "null ctx");
    if (n_sms) *n_sms = ctx->n_sms;
    if (hbm_bytes) *hbm_bytes = ctx->hbm_bytes;
    if (sm_arch) *sm_arch = ctx->sm_arch;
    return SRB_OK;
    SRB_API_END
}

srb_status srb_ctx_trim(srb_ctx* ctx) {
    SRB_API_BEGIN
    if (!ctx) throw srb::Error(SRB_INVALID_ARGUMENT, "null ctx");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    ctx->pool_clear();
    return SRB_OK;
    SRB_API_END
}

void* srb_op_stream(srb_op* op) { return op ? (void*)op->stream : nullptr; }

srb_status srb_ctx_alloc_stats(srb_ctx* ctx, int64_t* n_malloc, int64_t* n_free, int64_t* malloc_us, int64_t* free_us, int64_t* pooled_bytes) {
    SRB_API_BEGIN
    if (!ctx) throw srb::Error(SRB_INVALID_ARGUMENT, "null ctx");
    if (n_malloc) *n_malloc = ctx->n_malloc.load();
    if (n_free) *n_free = ctx->n_free.load();
    if (malloc_us) *malloc_us = ctx->malloc_us.load();
    if (free_us) *free_us = ctx->free_us.load();
    if (pooled_bytes) {
        std::lock_guard<std::mutex> g(ctx->pool_mu);
        *pooled_bytes = ctx->pool_bytes;
    }
    return SRB_OK;
    SRB_API_END
}

srb_status srb_jit_wait(void) {
    SRB_API_BEGIN
    srb::jit_wait_all();
    return SRB_OK;
    SRB_API_END
}

srb_status srb_jit_stats(int64_t* compiled, int64_t* from_disk_cache, int64_t* failed, int64_t* pending, char* last_error, int64_t last_error_len) {
    SRB_API_BEGIN
    int64_t st[4];
    std::string err;
    srb::jit_stats(st, &err);
    if (compiled) *compiled = st[0];
    if (from_disk_cache) *from_disk_cache = st[1];
    if (failed) *failed = st[2];
    if (pending) *pending = st[3];
    if (last_error && last_error_len > 0) snprintf(last_error, (size_t)last_error_len, "%s", err.c_str());
    return SRB_OK;
    SRB_API_END
}

srb_status srb_jit_check_plan(const srb_scan_agg_plan* plan, const srb_chunk* first_chunk, char* description, int64_t description_len) {
    SRB_API_BEGIN
    std::string d;
    srb::jit_check_plan(plan, first_chunk, &d);
    if (description && description_len > 0) snprintf(description, (size_t)description_len, "%s", d.c_str());
    return SRB_OK;
    SRB_API_END
}

srb_status srb_scan_agg_create(srb_ctx* ctx, const srb_scan_agg_plan* plan, srb_op** out) {
    SRB_API_BEGIN
    if (!ctx || !out) throw srb::Error(SRB_INVALID_ARGUMENT, "null argument");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    *out = srb::make_scan_agg_op(ctx, plan);
    return SRB_OK;
    SRB_API_END
}

srb_status srb_join_create(srb_ctx* ctx, const srb_join_plan* plan, srb_op** out) {
    SRB_API_BEGIN
    if (!ctx || !out) throw srb::Error(SRB_INVALID_ARGUMENT, "null argument");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    *out = srb::make_join_op(ctx, plan);
    return SRB_OK;
    SRB_API_END
}

srb_status srb_topn_create(srb_ctx* ctx, const srb_topn_plan* plan, srb_op** out) {
    SRB_API_BEGIN
    if (!ctx || !out) throw srb::Error(SRB_INVALID_ARGUMENT, "null argument");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    *out = srb::make_topn_op(ctx, plan);
    return SRB_OK;
    SRB_API_END
}

#define OP_GUARD(op)                                                   \
    if (!(op)) throw srb::Error(SRB_INVALID_ARGUMENT, "null operator"); \
    SRB_CUDA(cudaSetDevice((op)->ctx->cfg.device));

srb_status srb_op_push_chunk(srb_op* op, const srb_chunk* chunk) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->push_chunk(chunk);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_op_push_pages(srb_op* op, int32_t n_cols, const int32_t* slot_ids, const srb_page_column* cols, int64_t n_rows) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->push_pages(n_cols, slot_ids, cols, n_rows);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_op_set_finishing_merged(srb_op* op, srb_comm* comm) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->set_finishing_merged(comm);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_op_set_finishing(srb_op* op) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->set_finishing();
    return SRB_OK;
    SRB_API_END
}
int32_t srb_op_has_output(srb_op* op) { return op && !op->cancelled && op->has_output() ? 1 : 0; }
int32_t srb_op_need_input(srb_op* op) { return op && !op->cancelled && op->need_input() ? 1 : 0; }
int32_t srb_op_is_finished(srb_op* op) { return !op || op->cancelled || op->is_finished() ? 1 : 0; }
srb_status srb_op_pull_chunk(srb_op* op, srb_chunk* out) {
    SRB_API_BEGIN
    OP_GUARD(op)
    if (!out) throw srb::Error(SRB_INVALID_ARGUMENT, "null out chunk");
    return op->pull_chunk(out, false);
    SRB_API_END
}
srb_status srb_op_pull_all(srb_op* op, srb_chunk* out) {
    SRB_API_BEGIN
    OP_GUARD(op)
    if (!out) throw srb::Error(SRB_INVALID_ARGUMENT, "null out chunk");
    return op->pull_chunk(out, true);
    SRB_API_END
}
srb_status srb_op_pull_device(srb_op* op, srb_chunk* out) {
    SRB_API_BEGIN
    OP_GUARD(op)
    if (!out) throw srb::Error(SRB_INVALID_ARGUMENT, "null out chunk");
    return op->pull_device(out);
    SRB_API_END
}
srb_status srb_join_probe_pull_device(srb_op* op, srb_chunk* out) {
    SRB_API_BEGIN
    OP_GUARD(op)
    if (!out) throw srb::Error(SRB_INVALID_ARGUMENT, "null out chunk");
    return op->probe_pull_device(out);
    SRB_API_END
}
srb_status srb_op_cancel(srb_op* op) {
    SRB_API_BEGIN
    if (!op) throw srb::Error(SRB_INVALID_ARGUMENT, "null operator");
    op->cancelled = true;
    op->finished = true;
    return SRB_OK;
    SRB_API_END
}
void srb_op_close(srb_op* op) {
    if (!op) return;
    cudaSetDevice(op->ctx->cfg.device);
    delete op;
}
srb_status srb_op_sync(srb_op* op) {
    SRB_API_BEGIN
    OP_GUARD(op)
    SRB_CUDA(cudaStreamSynchronize(op->stream));
    return SRB_OK;
    SRB_API_END
}
srb_status srb_op_counters(srb_op* op, int64_t* rows_in, int64_t* rows_selected, int64_t* kernel_launches) {
    SRB_API_BEGIN
    if (!op) throw srb::Error(SRB_INVALID_ARGUMENT, "null operator");
    op->refresh_counters();
    if (rows_in) *rows_in = op->rows_in;
    if (rows_selected) *rows_selected = op->rows_selected;
    if (kernel_launches) *kernel_launches = op->launches;
    return SRB_OK;
    SRB_API_END
}

srb_status srb_op_profile(srb_op* op, double* kernel_ms, int64_t* batches, int64_t* static_launches, int32_t* shape) {
    SRB_API_BEGIN
    if (!op || !kernel_ms || !batches || !static_launches || !shape) throw srb::Error(SRB_INVALID_ARGUMENT, "null argument");
    op->profile(kernel_ms, batches, static_launches, shape);
    return SRB_OK;
    SRB_API_END
}

srb_status srb_join_probe_push_chunk(srb_op* op, const srb_chunk* chunk) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->probe_push_chunk(chunk);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_join_probe_set_finishing(srb_op* op) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->probe_set_finishing();
    return SRB_OK;
    SRB_API_END
}
int32_t srb_join_probe_has_output(srb_op* op) { return op && !op->cancelled && op->probe_has_output() ? 1 : 0; }
int32_t srb_join_probe_need_input(srb_op* op) { return op && !op->cancelled && op->probe_need_input() ? 1 : 0; }
int32_t srb_join_probe_is_finished(srb_op* op) { return !op || op->cancelled || op->probe_is_finished() ? 1 : 0; }
srb_status srb_join_probe_pull_chunk(srb_op* op, srb_chunk* out) {
    SRB_API_BEGIN
    OP_GUARD(op)
    if (!out) throw srb::Error(SRB_INVALID_ARGUMENT, "null out chunk");
    return op->probe_pull_chunk(out, false);
    SRB_API_END
}
srb_status srb_join_probe_pull_all(srb_op* op, srb_chunk* out) {
    SRB_API_BEGIN
    OP_GUARD(op)
    if (!out) throw srb::Error(SRB_INVALID_ARGUMENT, "null out chunk");
    return op->probe_pull_chunk(out, true);
    SRB_API_END
}

// ---- runtime filters ---------------------------------------------------------------------------------------------------------
srb_status srb_join_build_runtime_filter(srb_op* op, int32_t key_index, int64_t expected_rows, int32_t hash_mix, srb_runtime_filter** out) {
    SRB_API_BEGIN
    OP_GUARD(op)
    if (!out) throw srb::Error(SRB_INVALID_ARGUMENT, "null out");
    *out = op->build_runtime_filter(key_index, expected_rows, hash_mix);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_op_add_runtime_filter(srb_op* op, const srb_runtime_filter* rf, int32_t slot_id) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->add_runtime_filter(rf, slot_id);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_op_set_streaming_mode(srb_op* op, int32_t mode) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->set_streaming_mode(mode);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_op_set_streaming_memory_limit(srb_op* op, int64_t bytes) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->set_streaming_memory_limit(bytes);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_op_streaming_stats(srb_op* op, int64_t* rows_passed_through, int64_t* groups_in_table, int32_t* auto_state) {
    SRB_API_BEGIN
    OP_GUARD(op)
    op->streaming_stats(rows_passed_through, groups_in_table, auto_state);
    return SRB_OK;
    SRB_API_END
}
srb_status srb_runtime_filter_info(const srb_runtime_filter* rf, int64_t* n_buckets, int64_t* min_value, int64_t* max_value, int32_t* has_null,
                                   int64_t* inserted) {
    SRB_API_BEGIN
    if (!rf) throw srb::Error(SRB_INVALID_ARGUMENT, "null runtime filter");
    if (n_buckets) *n_buckets = (int64_t)1 << rf->log_num_buckets;
    if (min_value) *min_value = rf->min;
    if (max_value) *max_value = rf->max;
    if (has_null) *has_null = rf->has_null ? 1 : 0;
    if (inserted) *inserted = rf->inserted;
    return SRB_OK;
    SRB_API_END
}
srb_status srb_runtime_filter_directory(const srb_runtime_filter* rf, void* dst, int64_t bytes) {
    SRB_API_BEGIN
    if (!rf || !dst) throw srb::Error(SRB_INVALID_ARGUMENT, "null argument");
    if (bytes != (int64_t)rf->dir_bytes()) throw srb::Error(SRB_INVALID_ARGUMENT, "directory size mismatch (32 bytes per bucket)");
    SRB_CUDA(cudaSetDevice(rf->ctx->cfg.device));
    SRB_CUDA(cudaMemcpy(dst, rf->dir.p, (size_t)bytes, cudaMemcpyDeviceToHost));
    return SRB_OK;
    SRB_API_END
}
srb_status srb_runtime_filter_kind(const srb_runtime_filter* rf, int32_t* kind, int64_t* bitset_bytes) {
    SRB_API_BEGIN
    if (!rf || !kind) throw srb::Error(SRB_INVALID_ARGUMENT, "null argument");
    *kind = rf->is_bitset ? SRB_RF_KIND_BITSET : SRB_RF_KIND_BLOOM;
    if (bitset_bytes) *bitset_bytes = (int64_t)rf->bits_bytes();
    return SRB_OK;
    SRB_API_END
}
srb_status srb_runtime_filter_bitset(const srb_runtime_filter* rf, void* dst, int64_t bytes) {
    SRB_API_BEGIN
    if (!rf || !dst) throw srb::Error(SRB_INVALID_ARGUMENT, "null argument");
    if (!rf->is_bitset || bytes != (int64_t)rf->bits_bytes()) throw srb::Error(SRB_INVALID_ARGUMENT, "not a bitset filter / size mismatch");
    SRB_CUDA(cudaSetDevice(rf->ctx->cfg.device));
    SRB_CUDA(cudaMemcpy(dst, rf->bits.p, (size_t)bytes, cudaMemcpyDeviceToHost));
    return SRB_OK;
    SRB_API_END
}
void srb_runtime_filter_destroy(srb_runtime_filter* rf) {
    if (!rf) return;
    cudaSetDevice(rf->ctx->cfg.device);
    delete rf;
}

srb_status srb_device_alloc(srb_ctx* ctx, int64_t bytes, void** out) {
    SRB_API_BEGIN
    if (!ctx || !out || bytes < 0) throw srb::Error(SRB_INVALID_ARGUMENT, "bad argument");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    SRB_CUDA(cudaMalloc(out, (size_t)(bytes ? bytes : 16)));
    return SRB_OK;
    SRB_API_END
}
void srb_device_free(srb_ctx* ctx, void* p) {
    if (ctx) cudaSetDevice(ctx->cfg.device);
    if (p) cudaFree(p);
}
srb_status srb_pinned_alloc(srb_ctx* ctx, int64_t bytes, void** out) {
    SRB_API_BEGIN
    if (!ctx || !out || bytes < 0) throw srb::Error(SRB_INVALID_ARGUMENT, "bad argument");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    SRB_CUDA(cudaMallocHost(out, (size_t)(bytes ? bytes : 16)));
    return SRB_OK;
    SRB_API_END
}
void srb_pinned_free(srb_ctx* ctx, void* p) {
    (void)ctx;
    if (p) cudaFreeHost(p);
}
srb_status srb_memcpy_h2d(srb_ctx* ctx, void* dst, const void* src, int64_t bytes) {
    SRB_API_BEGIN
    if (!ctx) throw srb::Error(SRB_INVALID_ARGUMENT, "null ctx");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    SRB_CUDA(cudaMemcpy(dst, src, (size_t)bytes, cudaMemcpyHostToDevice));
    return SRB_OK;
    SRB_API_END
}
srb_status srb_memcpy_d2h(srb_ctx* ctx, void* dst, const void* src, int64_t bytes) {
    SRB_API_BEGIN
    if (!ctx) throw srb::Error(SRB_INVALID_ARGUMENT, "null ctx");
    SRB_CUDA(cudaSetDevice(ctx->cfg.device));
    SRB_CUDA(cudaMemcpy(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost));
    return SRB_OK;
    SRB_API_END
}

} // extern "C"
