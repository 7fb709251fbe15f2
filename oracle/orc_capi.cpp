// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_arith.h header). extern "C" surface for ctypes.
#include <string>

#include "orc_agg.h"
#include "orc_rf.h"

namespace orc {
Result* run_scan_agg(const srb_scan_agg_plan* plan, const srb_chunk* in, int n_threads, int chunk_size);
Result* run_scan_agg_streaming(const srb_scan_agg_plan* plan, const srb_chunk* in, int chunk_size, int mode, int64_t* stats);
Result* run_sort(const srb_chunk* in, int n_keys, const int32_t* key_slots, const uint8_t* is_asc, const uint8_t* nulls_first, int64_t offset, int64_t limit);
Result* run_join(const srb_join_plan* plan, const srb_chunk* build, const srb_chunk* probe, int n_threads);
void join_build_raw(const void* keys, int width, uint32_t n, uint32_t* bucket_size, uint32_t* first, uint32_t* next);
int pack_keys_raw(const srb_chunk* ch, const int32_t* slots, int n_keys, void* out16);
} // namespace orc

static thread_local std::string g_err;

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

// returns NULL on error
void* orc_scan_agg(const srb_scan_agg_plan* plan, const srb_chunk* in, int n_threads, int chunk_size) {
    try {
        return orc::run_scan_agg(plan, in, n_threads, chunk_size);
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}
// streaming pre-aggregation sink; mode = TStreamingPreaggregationMode (0 AUTO, 1 FORCE_STREAMING, 2 FORCE_PREAGGREGATION);
// stats[9]: rows passed through, groups in the table, chunks seen in each of the six AggrAutoState states, final state
void* orc_scan_agg_streaming(const srb_scan_agg_plan* plan, const srb_chunk* in, int chunk_size, int mode, int64_t* stats) {
    try {
        return orc::run_scan_agg_streaming(plan, in, chunk_size, mode, stats);
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}
// The streaming sink's table lookups on a sequence of key chunks (be/test/exec/agg_hash_map_test.cpp:232-275): ops[i] = 0 aggregate
// the chunk (build_hash_map), 1 build_hash_map_with_selection, 2 build_hash_map_with_selection_and_allocation; not_found receives
// the concatenated per-row flags of the chunks run with ops 1 and 2
int orc_agg_lookup_sequence(const srb_scan_agg_plan* plan, int n_chunks, const srb_chunk* const* chunks, const int32_t* ops, uint8_t* not_found) {
    try {
        orc::Aggregator ag(plan);
        size_t out = 0;
        for (int c = 0; c < n_chunks; c++) {
            if (ops[c] == 0) {
                ag.push_rows(*chunks[c], 0, (size_t)chunks[c]->n_rows);
                continue;
            }
            std::vector<uint8_t> nf;
            ag.lookup_selection(*chunks[c], ops[c] == 2, nf);
            for (uint8_t x : nf) not_found[out++] = x;
        }
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}
// ORDER BY key_slots (is_asc / nulls_first per key) OFFSET offset LIMIT limit (limit < 0: every row)
void* orc_sort(const srb_chunk* in, int n_keys, const int32_t* key_slots, const uint8_t* is_asc, const uint8_t* nulls_first, int64_t offset,
               int64_t limit) {
    try {
        return orc::run_sort(in, n_keys, key_slots, is_asc, nulls_first, offset, limit);
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}
void* orc_join_mt(const srb_join_plan* plan, const srb_chunk* build, const srb_chunk* probe, int n_threads);
void* orc_join(const srb_join_plan* plan, const srb_chunk* build, const srb_chunk* probe) { return orc_join_mt(plan, build, probe, 1); }
void* orc_join_mt(const srb_join_plan* plan, const srb_chunk* build, const srb_chunk* probe, int n_threads) {
    try {
        return orc::run_join(plan, build, probe, n_threads);
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}
const srb_chunk* orc_result_chunk(void* r) { return &((orc::Result*)r)->chunk; }
void orc_result_counters(void* r, int64_t* rows_in, int64_t* rows_selected) {
    *rows_in = ((orc::Result*)r)->rows_in;
    *rows_selected = ((orc::Result*)r)->rows_selected;
}
void orc_result_free(void* r) { delete (orc::Result*)r; }

// ---- scalar entry points for the golden-vector tests ------------------------------------------------
int32_t orc_from_date(int y, int m, int d) { return orc::from_date(y, m, d); }
void orc_to_date(int32_t j, int* y, int* m, int* d) { orc::to_date(j, y, m, d); }

// width: 4, 8, 16; op: SRB_EXPR_ADD/SUB/MUL; values passed as (lo, hi) of a sign-extended int128
int orc_checked_binop(int width, int op, int64_t alo, int64_t ahi, int64_t blo, int64_t bhi, int64_t* rlo, int64_t* rhi) {
    orc::int128_t a = ((orc::int128_t)ahi << 64) | (uint64_t)alo, b = ((orc::int128_t)bhi << 64) | (uint64_t)blo, r;
    bool ov = width == 16 ? orc::checked_binop<orc::int128_t>(op, a, b, &r)
                          : (width == 8 ? orc::checked_binop<int64_t>(op, a, b, &r) : orc::checked_binop<int32_t>(op, a, b, &r));
    *rlo = (int64_t)r;
    *rhi = (int64_t)(r >> 64);
    return ov;
}
void orc_decimal_div_integer(int64_t slo, int64_t shi, int64_t count, int scale, int64_t* rlo, int64_t* rhi) {
    orc::int128_t s = ((orc::int128_t)shi << 64) | (uint64_t)slo;
    orc::int128_t r = orc::decimal_div_integer(s, (orc::int128_t)count, scale);
    *rlo = (int64_t)r;
    *rhi = (int64_t)(r >> 64);
}
void orc_div_round128(int64_t alo, int64_t ahi, int64_t blo, int64_t bhi, int64_t* rlo, int64_t* rhi) {
    orc::int128_t a = ((orc::int128_t)ahi << 64) | (uint64_t)alo, b = ((orc::int128_t)bhi << 64) | (uint64_t)blo;
    orc::int128_t r = orc::div_round<orc::int128_t>(a, b);
    *rlo = (int64_t)r;
    *rhi = (int64_t)(r >> 64);
}
int64_t orc_div_round64(int64_t a, int64_t b) { return orc::div_round<int64_t>(a, b); }

uint32_t orc_join_key_hash32(uint32_t v, uint32_t log_buckets) { return orc::join_key_hash32(v, log_buckets); }
uint32_t orc_join_key_hash64(uint64_t v, uint32_t log_buckets) { return orc::join_key_hash64(v, log_buckets); }
uint32_t orc_join_key_hash_bytes(const void* p, int32_t n, uint32_t num_buckets) { return orc::join_key_hash_bytes(p, n, num_buckets); }
uint32_t orc_calc_bucket_size(uint32_t size) { return orc::calc_bucket_size(size); }
void orc_join_build_raw(const void* keys, int width, uint32_t n, uint32_t* bucket_size, uint32_t* first, uint32_t* next) {
    orc::join_build_raw(keys, width, n, bucket_size, first, next);
}
int orc_pack_keys(const srb_chunk* ch, const int32_t* slots, int n_keys, void* out16) {
    try {
        return orc::pack_keys_raw(ch, slots, n_keys, out16);
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1;
    }
}
uint32_t orc_crc_hash_32(const void* p, int32_t n, uint32_t seed) { return orc::crc_hash_32(p, n, seed); }
uint64_t orc_crc_hash_64(const void* p, int32_t n, uint64_t seed) { return orc::crc_hash_64(p, n, seed); }
uint64_t orc_phmap_mix8(uint64_t v, int seed) { return orc::phmap_mix8(v, seed); }
uint64_t orc_hash128_with_seed(int64_t lo, int64_t hi, int seed) {
    return orc::hash128_with_seed(((orc::int128_t)hi << 64) | (uint64_t)lo, seed);
}
uint32_t orc_fnv_hash(const void* p, int32_t n, uint32_t seed) { return orc::fnv_hash(p, n, seed); }
uint32_t orc_xorshift32(uint32_t x) { return orc::xorshift32(x); }
// ExchangeSinkOperator HASH_PARTITIONED channel of one fixed-width key (exchange_sink_operator.cpp:581-610,
// shuffler.h:70-85): FNV-1 over the value bytes from FNV_SEED, then ReduceOp
uint32_t orc_shuffle_channel(const void* value, int32_t width, uint32_t n_channels) {
    return orc::reduce_op(orc::fnv_hash(value, width, orc::FNV_SEED), n_channels);
}
// ---- runtime filter (orc_rf.h) ------------------------------------------------------------------------------------------------
void* orc_rf_create(uint64_t nums, int mix) {
    auto* f = new orc::RuntimeBloomFilter();
    f->mix = mix != 0;
    f->bf.init((size_t)nums);
    return f;
}
void orc_rf_free(void* f) { delete (orc::RuntimeBloomFilter*)f; }
void orc_rf_insert_hash(void* f, uint64_t hash) { ((orc::RuntimeBloomFilter*)f)->bf.insert_hash(hash); }
int orc_rf_test_hash(void* f, uint64_t hash) { return ((orc::RuntimeBloomFilter*)f)->bf.test_hash(hash) ? 1 : 0; }
// keys: n values of `width` bytes (1/2/4/8, signed), nulls optional
void orc_rf_insert_column(void* f, const void* keys, int width, int64_t n, const uint8_t* nulls) {
    auto* rf = (orc::RuntimeBloomFilter*)f;
    for (int64_t i = 0; i < n; i++) {
        if (nulls && nulls[i]) {
            rf->insert_null();
            continue;
        }
        int64_t v = width == 1 ? ((const int8_t*)keys)[i] : width == 2 ? ((const int16_t*)keys)[i] : width == 4 ? ((const int32_t*)keys)[i] : ((const int64_t*)keys)[i];
        rf->insert(v);
    }
}
// selection[i] = evaluate(selection[i], null, key) for every row (RuntimeBloomFilter::_t_evaluate)
void orc_rf_evaluate(void* f, const void* keys, int width, int64_t n, const uint8_t* nulls, uint8_t* selection) {
    auto* rf = (orc::RuntimeBloomFilter*)f;
    for (int64_t i = 0; i < n; i++) {
        int64_t v = width == 1 ? ((const int8_t*)keys)[i] : width == 2 ? ((const int16_t*)keys)[i] : width == 4 ? ((const int32_t*)keys)[i] : ((const int64_t*)keys)[i];
        selection[i] = rf->evaluate(selection[i], nulls && nulls[i], v);
    }
}
int orc_rf_test_data(void* f, int64_t v) { return ((orc::RuntimeBloomFilter*)f)->test_data(v) ? 1 : 0; }
int64_t orc_rf_info(void* f, int what) { // 0 n_buckets, 1 min, 2 max, 3 has_null, 4 inserted
    auto* rf = (orc::RuntimeBloomFilter*)f;
    switch (what) {
    case 0: return (int64_t)1 << rf->bf.log_num_buckets;
    case 1: return rf->min;
    case 2: return rf->max;
    case 3: return rf->has_null ? 1 : 0;
    default: return rf->inserted;
    }
}
void orc_rf_directory(void* f, uint32_t* out) {
    auto* rf = (orc::RuntimeBloomFilter*)f;
    std::copy(rf->bf.directory.begin(), rf->bf.directory.end(), out);
}

// RuntimeBitsetFilter (orc_rf.h)
int orc_rf_choose_bitset(int64_t min, int64_t max, uint64_t row_count, uint64_t l2_cache, uint64_t l3_cache) {
    return orc::choose_bitset_filter(min, max, row_count, l2_cache, l3_cache) ? 1 : 0;
}
void* orc_bitset_create(int64_t min, int64_t max) {
    auto* f = new orc::RuntimeBitsetFilter();
    f->set_min_max(min, max);
    f->init();
    return f;
}
void orc_bitset_free(void* f) { delete (orc::RuntimeBitsetFilter*)f; }
void orc_bitset_insert_column(void* f, const void* keys, int width, int64_t n, const uint8_t* nulls) {
    auto* rf = (orc::RuntimeBitsetFilter*)f;
    for (int64_t i = 0; i < n; i++) {
        if (nulls && nulls[i]) {
            rf->has_null = true;
            continue;
        }
        rf->insert(width == 1 ? ((const int8_t*)keys)[i] : width == 2 ? ((const int16_t*)keys)[i] : width == 4 ? ((const int32_t*)keys)[i] : ((const int64_t*)keys)[i]);
    }
}
void orc_bitset_evaluate(void* f, const void* keys, int width, int64_t n, const uint8_t* nulls, uint8_t* selection) {
    auto* rf = (orc::RuntimeBitsetFilter*)f;
    for (int64_t i = 0; i < n; i++) {
        int64_t v = width == 1 ? ((const int8_t*)keys)[i] : width == 2 ? ((const int16_t*)keys)[i] : width == 4 ? ((const int32_t*)keys)[i] : ((const int64_t*)keys)[i];
        selection[i] = rf->evaluate(selection[i], nulls && nulls[i], v);
    }
}
int64_t orc_bitset_bytes(void* f) { return (int64_t)((orc::RuntimeBitsetFilter*)f)->bitset.size(); }
void orc_bitset_data(void* f, uint8_t* out) {
    auto* rf = (orc::RuntimeBitsetFilter*)f;
    std::copy(rf->bitset.begin(), rf->bitset.end(), out);
}

#if defined(__SSE4_2__)
// hardware cross-check of the software CRC32C
uint32_t orc_hw_crc32c_u64(uint32_t crc, uint64_t v) { return (uint32_t)_mm_crc32_u64(crc, v); }
uint32_t orc_hw_crc32c_u8(uint32_t crc, uint8_t v) { return _mm_crc32_u8(crc, v); }
int orc_has_hw_crc() { return 1; }
#else
int orc_has_hw_crc() { return 0; }
#endif
}
