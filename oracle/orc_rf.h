// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_arith.h header): CPU restatement of the join runtime filter.
//
//   SimdBlockFilter                 be/src/exprs/runtime_filter.h:79-222 (insert_hash / test_hash, scalar branch :113-117,152-159),
//                                   be/src/exprs/runtime_filter.cpp:26-35 (init), :114-123 (make_mask), SALT runtime_filter.h:58-59
//   RuntimeBloomFilter::compute_hash / insert / test_data / _t_evaluate     be/src/exprs/runtime_filter.h:1192-1205,1322-1329,1369-1422
//   min / max of the inserted values (evaluated before the bloom test)      be/src/exprs/runtime_filter.h (MinMaxRuntimeFilter)
//
// compute_hash = phmap_mix<sizeof(size_t)>()(std::hash<T>()(v)). ENABLE_PHMAP_MIX is not defined anywhere in the reference tree,
// so phmap_mix<8> is the identity (be/src/util/phmap/phmap_utils.h:112-124) and std::hash of an integer is its value converted
// to size_t (sign-extended): the hash of an integer key is the key.
// Pinned by be/test/exprs/runtime_filter_test.cpp:255-265 (TestSimdBlockFilter) and :342-380 (TestJoinRuntimeFilter),
// restated in tests/test_oracle_golden.py.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

namespace orc {

static const uint32_t RF_SALT[8] = {0x47b6137b, 0x44974d91, 0x8824ad5b, 0xa2b7289d, 0x705495c7, 0x2df1424b, 0x9efc4947, 0x5c6bfb31};

struct SimdBlockFilter {
    static constexpr int LOG_BUCKET_BYTE_SIZE = 5;
    int log_num_buckets = 0;
    uint32_t directory_mask = 0;
    std::vector<uint32_t> directory; // [n_buckets][8]

    void init(size_t nums) { // runtime_filter.cpp:26-35
        nums = std::max<size_t>(1, nums);
        int log_heap_space = (int)std::ceil(std::log2((double)nums));
        log_num_buckets = std::max(1, log_heap_space - LOG_BUCKET_BYTE_SIZE);
        directory_mask = (uint32_t)((1ull << std::min(63, log_num_buckets)) - 1);
        directory.assign((size_t)8 << log_num_buckets, 0u);
    }
    static void make_mask(uint32_t key, uint32_t* masks) { // runtime_filter.cpp:114-123
        for (int i = 0; i < 8; ++i) {
            masks[i] = key * RF_SALT[i];
            masks[i] = masks[i] >> 27;
            masks[i] = 0x1u << masks[i];
        }
    }
    void insert_hash(uint64_t hash) { // runtime_filter.h:97-118
        const uint32_t bucket_idx = (uint32_t)(hash & directory_mask);
        uint32_t masks[8];
        make_mask((uint32_t)(hash >> log_num_buckets), masks);
        for (int i = 0; i < 8; ++i) directory[(size_t)bucket_idx * 8 + i] |= masks[i];
    }
    bool test_hash(uint64_t hash) const { // runtime_filter.h:121-160
        const uint32_t bucket_idx = (uint32_t)(hash & directory_mask);
        uint32_t masks[8];
        make_mask((uint32_t)(hash >> log_num_buckets), masks);
        for (int i = 0; i < 8; ++i)
            if ((directory[(size_t)bucket_idx * 8 + i] & masks[i]) == 0) return false;
        return true;
    }
};

// RuntimeBloomFilter over integer keys (any width up to 8 bytes, sign-extended to int64)
struct RuntimeBloomFilter {
    SimdBlockFilter bf;
    int64_t min = INT64_MAX, max = INT64_MIN;
    bool has_null = false;
    int64_t inserted = 0;
    bool mix = false; // true: phmap_mix<8> as compiled with ENABLE_PHMAP_MIX (phmap_utils.h:85-95): hi + lo of v * 0xde5fb9d2630458e9
    uint64_t compute_hash(int64_t v) const {
        if (!mix) return (uint64_t)v; // identity mix of the identity std::hash
        unsigned __int128 r = (unsigned __int128)(uint64_t)v * 0xde5fb9d2630458e9ULL;
        return (uint64_t)(r >> 64) + (uint64_t)r;
    }
    void insert(int64_t v) {
        bf.insert_hash(compute_hash(v));
        min = std::min(min, v);
        max = std::max(max, v);
        inserted++;
    }
    void insert_null() { has_null = true; }
    bool test_data(int64_t v) const { return bf.test_hash(compute_hash(v)); }
    // _t_evaluate on one row: selection &= (in [min,max]) && test_data; NULL rows: selection = has_null
    uint8_t evaluate(uint8_t selected, bool is_null, int64_t v) const {
        if (is_null) return has_null ? 1 : 0;
        if (!selected) return 0;
        if (v < min || v > max) return 0;
        return test_data(v) ? 1 : 0;
    }
};

// RuntimeBitsetFilter over integer keys: Bitset (be/src/exprs/runtime_filter.h:1587-1655, runtime_filter.cpp:287-292,350-354) behind the
// ComposedRuntimeFilter's min / max test (contains<CheckRange>); NULL rows: selection = has_null (contains_batch<.., NullIsTrue>).
// Pinned by be/test/exprs/runtime_filter_test.cpp:1409-1455 (TestEvaluateBitsetFilter), restated in tests/test_oracle_golden.py.
struct RuntimeBitsetFilter {
    int64_t min = 0, max = -1;
    bool has_null = false;
    std::vector<uint8_t> bitset;
    void set_min_max(int64_t mn, int64_t mx) {
        min = mn;
        max = mx;
    }
    void init() { // value_interval() + 1 values, eight per byte
        const uint64_t key_interval = ((uint64_t)max - (uint64_t)min) + 1;
        bitset.assign((size_t)((key_interval + 7) / 8), 0);
    }
    void insert(int64_t v) {
        const uint64_t bucket = (uint64_t)v - (uint64_t)min;
        bitset[bucket / 8] |= (uint8_t)(1 << (bucket % 8));
    }
    uint8_t evaluate(uint8_t selected, bool is_null, int64_t v) const {
        if (is_null) return has_null ? 1 : 0;
        if (!selected || v < min || v > max) return 0;
        const uint64_t bucket = (uint64_t)v - (uint64_t)min;
        return (bitset[bucket / 8] >> (bucket % 8)) & 1;
    }
};
// try_create_runtime_bitset_filter (be/src/exprs/runtime_filter_bank.cpp:160-200): the bitset replaces the bloom filter when it needs no more
// memory than the bloom filter's budget (row_count bytes), or at most 4x / 2x that and fits the L2 / L3 cache
inline bool choose_bitset_filter(int64_t min, int64_t max, uint64_t row_count, uint64_t l2_cache, uint64_t l3_cache) {
    const uint64_t value_interval = (uint64_t)max - (uint64_t)min + 1;
    if (min > max || value_interval == 0) return false;
    const uint64_t bitset_mem = (value_interval + 7) / 8, bloom_mem = row_count;
    return bitset_mem <= bloom_mem || (bitset_mem <= bloom_mem * 4 && bitset_mem <= l2_cache) || (bitset_mem <= bloom_mem * 2 && bitset_mem <= l3_cache);
}

} // namespace orc
