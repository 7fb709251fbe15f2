// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_arith.h header). Hash functions of the reference, restated.
// Pinned by be/test/exec/join_hash_map_test.cpp:914-1023 (JoinKeyHash / calc_bucket_num vectors);
// phmap_mix and FNV have no known-answer test in the reference ("parity unpinned" for those two; neither is
// observable in order-normalised results).
#pragma once
#include <cstdint>
#include <cstring>
#if defined(__SSE4_2__)
#include <nmmintrin.h>
#endif

namespace orc {

// CRC32C (Castagnoli), bit-reflected, as computed by the x86 crc32 instruction the reference calls through
// _mm_crc32_u8/u32/u64 (be/src/util/hash.h:98-170). Software form so the oracle does not depend on -msse4.2;
// cross-checked against the intrinsic when it is available (tests).
inline uint32_t crc32c_byte(uint32_t crc, uint8_t b) {
    crc ^= b;
    for (int k = 0; k < 8; k++) crc = (crc >> 1) ^ (0x82F63B78u & (0u - (crc & 1u)));
    return crc;
}
inline uint32_t crc32c_u32(uint32_t crc, uint32_t v) {
    for (int i = 0; i < 4; i++) crc = crc32c_byte(crc, (uint8_t)(v >> (8 * i)));
    return crc;
}
inline uint64_t crc32c_u64(uint64_t crc, uint64_t v) { // _mm_crc32_u64: uses the low 32 bits of crc, zero-extends
    uint32_t c = (uint32_t)crc;
    for (int i = 0; i < 8; i++) c = crc32c_byte(c, (uint8_t)(v >> (8 * i)));
    return (uint64_t)c;
}

// be/src/util/hash.h:98-133
inline uint32_t crc_hash_32(const void* data, int32_t bytes, uint32_t hash) {
    uint32_t words = bytes / 4;
    bytes = bytes % 4;
    const uint8_t* p = (const uint8_t*)data;
    while (words--) {
        uint32_t w;
        memcpy(&w, p, 4);
        hash = crc32c_u32(hash, w);
        p += 4;
    }
    while (bytes--) {
        hash = crc32c_byte(hash, *p);
        ++p;
    }
    hash = (hash << 16u) | (hash >> 16u);
    return hash;
}

// be/src/util/hash.h:136-170
inline uint64_t crc_hash_64(const void* data, int32_t length, uint64_t hash) {
    if (length < 8) return crc_hash_32(data, length, (uint32_t)hash);
    uint64_t words = length / 8;
    const uint8_t* p = (const uint8_t*)data;
    const uint8_t* end = p + length;
    while (words--) {
        uint64_t w;
        memcpy(&w, p, 8);
        hash = crc32c_u64(hash, w);
        p += 8;
    }
    p = end - 8;
    uint64_t w;
    memcpy(&w, p, 8);
    hash = crc32c_u64(hash, w);
    return hash;
}

static const uint32_t CRC_HASH_SEED1 = 0x811C9DC5; // be/src/util/hash.h:172-176
static const uint32_t CRC_HASH_SEED2 = 0x811C9DD7;

// be/src/util/hash.h:15-19,64-96 — phmap_mix_with_seed<8, seed>
inline uint64_t phmap_mix8(uint64_t a, int seed /*1 or 2*/) {
    const uint64_t k = (seed == 1) ? 0xde5fb9d2630458e9ULL : 0xde5fb9d263046000ULL;
    unsigned __int128 r = (unsigned __int128)a * k;
    return (uint64_t)(r >> 64) + (uint64_t)r;
}
// StdHashWithSeed<T,seed> for integral T: std::hash is the identity on the sign-extended value (hash.h:211-216)
inline uint64_t std_hash_with_seed(int64_t v, int seed) { return phmap_mix8((uint64_t)v, seed); }

// be/src/column/column_hash.h:114-147 — hash_combine / hash_128 / Hash128WithSeed
inline void hash_combine_u64(uint64_t& seed, uint64_t val) {
    seed ^= val + 0x9e3779b97f4a7c15ULL + (seed << 12) + (seed >> 4);
}
inline uint64_t hash_128(uint64_t seed, __int128 val) {
    hash_combine_u64(seed, (uint64_t)val);
    hash_combine_u64(seed, (uint64_t)(val >> 64));
    return seed;
}
inline uint64_t hash128_with_seed(__int128 v, int seed /*PhmapSeed enum value + 1*/) {
    // Hash128WithSeed passes the enum (PhmapSeed1 = 0, PhmapSeed2 = 1) as the hash_128 seed (column_hash.h:143-147)
    return phmap_mix8(hash_128((uint64_t)(seed - 1), v), seed);
}

// be/src/util/hash_util.hpp:195-214 — FNV-1 used by the shuffle (column/column.h:376)
static const uint32_t FNV_PRIME = 0x01000193;
static const uint32_t FNV_SEED = 0x811C9DC5;
inline uint32_t fnv_hash(const void* data, int32_t bytes, uint32_t hash) {
    const uint8_t* ptr = (const uint8_t*)data;
    while (bytes--) {
        hash = (*ptr ^ hash) * FNV_PRIME;
        ++ptr;
    }
    return hash;
}
// be/src/util/hash_util.hpp:326-331, 342-344 and exec/pipeline/exchange/shuffler.h:70-85
inline uint32_t xorshift32(uint32_t x) {
    x ^= x << 13;
    x ^= x >> 17;
    x ^= x << 5;
    return x;
}
inline uint32_t reduce_op(uint32_t l, uint32_t r) { return (uint32_t)(((uint64_t)l * (uint64_t)r) >> 32); }

// ---- join hashing: be/src/exec/join_hash_map.h:310-364 ---------------------------------------------
inline uint32_t join_key_hash32(uint32_t v, uint32_t log_buckets) {
    v ^= v >> (32 - log_buckets);
    return (uint32_t)(v * 2654435761u) >> (32 - log_buckets);
}
inline uint32_t join_key_hash64(uint64_t v, uint32_t log_buckets) {
    v ^= v >> (64 - log_buckets);
    return (uint32_t)((v * 11400714819323198485ull) >> (64 - log_buckets));
}
inline uint32_t join_key_hash_bytes(const void* p, int32_t n, uint32_t num_buckets) {
    return crc_hash_32(p, n, 0x811C9DC5u) & (num_buckets - 1);
}
// phmap::priv::NormalizeCapacity(n) = n ? ~size_t{} >> countl_zero(n) : 1  (be/src/util/phmap/phmap.h)
inline uint64_t normalize_capacity(uint64_t n) { return n ? (~0ULL >> __builtin_clzll(n)) : 1; }
// JoinHashMapHelper::calc_bucket_size (join_hash_map.h:357-364); size = row_count + 1
inline uint32_t calc_bucket_size(uint32_t size) {
    uint64_t expect = (uint64_t)size + (size - 1) / 4;
    if (expect >= (1u << 31)) return 1u << 31;
    return (uint32_t)(normalize_capacity(expect) + 1);
}

} // namespace orc
