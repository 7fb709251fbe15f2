// srb_dev.cuh — device-side arithmetic and hash primitives shared by every kernel of libsrb.so.
//
// Everything here reproduces, bit for bit, the integer semantics of the reference BE on x86-64:
//   overflow-checked +,-,* on int32/int64      be/src/runtime/integer_overflow_arithmetics.h:23-138
//   int128 add/sub (OF flag of adc/sbb)         be/src/runtime/int128_arithmetics_x86_64.h:148-181
//   int128 multiply = multi3 on magnitudes      be/src/runtime/int128_arithmetics_x86_64.h:184-229
//   i64 x i64 -> i128 "fast mul"                be/src/exprs/arithmetic_expr.cpp:70-101
//   DecimalV3Arithmetics::div_round             be/src/runtime/decimalv3.h:86-100
//   decimal_div_integer (AVG finalize)          be/src/exprs/arithmetic_operation.h:409-441
//   phmap_mix / StdHashWithSeed / hash_128      be/src/util/hash.h:64-96,211-216, column/column_hash.h:114-147
//   crc_hash_32 / crc_hash_64 (CRC32C)          be/src/util/hash.h:98-170
//   JoinKeyHash / calc_bucket_size              be/src/exec/join_hash_map.h:310-364
//   FNV-1 shuffle hash + ReduceOp               be/src/util/hash_util.hpp:195-214, exec/pipeline/exchange/shuffler.h:70-85
// Functions are __host__ __device__ so the host operators (AVG finalize of a handful of groups, bucket sizing)
// run the very same code; nothing here is a CPU fallback for a kernel.
#pragma once
#if defined(__CUDACC_RTC__) // compiled at run time by NVRTC (srb_jit.h): no host headers, the fixed-width types are spelled out
typedef signed char int8_t;
typedef short int16_t;
typedef int int32_t;
typedef long long int64_t;
typedef unsigned char uint8_t;
typedef unsigned short uint16_t;
typedef unsigned int uint32_t;
typedef unsigned long long uint64_t;
typedef unsigned long uintptr_t;
#define INT32_MIN (-2147483647 - 1)
#define INT32_MAX 2147483647
#define INT64_MIN (-9223372036854775807LL - 1)
#define INT64_MAX 9223372036854775807LL
#define UINT32_MAX 4294967295U
#define UINT64_MAX 18446744073709551615ULL
#else
#include <cstdint>
#endif

#if defined(__CUDACC__)
#define SRB_HD __host__ __device__ __forceinline__
#define SRB_D __device__ __forceinline__
#else
#define SRB_HD inline
#define SRB_D inline
#endif

namespace srb {

typedef __int128 i128;
typedef unsigned __int128 u128;

SRB_HD i128 make_i128(uint64_t lo, int64_t hi) { return (i128)(((u128)(uint64_t)hi << 64) | (u128)lo); }
SRB_HD uint64_t lo64(i128 x) { return (uint64_t)x; }
SRB_HD int64_t hi64(i128 x) { return (int64_t)((u128)x >> 64); }

SRB_HD uint64_t umulhi64(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (uint64_t)(((u128)a * b) >> 64);
#endif
}

// ---- overflow-checked arithmetic ------------------------------------------------------------------
// width-4 / width-8: the value is carried sign-extended in an i128; the result must fit the narrow type.
template <int W>
SRB_HD bool fits(i128 x) {
    if (W == 4) return x >= (i128)INT32_MIN && x <= (i128)INT32_MAX;
    if (W == 8) return x >= (i128)INT64_MIN && x <= (i128)INT64_MAX;
    return true;
}
template <int W>
SRB_HD i128 wrap_to(i128 x) {
    if (W == 1) return (i128)(int8_t)lo64(x);
    if (W == 2) return (i128)(int16_t)lo64(x);
    if (W == 4) return (i128)(int32_t)lo64(x);
    if (W == 8) return (i128)(int64_t)lo64(x);
    return x;
}

SRB_HD bool add_overflow128(i128 a, i128 b, i128* c) {
    *c = (i128)((u128)a + (u128)b);
    return ((a < 0) == (b < 0)) && ((*c < 0) != (a < 0));
}
SRB_HD bool sub_overflow128(i128 a, i128 b, i128* c) {
    *c = (i128)((u128)a - (u128)b);
    return ((a < 0) != (b < 0)) && ((*c < 0) != (a < 0));
}

// multi3 on magnitudes (int128_arithmetics_x86_64.h:184-229): overflow is flagged when a partial product spills
// past bit 127, or the 64-bit accumulation of the high word carries or sets its top bit. |x|*|y| == 2^127 is
// therefore an overflow although -2^127 is representable; __builtin_mul_overflow would disagree.
SRB_HD bool mul_overflow128(i128 x, i128 y, i128* res) {
    // common case first: both operands fit in int64 -> one signed 64x64->128 product, can never overflow
    // (|x|,|y| < 2^63 so xh = yh = 0 and the product magnitude is < 2^126: every multi3 flag is clear)
    if (fits<8>(x) && fits<8>(y) && (int64_t)lo64(x) != INT64_MIN && (int64_t)lo64(y) != INT64_MIN) {
        int64_t a = (int64_t)lo64(x), b = (int64_t)lo64(y);
        uint64_t lo = (uint64_t)a * (uint64_t)b;
        uint64_t hi = umulhi64((uint64_t)a, (uint64_t)b);
        hi -= (a < 0 ? (uint64_t)b : 0) + (b < 0 ? (uint64_t)a : 0);
        *res = make_i128(lo, (int64_t)hi);
        return false;
    }
    i128 sx = x >> 127, sy = y >> 127;
    u128 ax = (u128)((x ^ sx) - sx), ay = (u128)((y ^ sy) - sy);
    uint64_t xl = (uint64_t)ax, xh = (uint64_t)(ax >> 64), yl = (uint64_t)ay, yh = (uint64_t)(ay >> 64);
    if (!((xl | xh) && (yl | yh))) {
        // the reference returns 0 without flags only when BOTH are zero (:185-189); with one zero operand every
        // partial product is zero as well, so the general path below gives the same answer. Short-circuit both.
        *res = 0;
        return false;
    }
    bool ov = (xh != 0 && yh != 0);
    uint64_t lo = xl * yl, hi = umulhi64(xl, yl);
    uint64_t t0l = xl * yh, t0h = umulhi64(xl, yh);
    uint64_t t1l = yl * xh, t1h = umulhi64(yl, xh);
    ov |= (t0h != 0) | (t1h != 0);
    uint64_t s = hi + t0l;
    ov |= (s < hi) | ((s >> 63) != 0);
    uint64_t s2 = s + t1l;
    ov |= (s2 < s) | ((s2 >> 63) != 0);
    i128 sgn = sx ^ sy;
    u128 w = ((u128)s2 << 64) | lo;
    *res = (i128)((w ^ (u128)sgn) - (u128)sgn);
    return ov;
}

// checked op at result width W (4, 8, 16). a and b are sign-extended values of the W-byte type.
// op codes are SRB_EXPR_ADD/SUB/MUL (4/5/6).
template <int W>
SRB_HD bool checked_binop(int op, i128 a, i128 b, i128* out) {
    if (W == 16) {
        if (op == 4) return add_overflow128(a, b, out);
        if (op == 5) return sub_overflow128(a, b, out);
        return mul_overflow128(a, b, out);
    }
    // narrow types: operands fit in W bytes, so the exact result fits in an i128 (|a*b| < 2^126)
    i128 r = op == 4 ? a + b : (op == 5 ? a - b : a * b);
    *out = wrap_to<W>(r);
    return !fits<W>(r);
}

SRB_HD i128 mul_i64_i64(int64_t a, int64_t b) {
    uint64_t lo = (uint64_t)a * (uint64_t)b;
    uint64_t hi = umulhi64((uint64_t)a, (uint64_t)b);
    hi -= (a < 0 ? (uint64_t)b : 0) + (b < 0 ? (uint64_t)a : 0);
    return make_i128(lo, (int64_t)hi);
}

SRB_HD i128 exp10_i128(int n) {
    i128 v = 1;
    for (int i = 0; i < n; i++) v *= 10;
    return v;
}

// DecimalV3Arithmetics<int128,false>::div_round (decimalv3.h:86-100): truncating divide, then round half away
// from zero. b == 0 never reaches here (count > 0 is checked by the caller, avg.h:226-229).
SRB_HD i128 div_round128(i128 a, i128 b) {
    i128 c = a / b;
    i128 r = a % b;
    i128 carry = ((a ^ b) >> 127) | 1;
    i128 abs_b = b < 0 ? -b : b;
    i128 abs_r = r < 0 ? -r : r;
    bool need_carry = ((abs_b >> 1) + (abs_b & 1)) <= abs_r;
    c += carry & -(i128)need_carry;
    return c;
}
SRB_HD int div_result_scale(int lhs_scale) {
    if (lhs_scale <= 6) return lhs_scale + 6;
    if (lhs_scale <= 12) return 12;
    return lhs_scale;
}
SRB_HD i128 decimal_div_integer(i128 dividend, i128 count, int dividend_scale) {
    int adjust = div_result_scale(dividend_scale) - dividend_scale;
    i128 scaled = (i128)((u128)dividend * (u128)exp10_i128(adjust)); // unchecked scale_up wraps (decimalv3.h:252-255)
    return div_round128(scaled, count);
}

// ---- hashes ----------------------------------------------------------------------------------------
SRB_HD uint64_t phmap_mix8(uint64_t a, int seed /*1 = phase 1, 2 = phase 2*/) {
    const uint64_t k = (seed == 1) ? 0xde5fb9d2630458e9ULL : 0xde5fb9d263046000ULL;
    return umulhi64(a, k) + a * k;
}
SRB_HD void hash_combine_u64(uint64_t& seed, uint64_t val) { seed ^= val + 0x9e3779b97f4a7c15ULL + (seed << 12) + (seed >> 4); }
SRB_HD uint64_t hash128_with_seed(uint64_t lo, uint64_t hi, int seed) {
    uint64_t h = (uint64_t)(seed - 1);
    hash_combine_u64(h, lo);
    hash_combine_u64(h, hi);
    return phmap_mix8(h, seed);
}

// CRC32C (Castagnoli, reflected) — what the x86 crc32 instruction computes. 4-bit table keeps it register/constant
// friendly on the device; only slice keys and 16-byte join keys hash through it.
SRB_HD uint32_t crc32c_byte(uint32_t crc, uint8_t b) {
    crc ^= b;
#pragma unroll
    for (int k = 0; k < 8; k++) crc = (crc >> 1) ^ (0x82F63B78u & (0u - (crc & 1u)));
    return crc;
}
SRB_HD uint32_t crc32c_u32(uint32_t crc, uint32_t v) {
#pragma unroll
    for (int i = 0; i < 4; i++) crc = crc32c_byte(crc, (uint8_t)(v >> (8 * i)));
    return crc;
}
SRB_HD uint32_t crc32c_u64(uint32_t crc, uint64_t v) {
#pragma unroll
    for (int i = 0; i < 8; i++) crc = crc32c_byte(crc, (uint8_t)(v >> (8 * i)));
    return crc;
}
// crc_hash_32 over a key held in registers/local bytes (util/hash.h:98-133)
SRB_HD uint32_t crc_hash_32(const uint8_t* p, int bytes, uint32_t hash) {
    int words = bytes / 4;
    bytes = bytes % 4;
    while (words--) {
        uint32_t w = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
        hash = crc32c_u32(hash, w);
        p += 4;
    }
    while (bytes--) hash = crc32c_byte(hash, *p++);
    return (hash << 16) | (hash >> 16);
}
SRB_HD uint64_t load_u64_le(const uint8_t* p) {
    uint64_t w = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) w |= (uint64_t)p[i] << (8 * i);
    return w;
}
// crc_hash_64 (util/hash.h:136-170): _mm_crc32_u64 keeps only the low 32 bits of the running value
SRB_HD uint64_t crc_hash_64(const uint8_t* p, int length, uint64_t hash) {
    if (length < 8) return crc_hash_32(p, length, (uint32_t)hash);
    int words = length / 8;
    const uint8_t* end = p + length;
    uint32_t h = (uint32_t)hash;
    while (words--) {
        h = crc32c_u64(h, load_u64_le(p));
        p += 8;
    }
    h = crc32c_u64(h, load_u64_le(end - 8));
    return (uint64_t)h;
}

// JoinKeyHash (join_hash_map.h:310-350)
SRB_HD uint32_t join_key_hash32(uint32_t v, uint32_t log_buckets) {
    v ^= v >> (32 - log_buckets);
    return (uint32_t)(v * 2654435761u) >> (32 - log_buckets);
}
SRB_HD uint32_t join_key_hash64(uint64_t v, uint32_t log_buckets) {
    v ^= v >> (64 - log_buckets);
    return (uint32_t)((v * 11400714819323198485ull) >> (64 - log_buckets));
}
SRB_HD uint32_t join_key_hash128(uint64_t lo, uint64_t hi, uint32_t num_buckets) {
    uint32_t h = 0x811C9DC5u; // crc_hash_32 over the 16 key bytes as four little-endian words
    h = crc32c_u32(h, (uint32_t)lo);
    h = crc32c_u32(h, (uint32_t)(lo >> 32));
    h = crc32c_u32(h, (uint32_t)hi);
    h = crc32c_u32(h, (uint32_t)(hi >> 32));
    h = (h << 16) | (h >> 16);
    return h & (num_buckets - 1);
}
SRB_HD uint64_t normalize_capacity(uint64_t n) {
#if defined(__CUDA_ARCH__)
    return n ? (~0ULL >> __clzll((long long)n)) : 1;
#else
    return n ? (~0ULL >> __builtin_clzll(n)) : 1;
#endif
}
SRB_HD uint32_t calc_bucket_size(uint32_t size) {
    uint64_t expect = (uint64_t)size + (size - 1) / 4;
    if (expect >= (1u << 31)) return 1u << 31;
    return (uint32_t)(normalize_capacity(expect) + 1);
}

// FNV-1 over the value bytes, as Column::fnv_hash does for fixed-length columns (hash_util.hpp:195-214)
SRB_HD uint32_t fnv_hash_value(uint64_t lo, uint64_t hi, int width, uint32_t hash) {
    for (int i = 0; i < width; i++) {
        uint8_t b = (uint8_t)((i < 8 ? lo >> (8 * i) : hi >> (8 * (i - 8))) & 0xff);
        hash = (b ^ hash) * 0x01000193u;
    }
    return hash;
}
// ---- runtime bloom filter: SimdBlockFilter (exprs/runtime_filter.h:79-222, runtime_filter.cpp:26-35,114-123) ----------------------
// A directory of 2^log_num_buckets buckets of 8 x uint32; a hash sets / tests one bit in each of the 8 words of ONE bucket:
//   bucket = hash & directory_mask;  word i: bit ((uint32)(hash >> log_num_buckets) * SALT[i]) >> 27.
// RuntimeBloomFilter::compute_hash (runtime_filter.h:1192-1198) = phmap_mix<8>(std::hash<T>(v)); ENABLE_PHMAP_MIX is not defined
// anywhere in the reference tree, so phmap_mix<8> is the identity (util/phmap/phmap_utils.h:112-124) and the hash of an
// integer key is its sign-extended value.
struct RfView {
    uint32_t* dir; // [n_buckets][8]
    uint32_t directory_mask;
    int32_t log_num_buckets;
    int64_t min, max; // over the inserted non-NULL keys (min > max: nothing inserted)
    int32_t has_null;
    int32_t mix; // 0: hash = key (the reference tree as it builds: phmap_mix is the identity); 1: hash = phmap_mix<8>(key), what the
                 // same source computes when ENABLE_PHMAP_MIX is defined (phmap_utils.h:85-95)
    // RuntimeBitsetFilter (exprs/runtime_filter.h:1583-1700): one bit per value of [min, max], bit (x - min) of a little-endian bit string
    // (Bitset::insert, runtime_filter.cpp:287-292); nullptr: the filter is a bloom filter
    const uint32_t* bits;
};
SRB_HD uint32_t rf_salt(int i) {
    switch (i) {
    case 0: return 0x47b6137bu;
    case 1: return 0x44974d91u;
    case 2: return 0x8824ad5bu;
    case 3: return 0xa2b7289du;
    case 4: return 0x705495c7u;
    case 5: return 0x2df1424bu;
    case 6: return 0x9efc4947u;
    default: return 0x5c6bfb31u;
    }
}
SRB_HD uint32_t rf_word_mask(uint32_t key, int i) { return 1u << ((key * rf_salt(i)) >> 27); }
SRB_HD uint64_t rf_hash(long long key, int mix) { return mix ? phmap_mix8((uint64_t)key, 1) : (uint64_t)key; }
SRB_HD int rf_log_num_buckets(uint64_t nums) { // SimdBlockFilter::init: max(1, ceil(log2(max(1, nums))) - 5)
    if (nums < 1) nums = 1;
    int log_heap = 0;
    while (((uint64_t)1 << log_heap) < nums) log_heap++;
    return log_heap - 5 > 1 ? log_heap - 5 : 1;
}

SRB_HD uint32_t reduce_op(uint32_t h, uint32_t n) { return (uint32_t)(((uint64_t)h * (uint64_t)n) >> 32); }
SRB_HD uint32_t xorshift32(uint32_t x) {
    x ^= x << 13;
    x ^= x >> 17;
    x ^= x << 5;
    return x;
}

} // namespace srb
