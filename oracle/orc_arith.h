// ORACLE — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's arithmetic for the
// scan -> filter -> aggregate / join path. Nothing under starrocks_b200/ may include, link or call this.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
//
// Parity status: PINNED by the reference's own known-answer tests (see tests/test_oracle_golden.py):
//   be/test/runtime/decimalv3_test.cpp:402-491, be/test/exprs/agg/aggregate_test.cpp:331-468,
//   be/test/exec/join_hash_map_test.cpp:914-1056, be/test/column/date_value_test.cpp.
// The reference BE itself cannot be compiled here (needs glog/gflags/boost/brpc/thrift codegen), so there is
// no oracle/_ref; see DESIGN.md.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>

namespace orc {

typedef __int128 int128_t;
typedef unsigned __int128 uint128_t;

// ---- dates: be/src/runtime/time_types.cpp:474-491 (date::from_date) -------------------------------
inline int32_t from_date(int year, int month, int day) {
    int century, julian;
    if (month > 2) {
        month += 1;
        year += 4800;
    } else {
        month += 13;
        year += 4799;
    }
    century = year / 100;
    julian = year * 365 - 32167;
    julian += year / 4 - century + century / 4;
    julian += 7834 * month / 256 + day;
    return julian;
}

// inverse: be/src/runtime/time_types.h:469-497 (date::to_date; the ZERO_EPOCH special case is not needed here)
inline void to_date(int32_t julian, int* year, int* month, int* day) {
    int quad, extra, y;
    julian += 32044;
    quad = julian / 146097;
    extra = (julian - quad * 146097) * 4 + 3;
    julian += 60 + quad * 3 + extra / 146097;
    quad = julian / 1461;
    julian -= quad * 1461;
    y = julian * 4 / 1461;
    julian = ((y != 0) ? ((julian + 305) % 365) : ((julian + 306) % 366)) + 123;
    y += quad * 4;
    *year = y - 4800;
    quad = julian * 2141 / 65536;
    *day = julian - 7834 * quad / 256;
    *month = (quad + 10) % 12 + 1;
}

// ---- overflow-checked integer arithmetic: be/src/runtime/integer_overflow_arithmetics.h:23-138 ----
inline bool add_overflow(int32_t a, int32_t b, int32_t* c) { return __builtin_add_overflow(a, b, c); }
inline bool add_overflow(int64_t a, int64_t b, int64_t* c) { return __builtin_add_overflow(a, b, c); }
inline bool sub_overflow(int32_t a, int32_t b, int32_t* c) { return __builtin_sub_overflow(a, b, c); }
inline bool sub_overflow(int64_t a, int64_t b, int64_t* c) { return __builtin_sub_overflow(a, b, c); }
inline bool mul_overflow(int32_t a, int32_t b, int32_t* c) { return __builtin_mul_overflow(a, b, c); }
inline bool mul_overflow(int64_t a, int64_t b, int64_t* c) { return __builtin_mul_overflow(a, b, c); }

// int128 add/sub: asm_add_overflow / asm_sub_overflow report the CPU OF flag of the high-word adc/sbb
// (be/src/runtime/int128_arithmetics_x86_64.h:148-181) == signed 128-bit overflow.
inline bool add_overflow(int128_t a, int128_t b, int128_t* c) {
    *c = (int128_t)((uint128_t)a + (uint128_t)b);
    return ((a < 0) == (b < 0)) && ((*c < 0) != (a < 0));
}
inline bool sub_overflow(int128_t a, int128_t b, int128_t* c) {
    *c = (int128_t)((uint128_t)a - (uint128_t)b);
    return ((a < 0) != (b < 0)) && ((*c < 0) != (a < 0));
}

// int128 multiply on x86-64 is multi3 (int128_arithmetics_x86_64.h:184-229), NOT __builtin_mul_overflow:
// it multiplies |x|*|y| as unsigned words and flags overflow when any partial product spills past bit 127
// or the unsigned high-word accumulation carries or sets bit 63 (asm_add's "setc | sets", :62-75).
// So a product whose magnitude is exactly 2^127 is an overflow even when the signed result would fit.
inline int multi3_abs(uint64_t xl, uint64_t xh, uint64_t yl, uint64_t yh, uint64_t* rl, uint64_t* rh) {
    if (!((xl | xh) || (yl | yh))) { // both zero (:185-189)
        *rl = 0;
        *rh = 0;
        return 0;
    }
    int overflow = (xh != 0 && yh != 0);
    uint128_t p = (uint128_t)xl * yl; // first asm_mul: its flag is ignored (:193)
    uint64_t lo = (uint64_t)p, hi = (uint64_t)(p >> 64);
    uint128_t t0 = (uint128_t)xl * yh; // unsigned mul sets OF=CF when the high half is non-zero (:99-108)
    overflow |= ((uint64_t)(t0 >> 64) != 0);
    uint128_t t1 = (uint128_t)yl * xh;
    overflow |= ((uint64_t)(t1 >> 64) != 0);
    // asm_add<uint64,uint64>: overflow = CF | SF of the 64-bit sum (:62-75)
    uint64_t s = hi + (uint64_t)t0;
    overflow |= (s < hi) | (int)(s >> 63);
    uint64_t s2 = s + (uint64_t)t1;
    overflow |= (s2 < s) | (int)(s2 >> 63);
    *rl = lo;
    *rh = s2;
    return overflow;
}
inline bool mul_overflow(int128_t x, int128_t y, int128_t* res) {
    int128_t sx = x >> 127, sy = y >> 127;
    uint128_t ax = (uint128_t)((x ^ sx) - sx), ay = (uint128_t)((y ^ sy) - sy);
    uint64_t rl, rh;
    int ov = multi3_abs((uint64_t)ax, (uint64_t)(ax >> 64), (uint64_t)ay, (uint64_t)(ay >> 64), &rl, &rh);
    sx ^= sy;
    uint128_t w = ((uint128_t)rh << 64) | rl;
    *res = (int128_t)((w ^ (uint128_t)sx) - (uint128_t)sx);
    return ov != 0;
}

// i64 x i64 -> i128 "fast mul": one signed imul, cannot overflow (int128_arithmetics_x86_64.h:198-202)
inline int128_t i64_x_i64_produce_i128(int64_t a, int64_t b) { return (int128_t)a * (int128_t)b; }

// ---- DecimalV3Arithmetics<T,check>::div_round: be/src/runtime/decimalv3.h:86-100 --------------------
template <typename T>
inline T abs_t(T v) {
    return v < 0 ? -v : v;
}
template <typename T>
inline T div_round(T a, T b) {
    T c = a / b;
    T r = a % b;
    T carry = ((a ^ b) >> (sizeof(T) * 8 - 1)) | 1;
    T abs_b = abs_t(b);
    T abs_r = abs_t(r);
    bool need_carry = ((abs_b >> 1) + (abs_b & 1)) <= abs_r;
    c += carry & -T(need_carry);
    return c;
}

inline int128_t exp10_128(int n) {
    int128_t v = 1;
    for (int i = 0; i < n; i++) v *= 10;
    return v;
}

// scale of a / b for decimals: be/src/exprs/arithmetic_operation.h:409-422
inline int div_result_scale(int lhs_scale) {
    if (lhs_scale <= 6) return lhs_scale + 6;
    if (lhs_scale <= 12) return 12;
    return lhs_scale;
}

// AVG finalize: be/src/exprs/arithmetic_operation.h:429-441 (decimal_div_integer<int128>)
inline int128_t decimal_div_integer(int128_t dividend, int128_t count, int dividend_scale) {
    int adjust_scale = div_result_scale(dividend_scale) + 0 - dividend_scale;
    int128_t f = exp10_128(adjust_scale);
    // DecimalV3Cast::to_decimal<T,T,T,true,false> == unchecked scale_up: plain wrapping multiply (decimalv3.h:252-255)
    int128_t scaled = (int128_t)((uint128_t)dividend * (uint128_t)f);
    return div_round<int128_t>(scaled, count);
}

} // namespace orc
