// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_arith.h header).
// Segment data pages of fixed-width columns, restating:
//   ForEncoder<T>::put_batch / bit_pack / bit_packing_one_frame_value / flush    be/src/util/frame_of_reference_coding.cpp:44-241
//   ForDecoder<T>::init / bit_unpack / decode_current_frame / get_batch          be/src/util/frame_of_reference_coding.cpp:243-410
//   bits() (significant bits of a value)                                         be/src/util/frame_of_reference_coding.h:50-71
//   FrameOfReferencePageBuilder / Decoder (a page IS the encoder's buffer)       be/src/storage/rowset/frame_of_reference_page.h:46-223
//   PlainPageBuilder / Decoder ([u32 count][values])                             be/src/storage/rowset/plain_page.h:50-160
// Page layout (frame_of_reference_coding.h:78-101): frames of 128 values, each [min: sizeof(T) bytes LE][values: bit_width bits each, packed
// MSB first], then the footer: (u8 storage_format, u8 bit_width) per frame, u8 128, u32 value count. storage_format 0: value - min;
// 1 (frame is non-decreasing): value - previous value, first delta 0; 2 (the subtraction would overflow): the value itself, bit_width =
// 8 * sizeof(T). T is the column's signed C++ type (int32_t for INT / DATE / DECIMAL32, int64_t for BIGINT / DECIMAL64, __int128 for
// LARGEINT / DECIMAL128): min / max / ordering are signed compares.
// Pinned by tests/test_oracle_golden.py to be/test/util/frame_of_reference_coding_test.cpp and
// be/test/storage/rowset/frame_of_reference_page_test.cpp (round trips of 64 / 128 / 256 / 320 values, the 27- and 123-byte pages,
// the 5-byte empty page, bits() of 0x06 / 0x0F / 0x100FF / 0x1FFFFFFFF / 0).
#include <cstdint>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <thread>
#include <vector>

namespace orc {

typedef __int128 i128_t;
typedef unsigned __int128 u128_t;

// frame_of_reference_coding.h:50-71: the argument is converted to uint64_t (sign-extending a negative narrow value) or taken as uint128
template <typename T>
static inline uint8_t bits_of(T v) {
    if (sizeof(T) <= 8) {
        const uint64_t u = (uint64_t)v;
        return u == 0 ? 0 : (uint8_t)(64 - __builtin_clzll(u));
    }
    const u128_t u = (u128_t)v;
    const uint64_t hi = (uint64_t)(u >> 64), lo = (uint64_t)u;
    if (hi) return (uint8_t)(128 - __builtin_clzll(hi));
    return lo ? (uint8_t)(64 - __builtin_clzll(lo)) : 0;
}

// bit_pack (frame_of_reference_coding.cpp:95-117): value i occupies bits [i * bw, (i + 1) * bw) of a big-endian bit stream, most
// significant bit first. Same stream, written a value at a time instead of a bit at a time.
template <typename U>
static void bit_pack(const U* in, int n, int bw, uint8_t* out) {
    if (n == 0 || bw == 0) return;
    const size_t bytes = ((size_t)n * bw + 7) / 8;
    memset(out, 0, bytes);
    size_t bit = 0;
    for (int i = 0; i < n; i++) {
        U v = in[i];
        if (bw < (int)sizeof(U) * 8) v &= (((U)1) << bw) - 1;
        int left = bw; // bits of v still to write, from the top
        while (left > 0) {
            const int room = 8 - (int)(bit & 7), take = left < room ? left : room;
            const uint8_t piece = (uint8_t)((v >> (left - take)) & (((U)1 << take) - 1));
            out[bit >> 3] |= (uint8_t)(piece << (room - take));
            bit += take;
            left -= take;
        }
    }
}

template <typename U>
static void bit_unpack(const uint8_t* in, int n, int bw, U* out) {
    size_t bit = 0;
    for (int i = 0; i < n; i++) {
        U v = 0;
        int left = bw;
        while (left > 0) {
            const int room = 8 - (int)(bit & 7), take = left < room ? left : room;
            const uint8_t piece = (uint8_t)((in[bit >> 3] >> (room - take)) & ((1u << take) - 1));
            v = (U)((v << take) | piece);
            bit += take;
            left -= take;
        }
        out[i] = v;
    }
}

template <typename T>
struct Unsigned;
template <>
struct Unsigned<int8_t> { typedef uint8_t type; };
template <>
struct Unsigned<int16_t> { typedef uint16_t type; };
template <>
struct Unsigned<int32_t> { typedef uint32_t type; };
template <>
struct Unsigned<int64_t> { typedef uint64_t type; };
template <>
struct Unsigned<i128_t> { typedef u128_t type; };

template <typename T>
static T type_max() {
    typedef typename Unsigned<T>::type U;
    return (T)(((U)~(U)0) >> 1);
}

// the minimum of a frame is stored in 4 bytes for types up to 32 bits wide, 8 or 16 bytes otherwise (put_fixed32/64/128_le)
template <typename T>
static constexpr int min_bytes() {
    return sizeof(T) == 16 ? 16 : (sizeof(T) == 8 ? 8 : 4);
}

template <typename T>
static void encode_frame(const T* in, int n, std::vector<uint8_t>& buf, std::vector<uint8_t>& fmts, std::vector<uint8_t>& bws) {
    typedef typename Unsigned<T>::type U;
    T mn = in[0], mx = in[0];
    bool asc = true, keep = false;
    uint8_t bw = 0;
    const T half = type_max<T>() >> 1;
    for (int i = 1; i < n; i++) {
        if (asc) {
            if (in[i] < in[i - 1]) {
                asc = false;
            } else if ((in[i] >> 1) - (in[i - 1] >> 1) > half) {
                keep = true;
            } else {
                const uint8_t b = bits_of<T>((T)(in[i] - in[i - 1]));
                if (b > bw) bw = b;
            }
        }
        if (in[i] < mn) {
            mn = in[i];
            continue;
        }
        if (in[i] > mx) mx = in[i];
    }
    if (!asc && (mx >> 1) - (mn >> 1) > half) keep = true;
    {
        uint8_t raw[16] = {0};
        if (sizeof(T) < 4) {
            const uint32_t m32 = (uint32_t)(int32_t)mn; // put_fixed32_le of the (sign-extended) narrow value
            memcpy(raw, &m32, 4);
        } else {
            memcpy(raw, &mn, sizeof(T));
        }
        buf.insert(buf.end(), raw, raw + min_bytes<T>());
    }
    std::vector<U> d((size_t)n);
    if (keep) {
        bw = (uint8_t)(sizeof(T) * 8);
        for (int i = 0; i < n; i++) d[i] = (U)in[i];
        const size_t at = buf.size();
        buf.resize(at + (size_t)n * bw); // (the reference reserves n * bit_width BYTES here: frame_of_reference_coding.cpp:176-180)
        bit_pack<U>(d.data(), n, bw, buf.data() + at);
    } else {
        if (asc) {
            d[0] = 0;
            for (int i = 1; i < n; i++) d[i] = (U)((U)in[i] - (U)in[i - 1]);
        } else {
            bw = bits_of<T>((T)(mx - mn));
            for (int i = 0; i < n; i++) d[i] = (U)((U)in[i] - (U)mn);
        }
        const size_t at = buf.size(), len = ((size_t)n * bw + 7) / 8;
        buf.resize(at + len);
        bit_pack<U>(d.data(), n, bw, buf.data() + at);
    }
    fmts.push_back(keep ? 2 : (asc ? 1 : 0));
    bws.push_back(bw);
}

template <typename T>
static std::vector<uint8_t> for_encode(const T* in, int64_t n) {
    std::vector<uint8_t> buf, fmts, bws;
    for (int64_t a = 0; a < n; a += 128) encode_frame<T>(in + a, (int)(n - a < 128 ? n - a : 128), buf, fmts, bws);
    for (size_t i = 0; i < fmts.size(); i++) {
        buf.push_back(fmts[i]);
        buf.push_back(bws[i]);
    }
    buf.push_back(128);
    const uint32_t cnt = (uint32_t)n;
    const uint8_t* c = (const uint8_t*)&cnt;
    buf.insert(buf.end(), c, c + 4);
    return buf;
}

template <typename T>
static int64_t for_decode(const uint8_t* page, int64_t len, T* out, int64_t cap) {
    typedef typename Unsigned<T>::type U;
    if (len < 5) return -1;
    const uint32_t frame = page[len - 5];
    uint32_t n;
    memcpy(&n, page + len - 4, 4);
    if (frame == 0) return n == 0 ? 0 : -1;
    const uint32_t frames = n / frame + (n % frame != 0);
    const uint32_t last = frame - (frame * frames - n);
    if ((int64_t)frames * 2 + 5 > len || (int64_t)n > cap) return -1;
    const uint8_t* meta = page + len - 5 - (size_t)frames * 2;
    size_t off = 0;
    std::vector<U> d(frame);
    for (uint32_t f = 0; f < frames; f++) {
        const uint8_t fmt = meta[2 * f], bw = meta[2 * f + 1];
        const int cnt = (int)(f == frames - 1 ? last : frame);
        T mn = 0;
        if (sizeof(T) < 4) {
            int32_t m32;
            memcpy(&m32, page + off, 4);
            mn = (T)m32;
        } else {
            memcpy(&mn, page + off, sizeof(T));
        }
        const uint8_t* body = page + off + min_bytes<T>();
        bit_unpack<U>(body, cnt, bw, d.data());
        T* o = out + (size_t)f * frame;
        if (fmt == 2) {
            for (int i = 0; i < cnt; i++) o[i] = (T)d[i];
        } else if (fmt == 1) {
            U pre = (U)mn;
            for (int i = 0; i < cnt; i++) {
                pre = (U)(pre + d[i]);
                o[i] = (T)pre;
            }
        } else {
            for (int i = 0; i < cnt; i++) o[i] = (T)((U)mn + d[i]);
        }
        off += (size_t)bw * frame / 8 + min_bytes<T>(); // ForDecoder::init: every frame but the last is full
    }
    return (int64_t)n;
}

} // namespace orc

extern "C" {
// -> page bytes (or -needed when `cap` is too small); width 1 / 2 / 4 / 8 / 16 = the signed column type
int64_t orc_for_encode(const void* values, int width, int64_t n, uint8_t* out, int64_t cap) {
    std::vector<uint8_t> b;
    switch (width) {
    case 1: b = orc::for_encode<int8_t>((const int8_t*)values, n); break;
    case 2: b = orc::for_encode<int16_t>((const int16_t*)values, n); break;
    case 4: b = orc::for_encode<int32_t>((const int32_t*)values, n); break;
    case 8: b = orc::for_encode<int64_t>((const int64_t*)values, n); break;
    case 16: b = orc::for_encode<orc::i128_t>((const orc::i128_t*)values, n); break;
    default: return 0;
    }
    if ((int64_t)b.size() > cap) return -(int64_t)b.size();
    memcpy(out, b.data(), b.size());
    return (int64_t)b.size();
}
int64_t orc_for_decode(const uint8_t* page, int64_t len, int width, void* out, int64_t cap_values) {
    switch (width) {
    case 1: return orc::for_decode<int8_t>(page, len, (int8_t*)out, cap_values);
    case 2: return orc::for_decode<int16_t>(page, len, (int16_t*)out, cap_values);
    case 4: return orc::for_decode<int32_t>(page, len, (int32_t*)out, cap_values);
    case 8: return orc::for_decode<int64_t>(page, len, (int64_t*)out, cap_values);
    case 16: return orc::for_decode<orc::i128_t>(page, len, (orc::i128_t*)out, cap_values);
    }
    return -1;
}
// many pages of one column, `threads` workers (the scan's DOP): page p = [offs[p], offs[p + 1]) of `data`, its values start at row rows[p]
int64_t orc_for_decode_pages(const uint8_t* data, const int64_t* offs, const int64_t* rows, int64_t n_pages, int width, void* out, int threads) {
    if (threads < 1) threads = 1;
    std::vector<int64_t> bad((size_t)threads, 0);
    auto work = [&](int t) {
        for (int64_t p = n_pages * t / threads; p < n_pages * (t + 1) / threads; p++) {
            const int64_t got = orc_for_decode(data + offs[p], offs[p + 1] - offs[p], width, (uint8_t*)out + (size_t)rows[p] * width, rows[p + 1] - rows[p]);
            if (got != rows[p + 1] - rows[p]) bad[(size_t)t]++;
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < threads; t++) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
    int64_t b = 0;
    for (auto x : bad) b += x;
    return b;
}
uint8_t orc_bits(int64_t lo, int64_t hi, int width) {
    switch (width) {
    case 1: return orc::bits_of<int8_t>((int8_t)lo);
    case 2: return orc::bits_of<int16_t>((int16_t)lo);
    case 4: return orc::bits_of<int32_t>((int32_t)lo);
    case 8: return orc::bits_of<int64_t>(lo);
    }
    return orc::bits_of<orc::i128_t>(((orc::i128_t)hi << 64) | (uint64_t)lo);
}
}
