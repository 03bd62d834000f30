// b200ans host-side prelude coder: normalised frequencies <-> bytes.
//
// Produces / consumes exactly the reference's prelude so that a model written by either
// side loads on the other:
//   ans_serialize_interp   <- /root/reference/include/ans_util.hpp:46-63
//   ans_load_interp        <- ans_util.hpp:25-42
//   interp_encode/decode   <- interp.hpp:100-118 (interpolative_internal::encode/decode),
//                             recursion of interp.hpp:67-98, centred minimal binary codes of
//                             interp.hpp:28-65, LSB-first 32-bit word packing of bits.hpp:146-218
//   vbyte_encode/decode    <- vbyte.hpp:57-95
// north_star: "The prelude (interp.hpp interpolative coding) stays on the host."
//
// One deliberate difference: the reference's bit_stream never initialises its 3-word buffer
// (bits.hpp:148), so the unused high bits of its LAST word are stack garbage.  Here they are
// zero.  Every decoder ignores those bits.
#pragma once

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

namespace b200ans {

inline void vbyte_encode_u32(uint8_t*& out, uint32_t x)
{
    for (; x >= 0x80; x >>= 7)
        *out++ = uint8_t(x | 0x80);
    *out++ = uint8_t(x);
}

inline uint32_t vbyte_decode_u32(const uint8_t*& in)
{
    uint32_t x = 0;
    for (uint32_t shift = 0;; shift += 7) {
        const uint8_t c = *in++;
        x += uint32_t(c & 0x7F) << shift;
        if (c < 0x80)
            return x;
    }
}

// The same for untrusted input: never reads at or beyond `end`, never shifts by 32 or more (a u32 has at most five
// groups); `ok` is cleared when the number is cut off or too long.
inline uint32_t vbyte_decode_u32(const uint8_t*& in, const uint8_t* end, bool& ok)
{
    uint32_t x = 0;
    for (uint32_t shift = 0; shift < 35; shift += 7) {
        if (in >= end)
            break;
        const uint8_t c = *in++;
        x += uint32_t(c & 0x7F) << shift;
        if (c < 0x80)
            return x;
    }
    ok = false;
    return 0;
}

namespace detail {

    // LSB-first packer of (value, width<=32) fields into little-endian 32-bit words at an
    // arbitrarily aligned address.
    class word_packer {
    public:
        explicit word_packer(uint8_t* dst) : out_(dst) {}
        void put(uint64_t value, unsigned width)
        {
            if (width == 0)
                return;
            acc_ |= (value & (~uint64_t(0) >> (64 - width))) << fill_;
            fill_ += width;
            if (fill_ >= 32)
                spill();
        }
        size_t finish()
        {
            if (fill_ != 0) {
                fill_ = 32;
                spill();
                acc_ = 0;
                fill_ = 0;
            }
            return bytes_;
        }

    private:
        void spill()
        {
            const uint32_t w = uint32_t(acc_);
            std::memcpy(out_, &w, sizeof w);
            out_ += sizeof w;
            bytes_ += sizeof w;
            acc_ >>= 32;
            fill_ -= 32;
        }
        uint8_t* out_;
        uint64_t acc_ = 0;
        unsigned fill_ = 0;
        size_t bytes_ = 0;
    };

    class word_unpacker {
    public:
        // `end` (optional): one past the last readable byte; a word that would cross it reads as zero and sets overrun()
        explicit word_unpacker(const uint8_t* src, const uint8_t* end = nullptr) : in_(src), end_(end) {}
        uint64_t get(unsigned width)
        {
            if (width == 0)
                return 0;
            if (fill_ < width) {
                uint32_t w = 0;
                if (end_ && size_t(end_ - in_) < sizeof w)
                    overrun_ = true;
                else
                    std::memcpy(&w, in_, sizeof w);
                in_ += sizeof w;
                words_++;
                acc_ |= uint64_t(w) << fill_;
                fill_ += 32;
            }
            const uint64_t v = acc_ & (~uint64_t(0) >> (64 - width));
            acc_ >>= width;
            fill_ -= width;
            return v;
        }
        size_t words() const { return words_; }
        bool overrun() const { return overrun_; }

    private:
        const uint8_t* in_;
        const uint8_t* end_;
        bool overrun_ = false;
        uint64_t acc_ = 0;
        unsigned fill_ = 0;
        size_t words_ = 0;
    };

    inline unsigned bit_width_minus1(uint64_t x) { return x ? 63u - unsigned(__builtin_clzll(x)) : 0u; }

    // Centred minimal binary code for val in [1, u]: the 2^b - u short codewords sit in the
    // middle of the range (rotation by u - d/2).
    inline void put_centered(word_packer& w, uint64_t val, uint64_t u)
    {
        if (u == 1)
            return;
        const unsigned b = bit_width_minus1(u - 1) + 1;
        const uint64_t span = uint64_t(1) << b;
        const uint64_t longs = 2 * u - span; // number of (b)-bit codewords
        val += u - (longs >> 1);
        if (val > u)
            val -= u;
        const uint32_t shorts = uint32_t(span - u);
        if (val <= shorts) {
            w.put(uint32_t(val - 1), b - 1);
        } else {
            val += shorts;
            w.put(uint32_t((val - 1) >> 1), b - 1);
            w.put((val - 1) & 1, 1);
        }
    }

    inline uint64_t get_centered(word_unpacker& r, uint64_t u)
    {
        if (u == 1)
            return 1; // (1 + 0) stays in range
        const unsigned b = bit_width_minus1(u - 1) + 1;
        const uint64_t span = uint64_t(1) << b;
        const uint64_t longs = 2 * u - span;
        const uint64_t shorts = span - u;
        uint64_t val = r.get(b - 1) + 1;
        if (val > shorts)
            val = 2 * val + r.get(1) - shorts - 1;
        val += longs >> 1;
        if (val > u)
            val -= u;
        return val;
    }

    // binary interpolative code over a strictly increasing sequence; values are coded +1
    inline void interp_put(word_packer& w, const uint32_t* v, size_t n, uint64_t low, uint64_t high)
    {
        while (n != 0) {
            const size_t h = (n + 1) >> 1, left = h - 1, right = n - h;
            const uint64_t mid = uint64_t(v[h - 1]) + 1;
            put_centered(w, mid - low - left + 1, high - right - low - left + 1);
            interp_put(w, v, left, low, mid - 1);
            // tail call on the right half
            v += h;
            n = right;
            low = mid + 1;
        }
    }

    inline void interp_get(word_unpacker& r, uint32_t* v, size_t n, uint64_t low, uint64_t high)
    {
        while (n != 0 && !r.overrun()) {
            const size_t h = (n + 1) >> 1, left = h - 1, right = n - h;
            const uint64_t mid = low + left - 1 + get_centered(r, high - right - low - left + 1);
            v[h - 1] = uint32_t(mid - 1);
            interp_get(r, v, left, low, mid - 1);
            v += h;
            n = right;
            low = mid + 1;
        }
    }
} // namespace detail

// n increasing values < u  ->  32-bit words at dst (may be unaligned). Returns bytes written.
inline size_t interp_encode(void* dst, const uint32_t* in, size_t n, size_t u)
{
    detail::word_packer w(static_cast<uint8_t*>(dst));
    detail::interp_put(w, in, n, 1, uint64_t(u) + 1);
    return w.finish();
}

// Returns the number of 32-bit words consumed.
inline size_t interp_decode(const void* src, uint32_t* out, size_t n, size_t u)
{
    detail::word_unpacker r(static_cast<const uint8_t*>(src));
    detail::interp_get(r, out, n, 1, uint64_t(u) + 1);
    return r.words();
}

// vbyte(max_sym) . u8(log2 frame) . interp(prefix sums of (nfreq+1)); advances out_u8 past all of
// it and returns the interp bytes only (like the reference).
inline size_t ans_serialize_interp(std::vector<uint32_t>& vec, size_t frame_size, uint8_t*& out_u8)
{
    const uint32_t max_sym = uint32_t(vec.size() - 1);
    vbyte_encode_u32(out_u8, max_sym);
    *out_u8++ = uint8_t(63 - __builtin_clzll(uint64_t(frame_size))); // frame_size is a power of two
    std::vector<uint32_t> climbing(vec.size());
    uint32_t run = vec[0];
    climbing[0] = run;
    for (size_t s = 1; s < vec.size(); s++) {
        run += vec[s] + 1;
        climbing[s] = run;
    }
    const size_t bytes = interp_encode(out_u8, climbing.data(), climbing.size(), frame_size + vec.size() + 1);
    out_u8 += bytes;
    return bytes;
}

// Optional second argument: receives the number of prelude bytes consumed (the reference
// never needs it because its decoder walks the payload from the far end).
inline std::vector<uint32_t> ans_load_interp(const uint8_t* in_u8, size_t* consumed = nullptr)
{
    const uint8_t* p = in_u8;
    const uint32_t max_sym = vbyte_decode_u32(p);
    const uint64_t frame_size = uint64_t(1) << (*p++);
    std::vector<uint32_t> vec(size_t(max_sym) + 1);
    const size_t words = interp_decode(p, vec.data(), vec.size(), frame_size + vec.size() + 1);
    uint32_t below = vec[0];
    for (size_t s = 1; s < vec.size(); s++) {
        const uint32_t here = vec[s];
        vec[s] = here - below - 1;
        below = here;
    }
    if (consumed)
        *consumed = size_t(p - in_u8) + words * sizeof(uint32_t);
    return vec;
}

// ans_load_interp for untrusted bytes: nothing at or beyond `end` is read, the alphabet is limited to `max_symbols`
// and log2 M to `max_log2`; returns false (vec unspecified) when the prelude is cut off or implausible.  Whether the
// frequencies add up to the frame is the caller's check.
inline bool ans_load_interp_checked(const uint8_t* in_u8, const uint8_t* end, size_t max_symbols, unsigned max_log2,
    std::vector<uint32_t>& vec, size_t* consumed = nullptr)
{
    const uint8_t* p = in_u8;
    bool ok = true;
    const uint32_t max_sym = vbyte_decode_u32(p, end, ok);
    if (!ok || p >= end || size_t(max_sym) + 1 > max_symbols)
        return false;
    const unsigned lg = *p++;
    if (lg > max_log2)
        return false;
    const uint64_t frame_size = uint64_t(1) << lg;
    vec.assign(size_t(max_sym) + 1, 0u);
    detail::word_unpacker r(p, end);
    detail::interp_get(r, vec.data(), vec.size(), 1, frame_size + vec.size() + 2);
    if (r.overrun())
        return false;
    uint32_t below = vec[0];
    for (size_t s = 1; s < vec.size(); s++) {
        const uint32_t here = vec[s];
        if (here <= below) // prefix sums of (nfreq + 1) climb strictly
            return false;
        vec[s] = here - below - 1;
        below = here;
    }
    if (consumed)
        *consumed = size_t(p - in_u8) + r.words() * sizeof(uint32_t);
    return true;
}

} // namespace b200ans
