// b200ans_methods.hpp -- header-only C++17 adaptors over the C ABI of include/b200ans.h.
//
// Drop-in for the four ANS method structs of the reference (mpetri/ans-large-alphabet,
// include/methods.hpp) and for the coder / model free functions they call:
//
//   reference (file:line)                                   this header
//   ------------------------------------------------------  ---------------------------------------------
//   struct ANSint            methods.hpp:484-497            struct ANSint_b200            name() == "ANS"
//   struct ANSmsb            methods.hpp:499-512            struct ANSmsb_b200            name() == "ANSmsb"
//   template ANSfold<f>      methods.hpp:529-547            template ANSfold_b200<f>      "ANSfold-f"
//   template ANSrfold<f>     methods.hpp:549-567            template ANSrfold_b200<f>     "ANSrfold-f"
//   template ANSsint<H>      methods.hpp:584-601            template ANSsint_b200<H>      "ANSsint-H"
//   template ANSsmsb<H>      methods.hpp:603-620            template ANSsmsb_b200<H>      "ANSsmsb-H"
//   ans_int_compress / _decompress      ans_int.hpp:197,257             b200::ans_int_compress / _decompress
//   ans_fold_compress<f> / _decompress  ans_fold.hpp:238,283            b200::ans_fold_compress<f> / ...
//   ans_msb_compress / _decompress      ans_msb.hpp:235,294             b200::ans_msb_compress / ...
//   ans_reorder_fold_compress<f> / ...  ans_reorder_fold.hpp:312,357    b200::ans_reorder_fold_compress<f> / ...
//   adjust_freqs             ans_util.hpp:100-157           b200::adjust_freqs
//   ans_serialize_interp     ans_util.hpp:46-63             b200::ans_serialize_interp
//   ans_load_interp          ans_util.hpp:25-42             b200::ans_load_interp
//   next_power_of_two / is_power_of_two  ans_util.hpp:65,74 b200::next_power_of_two / is_power_of_two
//
// Same static signatures as the reference structs -- `static std::string name()`,
// `static size_t encode(const uint32_t* in, size_t n, uint8_t* out, size_t out_cap, uint8_t* buf = NULL)`,
// `static void decode(const uint8_t* in, size_t in_bytes, uint32_t* out, size_t n, uint8_t* buf = NULL)` -- so they
// slot into any `run<t_compressor>` template of the reference's drivers (table_efficiency.cpp:64,
// benchmark.cpp:64, table_effectiveness.cpp, pseudo_adaptive.cpp) unchanged.  `buf` is ignored, as the reference's
// ANS coders ignore it.  Differences a caller can observe:
//   * the bytes between encode() and decode() are a b200ans container (64-byte header, reference-format prelude,
//     block size table, block payloads that are each byte-identical to the reference coder run on that block),
//     unless B200ANS_BLOCK_SYMS=0 is set, which selects the reference's own wire format (one block);
//   * the destination capacity IS checked (the reference never looks at it);
//   * errors end the process through quit()-like behaviour (message on stderr, exit(EXIT_FAILURE)), exactly what the
//     reference does on its fatal paths (util.hpp:101-115); the C ABI underneath returns error codes instead.
// One context per thread and GPU (device from B200ANS_DEVICE, default 0) is created on first use.
#pragma once

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "b200ans.h"

namespace b200 {

[[noreturn]] inline void quit(const char* what, int code, b200ans_ctx* ctx)
{
    std::fprintf(stderr, "b200ans: %s: %s (%s)\n", what, ctx ? b200ans_last_error(ctx) : "", b200ans_strerror(code));
    std::exit(EXIT_FAILURE);
}

// thread-local context, created on first use
inline b200ans_ctx* context()
{
    struct holder {
        b200ans_ctx* ctx = nullptr;
        ~holder() { b200ans_destroy(ctx); }
    };
    static thread_local holder h;
    if (!h.ctx) {
        const char* dev = std::getenv("B200ANS_DEVICE");
        const int rc = b200ans_create(&h.ctx, dev ? std::atoi(dev) : 0, nullptr);
        if (rc != B200ANS_OK)
            quit("no usable CUDA device (there is no CPU fallback)", rc, nullptr);
        if (const char* bs = std::getenv("B200ANS_BLOCK_SYMS")) {
            const int rc2 = b200ans_set_block_syms(h.ctx, uint32_t(std::strtoul(bs, nullptr, 10)));
            if (rc2 != B200ANS_OK)
                quit("B200ANS_BLOCK_SYMS", rc2, h.ctx);
        }
    }
    return h.ctx;
}

inline size_t compress(int method, uint32_t fidelity, uint8_t* dst, size_t dst_capacity, const uint32_t* src, size_t n)
{
    b200ans_ctx* ctx = context();
    size_t bytes = 0;
    const int rc = b200ans_compress(ctx, method, fidelity, dst, dst_capacity, src, n, &bytes);
    if (rc != B200ANS_OK)
        quit("compress", rc, ctx);
    return bytes;
}

inline void decompress(int method, uint32_t fidelity, uint32_t* dst, size_t n, const uint8_t* src, size_t src_bytes)
{
    b200ans_ctx* ctx = context();
    const int rc = b200ans_decompress(ctx, method, fidelity, dst, n, src, src_bytes);
    if (rc != B200ANS_OK)
        quit("decompress", rc, ctx);
}

// ---- the L2 free functions (same shape as the reference's) ------------------------------------------------------
inline size_t ans_int_compress(uint8_t* dst, size_t dstCapacity, const uint32_t* src, size_t srcSize)
{
    return compress(B200ANS_METHOD_INT, 0, dst, dstCapacity, src, srcSize);
}
inline void ans_int_decompress(uint32_t* dst, size_t to_decode, const uint8_t* cSrc, size_t cSrcSize)
{
    decompress(B200ANS_METHOD_INT, 0, dst, to_decode, cSrc, cSrcSize);
}
inline size_t ans_msb_compress(uint8_t* dst, size_t dstCapacity, const uint32_t* src, size_t srcSize)
{
    return compress(B200ANS_METHOD_MSB, 0, dst, dstCapacity, src, srcSize);
}
inline void ans_msb_decompress(uint32_t* dst, size_t to_decode, const uint8_t* cSrc, size_t cSrcSize)
{
    decompress(B200ANS_METHOD_MSB, 0, dst, to_decode, cSrc, cSrcSize);
}
template <uint32_t fidelity> size_t ans_fold_compress(uint8_t* dst, size_t dstCapacity, const uint32_t* src, size_t srcSize)
{
    return compress(B200ANS_METHOD_FOLD, fidelity, dst, dstCapacity, src, srcSize);
}
template <uint32_t fidelity> void ans_fold_decompress(uint32_t* dst, size_t to_decode, const uint8_t* cSrc, size_t cSrcSize)
{
    decompress(B200ANS_METHOD_FOLD, fidelity, dst, to_decode, cSrc, cSrcSize);
}
template <uint32_t fidelity>
size_t ans_reorder_fold_compress(uint8_t* dst, size_t dstCapacity, const uint32_t* src, size_t srcSize)
{
    return compress(B200ANS_METHOD_RFOLD, fidelity, dst, dstCapacity, src, srcSize);
}
template <uint32_t fidelity>
void ans_reorder_fold_decompress(uint32_t* dst, size_t to_decode, const uint8_t* cSrc, size_t cSrcSize)
{
    decompress(B200ANS_METHOD_RFOLD, fidelity, dst, to_decode, cSrc, cSrcSize);
}

// ANSsint<H> / ANSsmsb<H> (ans_sint.hpp:197, ans_smsb.hpp:235): the same coders with adjust_freqs' H_approx = H
template <uint32_t H_approx> size_t ans_sint_compress(uint8_t* dst, size_t dstCapacity, const uint32_t* src, size_t srcSize)
{
    b200ans_set_h_approx(context(), H_approx);
    const size_t bytes = compress(B200ANS_METHOD_INT, 0, dst, dstCapacity, src, srcSize);
    b200ans_set_h_approx(context(), 1);
    return bytes;
}
inline void ans_sint_decompress(uint32_t* dst, size_t to_decode, const uint8_t* cSrc, size_t cSrcSize)
{
    decompress(B200ANS_METHOD_INT, 0, dst, to_decode, cSrc, cSrcSize);
}
template <uint32_t H_approx> size_t ans_smsb_compress(uint8_t* dst, size_t dstCapacity, const uint32_t* src, size_t srcSize)
{
    b200ans_set_h_approx(context(), H_approx);
    const size_t bytes = compress(B200ANS_METHOD_MSB, 0, dst, dstCapacity, src, srcSize);
    b200ans_set_h_approx(context(), 1);
    return bytes;
}
inline void ans_smsb_decompress(uint32_t* dst, size_t to_decode, const uint8_t* cSrc, size_t cSrcSize)
{
    decompress(B200ANS_METHOD_MSB, 0, dst, to_decode, cSrc, cSrcSize);
}

// ---- the ans_util.hpp model API (host only; no GPU needed) -------------------------------------------------------
inline uint64_t next_power_of_two(uint64_t x)
{
    if (x <= 1)
        return 1;
    uint64_t p = 1;
    while (p < x)
        p <<= 1;
    return p;
}
inline bool is_power_of_two(uint64_t x) { return x != 0 && (x & (x - 1)) == 0; }

inline std::vector<uint32_t> adjust_freqs(
    const std::vector<uint64_t>& freqs, uint32_t largest_sym, bool require_u16, uint32_t H_approx = 1)
{
    std::vector<uint32_t> out(size_t(largest_sym) + 1);
    const int rc = b200ans_adjust_freqs(freqs.data(), freqs.size(), largest_sym, require_u16 ? 1 : 0, H_approx, out.data());
    if (rc != B200ANS_OK)
        quit("adjust_freqs", rc, nullptr);
    return out;
}

// writes vbyte(max_sym), log2(frame), interpolative code; advances out_u8; returns the interp bytes (ans_util.hpp:46-63)
inline size_t ans_serialize_interp(std::vector<uint32_t>& vec, size_t frame_size, uint8_t*& out_u8)
{
    size_t bytes = 0;
    const size_t cap = vec.size() * 8 + 64;
    const int rc = b200ans_serialize_interp(vec.data(), vec.size(), frame_size, out_u8, cap, &bytes);
    if (rc != B200ANS_OK)
        quit("ans_serialize_interp", rc, nullptr);
    // header = vbyte(max_sym) + 1 byte: the reference returns the interp part only
    size_t header = 1;
    for (uint32_t m = uint32_t(vec.size() - 1); m >= 128; m >>= 7)
        header++;
    header++;
    out_u8 += bytes;
    return bytes - header;
}

inline std::vector<uint32_t> ans_load_interp(const uint8_t* in_u8)
{
    size_t count = 0, used = 0;
    int rc = b200ans_load_interp(in_u8, nullptr, 0, &count, &used);
    if (rc != B200ANS_OK)
        quit("ans_load_interp", rc, nullptr);
    std::vector<uint32_t> out(count);
    rc = b200ans_load_interp(in_u8, out.data(), out.size(), &count, &used);
    if (rc != B200ANS_OK)
        quit("ans_load_interp", rc, nullptr);
    return out;
}

} // namespace b200

// ---- the method structs (methods.hpp:484-567) ---------------------------------------------------------------------
struct ANSint_b200 {
    static std::string name() { return std::string("ANS"); }
    static size_t encode(const uint32_t* in_ptr, size_t in_size_u32, uint8_t* out_ptr, size_t out_size_u8, uint8_t* buf = NULL)
    {
        (void)buf;
        return b200::ans_int_compress(out_ptr, out_size_u8, in_ptr, in_size_u32);
    }
    static void decode(const uint8_t* in_ptr, size_t in_size_u8, uint32_t* out_ptr, size_t out_size_u32, uint8_t* buf = NULL)
    {
        (void)buf;
        b200::ans_int_decompress(out_ptr, out_size_u32, in_ptr, in_size_u8);
    }
};

struct ANSmsb_b200 {
    static std::string name() { return "ANSmsb"; }
    static size_t encode(const uint32_t* in_ptr, size_t in_size_u32, uint8_t* out_ptr, size_t out_size_u8, uint8_t* buf = NULL)
    {
        (void)buf;
        return b200::ans_msb_compress(out_ptr, out_size_u8, in_ptr, in_size_u32);
    }
    static void decode(const uint8_t* in_ptr, size_t in_size_u8, uint32_t* out_ptr, size_t out_size_u32, uint8_t* buf = NULL)
    {
        (void)buf;
        b200::ans_msb_decompress(out_ptr, out_size_u32, in_ptr, in_size_u8);
    }
};

template <uint32_t H_approx> struct ANSsint_b200 { // methods.hpp:584-601
    static std::string name() { return std::string("ANSsint-") + std::to_string(H_approx); }
    static size_t encode(const uint32_t* in_ptr, size_t in_size_u32, uint8_t* out_ptr, size_t out_size_u8, uint8_t* buf = NULL)
    {
        (void)buf;
        return b200::ans_sint_compress<H_approx>(out_ptr, out_size_u8, in_ptr, in_size_u32);
    }
    static void decode(const uint8_t* in_ptr, size_t in_size_u8, uint32_t* out_ptr, size_t out_size_u32, uint8_t* buf = NULL)
    {
        (void)buf;
        b200::ans_sint_decompress(out_ptr, out_size_u32, in_ptr, in_size_u8);
    }
};

template <uint32_t H_approx> struct ANSsmsb_b200 { // methods.hpp:603-620
    static std::string name() { return std::string("ANSsmsb-") + std::to_string(H_approx); }
    static size_t encode(const uint32_t* in_ptr, size_t in_size_u32, uint8_t* out_ptr, size_t out_size_u8, uint8_t* buf = NULL)
    {
        (void)buf;
        return b200::ans_smsb_compress<H_approx>(out_ptr, out_size_u8, in_ptr, in_size_u32);
    }
    static void decode(const uint8_t* in_ptr, size_t in_size_u8, uint32_t* out_ptr, size_t out_size_u32, uint8_t* buf = NULL)
    {
        (void)buf;
        b200::ans_smsb_decompress(out_ptr, out_size_u32, in_ptr, in_size_u8);
    }
};

template <uint32_t fidelity> struct ANSfold_b200 {
    static std::string name() { return std::string("ANSfold-") + std::to_string(fidelity); }
    static size_t encode(const uint32_t* in_ptr, size_t in_size_u32, uint8_t* out_ptr, size_t out_size_u8, uint8_t* buf = NULL)
    {
        (void)buf;
        return b200::ans_fold_compress<fidelity>(out_ptr, out_size_u8, in_ptr, in_size_u32);
    }
    static void decode(const uint8_t* in_ptr, size_t in_size_u8, uint32_t* out_ptr, size_t out_size_u32, uint8_t* buf = NULL)
    {
        (void)buf;
        b200::ans_fold_decompress<fidelity>(out_ptr, out_size_u32, in_ptr, in_size_u8);
    }
};

template <uint32_t fidelity> struct ANSrfold_b200 {
    static std::string name() { return std::string("ANSrfold-") + std::to_string(fidelity); }
    static size_t encode(const uint32_t* in_ptr, size_t in_size_u32, uint8_t* out_ptr, size_t out_size_u8, uint8_t* buf = NULL)
    {
        (void)buf;
        return b200::ans_reorder_fold_compress<fidelity>(out_ptr, out_size_u8, in_ptr, in_size_u32);
    }
    static void decode(const uint8_t* in_ptr, size_t in_size_u8, uint32_t* out_ptr, size_t out_size_u32, uint8_t* buf = NULL)
    {
        (void)buf;
        b200::ans_reorder_fold_decompress<fidelity>(out_ptr, out_size_u32, in_ptr, in_size_u8);
    }
};
