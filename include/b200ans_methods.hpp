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
//   ans_byte_compress / _decompress     ans_byte.hpp:236,279            b200::ans_byte_compress / _decompress
//   adjust_freqs             ans_util.hpp:100-157           b200::adjust_freqs
//   ans_serialize_interp     ans_util.hpp:46-63             b200::ans_serialize_interp
//   ans_load_interp          ans_util.hpp:25-42             b200::ans_load_interp
//   next_power_of_two / is_power_of_two  ans_util.hpp:65,74 b200::next_power_of_two / is_power_of_two
//   scale_freqs              ans_util.hpp:77-95             b200::scale_freqs
//   ans_fold_encode<f> / ans_fold_decode<f>   ans_fold.hpp:67-125,177-235     b200::ans_fold_encode<f> / _decode<f>
//   ans_int_encode / _decode, ans_msb_encode / _decode, ans_reorder_fold_encode<f> / _decode<f>     likewise
//     (create / serialize / load and the public fields; the per-symbol steps are GPU block coders here, see below)
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
#include <cstring>
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

// ans_byte_compress / _decompress (ans_byte.hpp:236,279): the byte coder of the vbyte+ANS baselines
inline size_t ans_byte_compress(void* dst, size_t dstCapacity, const void* src, size_t srcSize)
{
    b200ans_ctx* ctx = context();
    size_t bytes = 0;
    const int rc = b200ans_byte_compress(ctx, static_cast<uint8_t*>(dst), dstCapacity, static_cast<const uint8_t*>(src), srcSize, &bytes);
    if (rc != B200ANS_OK)
        quit("ans_byte_compress", rc, ctx);
    return bytes;
}
inline void ans_byte_decompress(void* dst, size_t to_decode, const void* cSrc, size_t cSrcSize)
{
    b200ans_ctx* ctx = context();
    const int rc = b200ans_byte_decompress(ctx, static_cast<uint8_t*>(dst), to_decode, static_cast<const uint8_t*>(cSrc), cSrcSize);
    if (rc != B200ANS_OK)
        quit("ans_byte_decompress", rc, ctx);
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

// one fit of the counts F to the frame M, symbols in `mapping` order; true when the fit FAILED (the reference's return
// value, ans_util.hpp:77-95: `return M != 0` with the frame left over)
inline bool scale_freqs(std::vector<uint32_t>& S, const std::vector<uint64_t>& F, std::vector<uint32_t>& mapping, int64_t M,
    size_t sigma, size_t freq_sum)
{
    int fits = 0;
    if (S.size() < F.size())
        S.resize(F.size());
    const int rc = b200ans_scale_freqs(S.data(), F.data(), F.size(), mapping.data(), sigma, M, freq_sum, &fits);
    if (rc != B200ANS_OK)
        quit("scale_freqs", rc, nullptr);
    return fits == 0;
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

// ---- the model structs (ans_fold.hpp:67-125,177-235; ans_int.hpp:38-110,112-195; ans_msb.hpp:75-128,177-233;
//      ans_reorder_fold.hpp:70-186,233-310) ------------------------------------------------------------------------
// Same shape as the reference's: `T::create(in_u32, n)` builds the model from the input (histogram on the GPU,
// adjust_freqs on the host), `serialize(out_u8)` writes the prelude, advances the pointer and returns what the
// reference's returns; `T::load(in_u8, in_bytes)` reads a prelude back.  Public fields as in the reference: nfreqs,
// table (encoder: {freq, base, sym_upper_bound}), frame_size, lower_bound, frame_mask, frame_log2, and for ANSrfold
// most_frequent and sigma.  What is NOT here, on purpose: the per-symbol steps encode_symbol / decode_sym /
// flush_state / init_state -- they would be a CPU coder, and this library has none.  Their place is taken by
// `encode(dst, capacity, in_u32, n)`, which runs the block coders on the GPU (a b200ans container; the prelude inside
// it is the one serialize() writes), and by b200::ans_X_decompress for the way back.
// create() / load() make the model the current encoder / decoder model of the thread's context.
struct enc_entry { // ans_fold.hpp:30-34 (freq widened to 32 bits so that plain ANS fits too, ans_int.hpp:32-36)
    uint32_t freq;
    uint32_t base;
    uint64_t sym_upper_bound;
};

template <int METHOD, uint32_t fidelity> struct encode_model {
    static encode_model create(const uint32_t* in_u32, size_t n)
    {
        b200ans_ctx* ctx = context();
        int rc = b200ans_model_create(ctx, METHOD, fidelity, in_u32, n);
        if (rc != B200ANS_OK)
            quit("create", rc, ctx);
        encode_model m;
        size_t count = 0;
        rc = b200ans_model_nfreqs(ctx, nullptr, 0, &count, &m.frame_size);
        if (rc == B200ANS_OK) {
            m.nfreqs.resize(count);
            rc = b200ans_model_nfreqs(ctx, m.nfreqs.data(), m.nfreqs.size(), &count, &m.frame_size);
        }
        if (rc != B200ANS_OK)
            quit("create", rc, ctx);
        m.lower_bound = 16 * m.frame_size; // constants::K * frame_size (ans_fold.hpp:91)
        m.table.resize(m.nfreqs.size());
        uint64_t base = 0;
        for (size_t s = 0; s < m.nfreqs.size(); s++) {
            m.table[s] = enc_entry { m.nfreqs[s], uint32_t(base), (uint64_t(16) << 32) * m.nfreqs[s] };
            base += m.nfreqs[s];
        }
        if (METHOD == B200ANS_METHOD_RFOLD) {
            size_t k = 0;
            rc = b200ans_model_most_frequent(ctx, nullptr, 0, &k, &m.sigma);
            m.most_frequent.resize(k);
            if (rc == B200ANS_OK && k)
                rc = b200ans_model_most_frequent(ctx, m.most_frequent.data(), k, &k, &m.sigma);
            if (rc != B200ANS_OK)
                quit("create", rc, ctx);
        }
        size_t pb = 0;
        b200ans_model_prelude(ctx, nullptr, 0, &pb);
        m.prelude.resize(pb);
        rc = b200ans_model_prelude(ctx, m.prelude.data(), m.prelude.size(), &pb);
        if (rc != B200ANS_OK)
            quit("create", rc, ctx);
        return m;
    }

    // writes the prelude, advances out_u8; returns what the reference's serialize() returns: the interpolative part's
    // bytes (ans_util.hpp:46-63 -- not the vbyte + log2 M header), plus, for ANSrfold, the flag word and the
    // most_frequent list (ans_reorder_fold.hpp:132-154)
    size_t serialize(uint8_t*& out_u8)
    {
        std::memcpy(out_u8, prelude.data(), prelude.size());
        out_u8 += prelude.size();
        size_t header = 2; // vbyte(max_sym) + the log2 M byte
        for (uint32_t m = uint32_t(nfreqs.size() - 1); m >= 128; m >>= 7)
            header++;
        return prelude.size() - header;
    }

    uint64_t initial_state() const { return lower_bound; }

    // the block coders on the GPU; the model is a pure function of the input, so this is the one-shot call (it
    // rebuilds the same model on the device).  Writes a b200ans container, returns its size.
    size_t encode(uint8_t* dst, size_t dstCapacity, const uint32_t* in_u32, size_t n) const
    {
        return compress(METHOD, fidelity, dst, dstCapacity, in_u32, n);
    }

    std::vector<uint32_t> nfreqs;
    std::vector<enc_entry> table;
    std::vector<uint32_t> most_frequent; // ANSrfold
    uint64_t frame_size = 0;
    uint64_t lower_bound = 0;
    uint64_t sigma = 0;                  // ANSrfold
    std::vector<uint8_t> prelude;        // (not in the reference: the bytes serialize() writes)
};

template <int METHOD, uint32_t fidelity> struct decode_model {
    // the reference's load(in_u8) trusts the stream; here the number of readable bytes is part of the call
    static decode_model load(const uint8_t* in_u8, size_t in_bytes, size_t* consumed = nullptr)
    {
        b200ans_ctx* ctx = context();
        size_t used = 0;
        const int rc = b200ans_model_load(ctx, METHOD, fidelity, in_u8, in_bytes, &used); // bounded, validated
        if (rc != B200ANS_OK)
            quit("load", rc, ctx);
        decode_model m;
        const uint8_t* p = in_u8;
        if (METHOD == B200ANS_METHOD_RFOLD) { // flag [+ most_frequent] (ans_reorder_fold.hpp:238-252)
            uint32_t flag;
            std::memcpy(&flag, p, 4);
            p += 4;
            if (flag) {
                m.most_frequent.resize(size_t(1) << (fidelity + 7));
                std::memcpy(m.most_frequent.data(), p, 4 * m.most_frequent.size());
                p += 4 * m.most_frequent.size();
            }
        }
        m.nfreqs = ans_load_interp(p);
        for (uint32_t v : m.nfreqs)
            m.frame_size += v;
        m.frame_mask = m.frame_size - 1;
        while ((uint64_t(1) << m.frame_log2) < m.frame_size)
            m.frame_log2++;
        m.lower_bound = 16 * m.frame_size;
        if (consumed)
            *consumed = used;
        return m;
    }

    std::vector<uint32_t> nfreqs;
    std::vector<uint32_t> most_frequent; // ANSrfold
    uint64_t frame_size = 0;
    uint64_t frame_mask = 0;
    uint64_t frame_log2 = 0;
    uint64_t lower_bound = 0;
};

template <uint32_t fidelity> using ans_fold_encode = encode_model<B200ANS_METHOD_FOLD, fidelity>;
template <uint32_t fidelity> using ans_fold_decode = decode_model<B200ANS_METHOD_FOLD, fidelity>;
template <uint32_t fidelity> using ans_reorder_fold_encode = encode_model<B200ANS_METHOD_RFOLD, fidelity>;
template <uint32_t fidelity> using ans_reorder_fold_decode = decode_model<B200ANS_METHOD_RFOLD, fidelity>;
using ans_int_encode = encode_model<B200ANS_METHOD_INT, 0>;
using ans_int_decode = decode_model<B200ANS_METHOD_INT, 0>;
using ans_msb_encode = encode_model<B200ANS_METHOD_MSB, 0>;
using ans_msb_decode = decode_model<B200ANS_METHOD_MSB, 0>;

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
