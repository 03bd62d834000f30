// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// C-ABI shim around the UNMODIFIED reference headers (mpetri/ans-large-alphabet,
// /root/reference/include).  It is compiled by oracle/Makefile straight from the
// sources where they lie (-I/root/reference/include) into oracle/_ref/libansref.so
// and is used (a) to pin the C restatement in oracle/ans_oracle.c, (b) as the
// per-block / whole-stream oracle of the -m gpu parity tests and (c) as the CPU
// baseline ("kind": "reference") of bench.py.  No reference source is copied here:
// this file only *calls* the reference.
//
// Include-order note: namespace `constants` used by every coder
// lives in ans_byte.hpp, so that header comes first.
#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <random>
#include <string>
#include <vector>

#include "ans_byte.hpp"
#include "ans_fold.hpp"
#include "ans_int.hpp"
#include "ans_msb.hpp"
#include "ans_reorder_fold.hpp"
#include "ans_sint.hpp"
#include "ans_smsb.hpp"
#include "zipf_dist.hpp"

enum { M_INT = 0, M_FOLD = 1, M_MSB = 2, M_RFOLD = 3, M_BYTE = 4 }; // M_BYTE: ans_byte.hpp, the symbols are bytes handed over as uint32_t

static std::vector<uint8_t> to_bytes(const uint32_t* in, size_t n)
{
    std::vector<uint8_t> b(n);
    for (size_t i = 0; i < n; i++)
        b[i] = uint8_t(in[i]);
    return b;
}

#define FOLD_DISPATCH(f, EXPR)                                                                 \
    switch (f) {                                                                               \
    case 1: { constexpr uint32_t F = 1; EXPR; } break;                                         \
    case 2: { constexpr uint32_t F = 2; EXPR; } break;                                         \
    case 3: { constexpr uint32_t F = 3; EXPR; } break;                                         \
    case 4: { constexpr uint32_t F = 4; EXPR; } break;                                         \
    case 5: { constexpr uint32_t F = 5; EXPR; } break;                                         \
    case 6: { constexpr uint32_t F = 6; EXPR; } break;                                         \
    case 7: { constexpr uint32_t F = 7; EXPR; } break;                                         \
    case 8: { constexpr uint32_t F = 8; EXPR; } break;                                         \
    default: return -1;                                                                        \
    }

// ---------------------------------------------------------------------------
// generic per-block drivers: the loops of ans_fold.hpp:250-278 / :291-310 applied to a
// sub-range through the model's own members (SURVEY.md section 8(c) "per-block oracle").
// ---------------------------------------------------------------------------
template <class Enc>
static size_t encode_range(Enc& m, const uint32_t* in, size_t n, uint8_t* dst)
{
    uint8_t* out = dst;
    std::array<uint64_t, 4> st;
    for (auto& s : st)
        s = m.initial_state();
    size_t cur = 0;
    while ((n - cur) % 4 != 0) {
        m.encode_symbol(st[0], in[n - cur - 1], out);
        cur += 1;
    }
    while (cur != n) {
        m.encode_symbol(st[0], in[n - cur - 1], out);
        m.encode_symbol(st[1], in[n - cur - 2], out);
        m.encode_symbol(st[2], in[n - cur - 3], out);
        m.encode_symbol(st[3], in[n - cur - 4], out);
        cur += 4;
    }
    for (auto& s : st)
        m.flush_state(s, out);
    return out - dst;
}

// fold / msb / rfold decoders share decode_sym(state, in); ans_int needs the entry type.
template <class Dec, class Fn>
static void decode_range(Dec& m, const uint8_t* src, size_t bytes, uint32_t* out, size_t n, Fn sym)
{
    const uint8_t* in = src + bytes;
    std::array<uint64_t, 4> st;
    st[3] = m.init_state(in);
    st[2] = m.init_state(in);
    st[1] = m.init_state(in);
    st[0] = m.init_state(in);
    size_t i = 0, fast = n - (n % 4);
    while (i != fast) {
        out[i] = sym(m, st[3], in);
        out[i + 1] = sym(m, st[2], in);
        out[i + 2] = sym(m, st[1], in);
        out[i + 3] = sym(m, st[0], in);
        i += 4;
    }
    while (i != n)
        out[i++] = sym(m, st[0], in);
}

struct ref_model {
    int method;
    uint32_t f;
    void* enc; // ans_*_encode*
    void* dec; // ans_*_decode*
};

template <class T> static void del(void* p) { delete static_cast<T*>(p); }

// The same models from a HISTOGRAM instead of the input (2^30 / 2^32-symbol jobs, where a CPU counting pass would take
// minutes and the counts already exist): the members create() fills after its counting loop -- ans_fold.hpp:79-92,
// ans_msb.hpp:86-100, ans_int.hpp:50-66 -- are public, so they are filled here in the same order with the reference's
// own adjust_freqs; encode_symbol / flush_state / serialize are then the reference's.  `bins` must hold the
// reference's counts: 2^(f+9) bins (FOLD), 1280 (MSB), max value + 1 (INT).
template <class Enc> static Enc* enc_from_hist(const uint64_t* bins, size_t nbins, bool u16, bool skip_absent)
{
    auto* model = new Enc();
    std::vector<uint64_t> freqs(bins, bins + nbins);
    uint32_t max_sym = 0;
    for (size_t s = 0; s < nbins; s++)
        if (freqs[s])
            max_sym = uint32_t(s);
    model->nfreqs = adjust_freqs(freqs, max_sym, u16);
    model->frame_size = std::accumulate(std::begin(model->nfreqs), std::end(model->nfreqs), 0);
    uint64_t cur_base = 0;
    const uint64_t tmp = constants::K * constants::RADIX;
    model->table.resize(max_sym + 1);
    for (size_t sym = 0; sym < model->nfreqs.size(); sym++) {
        if (skip_absent && model->nfreqs[sym] == 0)
            continue;
        model->table[sym].freq = model->nfreqs[sym];
        model->table[sym].base = cur_base;
        model->table[sym].sym_upper_bound = tmp * model->nfreqs[sym];
        cur_base += model->nfreqs[sym];
    }
    model->lower_bound = constants::K * model->frame_size;
    return model;
}


extern "C" {

// ---- whole-stream reference coders (methods.hpp:484-567 call exactly these) ----------
long ref_compress(int method, uint32_t f, uint8_t* dst, size_t cap, const uint32_t* src, size_t n)
{
    switch (method) {
    case M_INT: return (long)ans_int_compress(dst, cap, src, n);
    case M_MSB: return (long)ans_msb_compress(dst, cap, src, n);
    case M_FOLD: FOLD_DISPATCH(f, return (long)ans_fold_compress<F>(dst, cap, src, n));
    case M_RFOLD: FOLD_DISPATCH(f, return (long)ans_reorder_fold_compress<F>(dst, cap, src, n));
    case M_BYTE: {
        const std::vector<uint8_t> b = to_bytes(src, n);
        return (long)ans_byte_compress(dst, cap, b.data(), n);
    }
    }
    return -1;
}

// ANSsint<H> / ANSsmsb<H> (methods.hpp:584-620): the entropy-approximation variants of benchmark.cpp:156-172
#define H_DISPATCH(h, EXPR)                                                                    \
    switch (h) {                                                                               \
    case 1: { constexpr uint32_t H = 1; EXPR; } break;                                         \
    case 5: { constexpr uint32_t H = 5; EXPR; } break;                                         \
    case 10: { constexpr uint32_t H = 10; EXPR; } break;                                       \
    case 20: { constexpr uint32_t H = 20; EXPR; } break;                                       \
    case 40: { constexpr uint32_t H = 40; EXPR; } break;                                       \
    case 80: { constexpr uint32_t H = 80; EXPR; } break;                                       \
    case 160: { constexpr uint32_t H = 160; EXPR; } break;                                     \
    case 320: { constexpr uint32_t H = 320; EXPR; } break;                                     \
    default: return -1;                                                                        \
    }
long ref_compress_h(int method, uint32_t h, uint8_t* dst, size_t cap, const uint32_t* src, size_t n)
{
    if (method == M_INT) {
        H_DISPATCH(h, return (long)ans_sint_compress<H>(dst, cap, src, n));
    } else if (method == M_MSB) {
        H_DISPATCH(h, return (long)ans_smsb_compress<H>(dst, cap, src, n));
    }
    return -1;
}
int ref_decompress_h(int method, uint32_t* dst, size_t n, const uint8_t* src, size_t bytes)
{
    if (method == M_INT)
        ans_sint_decompress(dst, n, src, bytes);
    else if (method == M_MSB)
        ans_smsb_decompress(dst, n, src, bytes);
    else
        return -1;
    return 0;
}

int ref_decompress(int method, uint32_t f, uint32_t* dst, size_t n, const uint8_t* src, size_t bytes)
{
    switch (method) {
    case M_INT: ans_int_decompress(dst, n, src, bytes); return 0;
    case M_MSB: ans_msb_decompress(dst, n, src, bytes); return 0;
    case M_FOLD: FOLD_DISPATCH(f, ans_fold_decompress<F>(dst, n, src, bytes); return 0);
    case M_RFOLD: FOLD_DISPATCH(f, ans_reorder_fold_decompress<F>(dst, n, src, bytes); return 0);
    case M_BYTE: {
        std::vector<uint8_t> b(n);
        ans_byte_decompress(b.data(), n, src, bytes);
        std::copy(b.begin(), b.end(), dst);
        return 0;
    }
    }
    return -1;
}

// ---- model API (ans_util.hpp) ----------------------------------------------------------
// out must hold largest_sym+1 entries.
int ref_adjust_freqs(const uint64_t* freqs, size_t nbins, uint32_t largest_sym, int require_u16,
    uint32_t H_approx, uint32_t* out)
{
    std::vector<uint64_t> F(freqs, freqs + nbins);
    auto s = adjust_freqs(F, largest_sym, require_u16 != 0, H_approx);
    std::copy(s.begin(), s.end(), out);
    return (int)s.size();
}

// scale_freqs (ans_util.hpp:77-95) as is: S in/out (nbins entries), returns the reference's bool ("the fit failed")
int ref_scale_freqs(uint32_t* S, const uint64_t* F, size_t nbins, const uint32_t* mapping, size_t sigma, int64_t M,
    uint64_t freq_sum)
{
    std::vector<uint32_t> s(S, S + nbins), map(mapping, mapping + sigma);
    std::vector<uint64_t> f(F, F + nbins);
    const bool failed = scale_freqs(s, f, map, M, sigma, freq_sum);
    std::copy(s.begin(), s.end(), S);
    return failed ? 1 : 0;
}

// returns total prelude bytes (vbyte + log2 byte + interp words) written to out
long ref_serialize_interp(const uint32_t* nfreqs, size_t count, uint64_t frame_size, uint8_t* out)
{
    std::vector<uint32_t> v(nfreqs, nfreqs + count);
    uint8_t* p = out;
    ans_serialize_interp(v, frame_size, p);
    return p - out;
}

// out must hold max_sym+1 entries (call with out==NULL to get the count)
long ref_load_interp(const uint8_t* in, uint32_t* out)
{
    auto v = ans_load_interp(in);
    if (out)
        std::copy(v.begin(), v.end(), out);
    return (long)v.size();
}

long ref_interp_encode(uint32_t* dst, const uint32_t* in, size_t n, size_t u)
{
    return (long)interpolative_internal::encode(dst, in, n, u);
}
long ref_interp_decode(const uint32_t* src, uint32_t* out, size_t n, size_t u)
{
    return (long)interpolative_internal::decode(src, out, n, u);
}

// ---- mappings (known-answer material of src/test.cpp:49-71) ----------------------------
int ref_fold_mapping(uint32_t f, uint32_t x, uint32_t* mapped)
{
    FOLD_DISPATCH(f, *mapped = ans_fold_mapping<F>(x); return 0);
    return -1;
}
int ref_fold_undo(uint32_t f, uint32_t s, uint32_t* value, uint32_t* bytes)
{
    FOLD_DISPATCH(f, *value = ans_fold_undo_mapping<F>(s); *bytes = ans_fold_exception_bytes<F>(s); return 0);
    return -1;
}
uint32_t ref_msb_mapping(uint32_t x) { return ans_msb_mapping(x); }
void ref_msb_undo(uint32_t s, uint32_t* value, uint32_t* bytes)
{
    *value = ans_msb_undo_mapping(s);
    *bytes = ans_msb_exception_bytes(s);
}

// ---- models + per-block oracle -----------------------------------------------------------
ref_model* ref_model_create(int method, uint32_t f, const uint32_t* in, size_t n)
{
    auto* m = new ref_model { method, f, nullptr, nullptr };
    switch (method) {
    case M_INT: m->enc = new ans_int_encode(ans_int_encode::create(in, n)); return m;
    case M_MSB: m->enc = new ans_msb_encode(ans_msb_encode::create(in, n)); return m;
    case M_BYTE: {
        const std::vector<uint8_t> b = to_bytes(in, n);
        m->enc = new ans_byte_encode(ans_byte_encode::create(b.data(), n));
        return m;
    }
    case M_FOLD:
        switch (f) {
#define C(F) case F: m->enc = new ans_fold_encode<F>(ans_fold_encode<F>::create(in, n)); return m;
            C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
        }
        break;
    case M_RFOLD:
        switch (f) {
#define C(F) case F: m->enc = new ans_reorder_fold_encode<F>(ans_reorder_fold_encode<F>::create(in, n)); return m;
            C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
        }
        break;
    }
    delete m;
    return nullptr;
}

ref_model* ref_model_from_hist(int method, uint32_t f, const uint64_t* bins, size_t nbins)
{
    auto* m = new ref_model { method, f, nullptr, nullptr };
    switch (method) {
    case M_INT: m->enc = enc_from_hist<ans_int_encode>(bins, nbins, false, true); return m;
    case M_MSB: m->enc = enc_from_hist<ans_msb_encode>(bins, nbins, true, false); return m;
    case M_FOLD:
        switch (f) {
#define C(F) case F: m->enc = enc_from_hist<ans_fold_encode<F>>(bins, nbins, true, false); return m;
            C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
        }
        break;
    }
    delete m;
    return nullptr;
}

#define WITH_ENC(m, EXPR)                                                                      \
    switch ((m)->method) {                                                                     \
    case M_INT: { auto& e = *static_cast<ans_int_encode*>((m)->enc); EXPR; } break;            \
    case M_MSB: { auto& e = *static_cast<ans_msb_encode*>((m)->enc); EXPR; } break;            \
    case M_BYTE: { auto& e = *static_cast<ans_byte_encode*>((m)->enc); EXPR; } break;          \
    case M_FOLD: FOLD_DISPATCH((m)->f, auto& e = *static_cast<ans_fold_encode<F>*>((m)->enc); EXPR); break; \
    case M_RFOLD: FOLD_DISPATCH((m)->f, auto& e = *static_cast<ans_reorder_fold_encode<F>*>((m)->enc); EXPR); break; \
    }

long ref_model_info(ref_model* m, uint64_t* frame_size, uint32_t* nfreqs_out)
{
    WITH_ENC(m, *frame_size = e.frame_size; if (nfreqs_out) std::copy(e.nfreqs.begin(), e.nfreqs.end(), nfreqs_out);
             return (long)e.nfreqs.size());
    return -1;
}

long ref_model_serialize(ref_model* m, uint8_t* out)
{
    uint8_t* p = out;
    WITH_ENC(m, e.serialize(p); return p - out);
    return -1;
}

// rfold only: most_frequent list (empty when sigma < 2^(f+7)) and sigma
long ref_model_rfold_info(ref_model* m, uint64_t* sigma, uint32_t* most_frequent_out)
{
    if (m->method != M_RFOLD)
        return -1;
    FOLD_DISPATCH(m->f, auto& e = *static_cast<ans_reorder_fold_encode<F>*>(m->enc); *sigma = e.sigma;
                  if (most_frequent_out) std::copy(e.most_frequent.begin(), e.most_frequent.end(), most_frequent_out);
                  return (long)e.most_frequent.size());
    return -1;
}

long ref_encode_block(ref_model* m, const uint32_t* in, size_t n, uint8_t* dst)
{
    WITH_ENC(m, return (long)encode_range(e, in, n, dst));
    return -1;
}

// decoder side: load() from a reference-format prelude
ref_model* ref_model_load(int method, uint32_t f, const uint8_t* prelude)
{
    auto* m = new ref_model { method, f, nullptr, nullptr };
    switch (method) {
    case M_INT: m->dec = new ans_int_decode(ans_int_decode::load(prelude)); return m;
    case M_MSB: m->dec = new ans_msb_decode(ans_msb_decode::load(prelude)); return m;
    case M_BYTE: m->dec = new ans_byte_decode(ans_byte_decode::load(prelude)); return m;
    case M_FOLD:
        switch (f) {
#define C(F) case F: m->dec = new ans_fold_decode<F>(ans_fold_decode<F>::load(prelude)); return m;
            C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
        }
        break;
    case M_RFOLD:
        switch (f) {
#define C(F) case F: m->dec = new ans_reorder_fold_decode<F>(ans_reorder_fold_decode<F>::load(prelude)); return m;
            C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
        }
        break;
    }
    delete m;
    return nullptr;
}

int ref_decode_block(ref_model* m, const uint8_t* src, size_t bytes, uint32_t* out, size_t n)
{
    switch (m->method) {
    case M_INT: {
        auto& d = *static_cast<ans_int_decode*>(m->dec);
        if (d.table_type == dec_table_type::SMALL)
            decode_range(d, src, bytes, out, n, [](ans_int_decode& mm, uint64_t& s, const uint8_t*& in) {
                return mm.decode_sym<dec_entry_int_small>(s, in);
            });
        else
            decode_range(d, src, bytes, out, n, [](ans_int_decode& mm, uint64_t& s, const uint8_t*& in) {
                return mm.decode_sym<dec_entry_int>(s, in);
            });
        return 0;
    }
    case M_MSB: {
        auto& d = *static_cast<ans_msb_decode*>(m->dec);
        decode_range(d, src, bytes, out, n,
            [](ans_msb_decode& mm, uint64_t& s, const uint8_t*& in) { return mm.decode_sym(s, in); });
        return 0;
    }
    case M_BYTE: {
        auto& d = *static_cast<ans_byte_decode*>(m->dec);
        decode_range(d, src, bytes, out, n,
            [](ans_byte_decode& mm, uint64_t& s, const uint8_t*& in) { return uint32_t(mm.decode_sym(s, in)); });
        return 0;
    }
    case M_FOLD:
        FOLD_DISPATCH(m->f, auto& d = *static_cast<ans_fold_decode<F>*>(m->dec);
                      decode_range(d, src, bytes, out, n,
                          [](ans_fold_decode<F>& mm, uint64_t& s, const uint8_t*& in) { return mm.decode_sym(s, in); });
                      return 0);
        break;
    case M_RFOLD:
        FOLD_DISPATCH(m->f, auto& d = *static_cast<ans_reorder_fold_decode<F>*>(m->dec);
                      decode_range(d, src, bytes, out, n,
                          [](ans_reorder_fold_decode<F>& mm, uint64_t& s, const uint8_t*& in) { return mm.decode_sym(s, in); });
                      return 0);
        break;
    }
    return -1;
}

void ref_model_free(ref_model* m)
{
    if (!m)
        return;
    if (m->enc) {
        switch (m->method) {
        case M_INT: del<ans_int_encode>(m->enc); break;
        case M_MSB: del<ans_msb_encode>(m->enc); break;
        case M_BYTE: del<ans_byte_encode>(m->enc); break;
        case M_FOLD:
            switch (m->f) {
#define C(F) case F: del<ans_fold_encode<F>>(m->enc); break;
                C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
            }
            break;
        case M_RFOLD:
            switch (m->f) {
#define C(F) case F: del<ans_reorder_fold_encode<F>>(m->enc); break;
                C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
            }
            break;
        }
    }
    if (m->dec) {
        switch (m->method) {
        case M_INT: del<ans_int_decode>(m->dec); break;
        case M_MSB: del<ans_msb_decode>(m->dec); break;
        case M_BYTE: del<ans_byte_decode>(m->dec); break;
        case M_FOLD:
            switch (m->f) {
#define C(F) case F: del<ans_fold_decode<F>>(m->dec); break;
                C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
            }
            break;
        case M_RFOLD:
            switch (m->f) {
#define C(F) case F: del<ans_reorder_fold_decode<F>>(m->dec); break;
                C(1) C(2) C(3) C(4) C(5) C(6) C(7) C(8)
#undef C
            }
            break;
        }
    }
    delete m;
}

// ---- input generators of src/generate_inputs.cpp:70-117 (mt19937 + libstdc++ / zipf_dist.hpp)
// kind: 0 = zipf(range, q) ; 1 = geometric(q) ; 2 = uniform [0, range) ; `sub` is subtracted
// from every value (SURVEY.md C5 uses zipf(2^20, 1.2) - 1).
int ref_generate(int kind, uint32_t seed, uint32_t range, double q, uint32_t sub, uint32_t* out, size_t n)
{
    std::mt19937 gen(seed);
    if (kind == 0) {
        zipf_distribution<uint32_t> d(range, q);
        for (size_t i = 0; i < n; i++)
            out[i] = d(gen) - sub;
    } else if (kind == 1) {
        std::geometric_distribution<uint32_t> d(q);
        for (size_t i = 0; i < n; i++)
            out[i] = d(gen) - sub;
    } else if (kind == 2) {
        std::uniform_int_distribution<uint32_t> d(0, range - 1);
        for (size_t i = 0; i < n; i++)
            out[i] = d(gen) - sub;
    } else
        return -1;
    return 0;
}

// table_efficiency.cpp:77-101 style timing: min of `runs`, ns. Round trip is verified.
// ---- one block of the pseudo-adaptive driver (src/pseudo_adaptive.cpp:83-125; the driver itself needs Boost) ------
// The driver's steps on one block, with the reference's own interpolative coder and T::encode (= ans_X_compress,
// methods.hpp:484-512) doing the coding: presence flags over [0, max], running count = 1-based rank, the list of
// present values turned into uint32_t running sums, [sigma][last + 1][interp], then the coder on the ranks.
long ref_pa_remap_block(int method, uint32_t f, const uint32_t* in, size_t n, uint8_t* dst, size_t cap,
    uint32_t* remapped_out, size_t* alphabet_bytes)
{
    if (n == 0 || !in || !dst)
        return -1;
    const uint32_t top = *std::max_element(in, in + n);
    std::vector<uint32_t> rank_of(size_t(top) + 1, 0);
    for (size_t k = 0; k < n; k++)
        rank_of[in[k]] = 1;
    std::vector<uint32_t> alphabet;
    for (size_t v = 0; v <= top; v++) {
        if (rank_of[v] == 1)
            alphabet.push_back(uint32_t(v));
        if (v)
            rank_of[v] += rank_of[v - 1];
    }
    std::vector<uint32_t> ranks(n);
    for (size_t k = 0; k < n; k++)
        ranks[k] = rank_of[in[k]];
    uint64_t run = 0;
    for (auto& a : alphabet) {
        run += a;
        if (run > 0xFFFFFFFEull)
            return -2; // the driver's uint32_t sums would wrap: outside its working domain
        a = uint32_t(run);
    }
    if (cap < 8 + 8 * alphabet.size() + 64)
        return -3;
    uint32_t* out32 = reinterpret_cast<uint32_t*>(dst);
    *out32++ = uint32_t(alphabet.size());
    *out32++ = alphabet.back() + 1;
    const size_t ib = interpolative_internal::encode(out32, alphabet.data(), alphabet.size(), alphabet.back() + 1);
    size_t at = 8 + ib;
    if (alphabet_bytes)
        *alphabet_bytes = at;
    if (alphabet.size() != 1) {
        const long eb = ref_compress(method, f, dst + at, cap - at, ranks.data(), n);
        if (eb < 0)
            return -1;
        at += size_t(eb);
    }
    if (remapped_out)
        std::copy(ranks.begin(), ranks.end(), remapped_out);
    return long(at);
}

int ref_time(int method, uint32_t f, const uint32_t* in, size_t n, int runs, double* enc_ns, double* dec_ns,
    size_t* bytes)
{
    std::vector<uint8_t> out(n * 8 + 4096);
    std::vector<uint32_t> rec(n + 16);
    double be = 1e300, bd = 1e300;
    long nb = 0;
    for (int r = 0; r < runs; r++) {
        auto t0 = std::chrono::high_resolution_clock::now();
        nb = ref_compress(method, f, out.data(), out.size(), in, n);
        auto t1 = std::chrono::high_resolution_clock::now();
        if (nb < 0)
            return -1;
        be = std::min(be, (double)std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count());
    }
    for (int r = 0; r < runs; r++) {
        auto t0 = std::chrono::high_resolution_clock::now();
        ref_decompress(method, f, rec.data(), n, out.data(), nb);
        auto t1 = std::chrono::high_resolution_clock::now();
        bd = std::min(bd, (double)std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count());
    }
    if (std::memcmp(rec.data(), in, n * sizeof(uint32_t)) != 0)
        return -2;
    *enc_ns = be;
    *dec_ns = bd;
    *bytes = (size_t)nb;
    return 0;
}

// the loop of pseudo_adaptive.cpp:76-141 around T::encode (values coded as they are, no alphabet remap), plus the
// matching decode: every `block` symbols are compressed on their own.  min of `runs`, ns, round trip verified.
int ref_time_blocks(int method, uint32_t f, const uint32_t* in, size_t n, size_t block, int runs, double* enc_ns,
    double* dec_ns, size_t* bytes)
{
    if (block == 0)
        return -1;
    const size_t nb = (n + block - 1) / block;
    std::vector<uint8_t> out(n * 8 + nb * 4096 + 4096);
    std::vector<size_t> at(nb + 1, 0);
    std::vector<uint32_t> rec(n + 16);
    double be = 1e300, bd = 1e300;
    for (int r = 0; r < runs; r++) {
        auto t0 = std::chrono::high_resolution_clock::now();
        size_t pos = 0;
        for (size_t b = 0; b < nb; b++) {
            const size_t len = std::min(block, n - b * block);
            const long w = ref_compress(method, f, out.data() + pos, out.size() - pos, in + b * block, len);
            if (w < 0)
                return -1;
            at[b] = pos;
            pos += (size_t)w;
        }
        at[nb] = pos;
        auto t1 = std::chrono::high_resolution_clock::now();
        be = std::min(be, (double)std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count());
    }
    for (int r = 0; r < runs; r++) {
        auto t0 = std::chrono::high_resolution_clock::now();
        for (size_t b = 0; b < nb; b++) {
            const size_t len = std::min(block, n - b * block);
            ref_decompress(method, f, rec.data() + b * block, len, out.data() + at[b], at[b + 1] - at[b]);
        }
        auto t1 = std::chrono::high_resolution_clock::now();
        bd = std::min(bd, (double)std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count());
    }
    if (std::memcmp(rec.data(), in, n * sizeof(uint32_t)) != 0)
        return -2;
    *enc_ns = be;
    *dec_ns = bd;
    *bytes = at[nb];
    return 0;
}

} // extern "C"
