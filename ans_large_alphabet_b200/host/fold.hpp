// b200ans: value -> (folded symbol, number of raw low bytes) and back, for host and device.
//
// Branch-free restatements of the reference's mapping rules (results identical for every
// 32-bit input; tests/test_mapping.py sweeps them against the reference):
//   fold_map        <- /root/reference/include/ans_fold.hpp:38-65 (ans_fold_mapping[_and_exceptions]<f>),
//                      ans_reorder_fold.hpp:36-64 (same rule applied to the reordered value)
//   fold_unmap      <- ans_fold.hpp:150-175 (ans_fold_undo_mapping<f>, ans_fold_exception_bytes<f>)
//   msb_map         <- ans_msb.hpp:41-73
//   msb_unmap       <- ans_msb.hpp:156-176
// Radix is 8 bits throughout (ans_fold.hpp:40).  `f` is the fidelity, 1..8.
#pragma once

#include <cstdint>

#if defined(__CUDACC__)
#define B200ANS_HD __host__ __device__ __forceinline__
#else
#define B200ANS_HD inline
#endif

namespace b200ans {

enum method_t : int { METHOD_INT = 0, METHOD_FOLD = 1, METHOD_MSB = 2, METHOD_RFOLD = 3, METHOD_BYTE = 4 };
// METHOD_BYTE (ans_byte.hpp): 256 symbols, its own normaliser and prelude; the coder is plain ANS's (no raw bytes)
B200ANS_HD int coder_of(int method) { return method == METHOD_BYTE ? METHOD_INT : method; }

constexpr uint32_t kByteSigma = 256;        // ans_byte.hpp:25  constants::MAX_SIGMA
constexpr uint32_t kByteMaxFrame = 4096;    // ans_byte.hpp:26  constants::MAX_FRAME_SIZE
constexpr uint32_t kByteFrameFactor = 64;   // ans_byte.hpp:27  constants::FRAME_FACTOR

constexpr uint32_t kAnsK = 16;         // ans_byte.hpp:30  constants::K
constexpr uint32_t kRadixLog2 = 32;    // ans_byte.hpp:28  constants::RADIX_LOG2
constexpr uint32_t kMsbMaxSigma = 1280; // ans_msb.hpp:29

B200ANS_HD uint32_t bit_length(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return 32u - uint32_t(__clz(int(x)));
#else
    return x ? 32u - uint32_t(__builtin_clz(x)) : 0u;
#endif
}

// Number of alphabet bins the folded histogram has: 2^(f+9) (ans_fold.hpp:70)
B200ANS_HD uint32_t fold_alphabet(uint32_t f) { return 1u << (f + 9); }

// x keeps its top (f+7) bits' worth of magnitude: every 8 bits above that become one raw byte
// and move the symbol up by 255 * 2^(f-1).
B200ANS_HD uint32_t fold_map(uint32_t x, uint32_t f, uint32_t& nbytes)
{
    const int excess = int(bit_length(x)) - int(f + 7);
    const uint32_t k = excess > 0 ? uint32_t(excess + 7) >> 3 : 0u;
    nbytes = k;
    return (x >> (8 * k)) + k * (255u << (f - 1));
}

B200ANS_HD uint32_t fold_unmap(uint32_t sym, uint32_t f, uint32_t& nbytes)
{
    const uint32_t thres = 1u << (f + 7);
    const uint32_t step = 255u << (f - 1);
    if (sym < thres) {
        nbytes = 0;
        return sym;
    }
    const uint32_t k = (sym - thres) / step + 1;
    nbytes = k;
    return (sym - step * k) << (8 * k);
}

B200ANS_HD uint32_t msb_map(uint32_t x, uint32_t& nbytes)
{
    const uint32_t k = uint32_t(x > 256u) + uint32_t(x > 65536u) + uint32_t(x > 16777216u);
    nbytes = k;
    return (x >> (8 * k)) + 256u * k;
}

B200ANS_HD uint32_t msb_unmap(uint32_t sym, uint32_t& nbytes)
{
    const uint32_t k = uint32_t(sym > 256u) + uint32_t(sym > 512u) + uint32_t(sym > 768u);
    nbytes = k;
    return (sym - 256u * k) << (8 * k);
}

} // namespace b200ans
