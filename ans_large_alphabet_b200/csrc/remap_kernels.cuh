// b200ans device code (sm_100a): the pseudo-adaptive driver's dense-rank alphabet remap (src/pseudo_adaptive.cpp:83-125).
//
// Per block the reference marks the values that occur in an array over the whole value range, turns the marks into
// running ranks (the smallest value present becomes 1), codes the block as those ranks with T::encode, and puts in front
//     [u32 sigma][u32 universe = last running sum + 1][interpolative code of the RUNNING SUMS of the sorted values]
// (its block_alphabet, summed in place, :100-112).  Here:
//   remap_block_kernel   one CTA per block: presence bitmap of the value range in shared memory (atomicOr), popcount
//                        prefix per 32-bit word, rank of a value = prefix + popc(bits at or below it); writes the ranks,
//                        the sorted distinct values and sigma
//   remap_alpha_kernel   one warp per block: running sums (a sum that would wrap 32 bits is outside the driver's working
//                        domain: error), then the alphabet header + interpolative code, bit-identical to interp.hpp
//   remap_alpha_decode_kernel / unremap_kernel: the way back (bounded reads, validated)
// The ranks then go through the per-block coders of the pseudo-adaptive mode (kernels.cuh / pa_kernels.cuh) as they are.
#pragma once

#include "kernels.cuh"

namespace b200ans {

constexpr int kRemapThreads = 1024;
constexpr uint32_t kRemapMaxValue = (1u << 20) + (1u << 17); // bitmap + prefixes of the value range must fit one SM's shared memory

// shared memory of remap_block_kernel for values 0..max_value
__host__ __device__ inline size_t remap_smem_bytes(uint32_t max_value)
{
    const size_t words = (size_t(max_value) >> 5) + 1;       // presence bits
    const size_t groups = (words + 31) / 32;                 // 1024 values each
    return words * 4 + groups * 4 + words * 2 + 64;          // bits, group prefix (u32), word prefix inside its group (u16)
}

__global__ void __launch_bounds__(kRemapThreads) remap_block_kernel(const uint32_t* __restrict__ in, uint64_t n, uint32_t model_syms,
    uint32_t n_models, uint32_t max_value, uint32_t* __restrict__ ranks, uint32_t* __restrict__ alphabet /* model_syms per block */,
    uint32_t* __restrict__ sigma)
{
    extern __shared__ __align__(16) uint8_t remap_smem[];
    const uint32_t words = (max_value >> 5) + 1u;
    const uint32_t groups = (words + 31u) / 32u;
    uint32_t* bits = reinterpret_cast<uint32_t*>(remap_smem);
    uint32_t* gpre = bits + words;
    uint16_t* wpre = reinterpret_cast<uint16_t*>(gpre + groups);
    __shared__ uint32_t s_warp[kRemapThreads / 32];
    __shared__ uint32_t s_carry;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, n_warps = kRemapThreads / 32;
    for (uint32_t w = threadIdx.x; w < words; w += blockDim.x) // once: every block clears the bits it set
        bits[w] = 0u;
    __syncthreads();
    for (uint32_t b = blockIdx.x; b < n_models; b += gridDim.x) {
        const uint64_t first = uint64_t(b) * model_syms;
        const uint32_t len = uint32_t(min(uint64_t(model_syms), n - first));
        if (threadIdx.x == 0)
            s_carry = 0u;
        // (blocks start 16-byte aligned: model_syms is a multiple of 128; the last block may end anywhere)
        const uint4* in4 = reinterpret_cast<const uint4*>(in + first);
        const uint32_t len4 = len >> 2;
        for (uint32_t i = threadIdx.x; i < len4; i += blockDim.x) {
            const uint4 q = __ldg(in4 + i);
            const uint32_t v4[4] = { min(q.x, max_value), min(q.y, max_value), min(q.z, max_value), min(q.w, max_value) }; // (the caller checked the maximum)
#pragma unroll
            for (int j = 0; j < 4; j++)
                atomicOr(bits + (v4[j] >> 5), 1u << (v4[j] & 31u));
        }
        if (threadIdx.x < (len & 3u)) {
            const uint32_t v = min(__ldg(in + first + (len4 << 2) + threadIdx.x), max_value);
            atomicOr(bits + (v >> 5), 1u << (v & 31u));
        }
        __syncthreads();
        // popcount prefix: inside each group of 32 words by a warp scan, then over the groups
        for (uint32_t g = warp; g < groups; g += n_warps) {
            const uint32_t w = g * 32u + lane;
            const uint32_t c = w < words ? uint32_t(__popc(bits[w])) : 0u;
            uint32_t incl = c;
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= uint32_t(o))
                    incl += up;
            }
            if (w < words)
                wpre[w] = uint16_t(incl - c);
            if (lane == 31u)
                gpre[g] = incl; // the group's total for now
        }
        __syncthreads();
        for (uint32_t base = 0; base < groups; base += kRemapThreads) { // exclusive scan of the group totals
            const uint32_t g = base + threadIdx.x;
            const uint32_t c = g < groups ? gpre[g] : 0u;
            uint32_t incl = c;
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= uint32_t(o))
                    incl += up;
            }
            if (lane == 31u)
                s_warp[warp] = incl;
            __syncthreads();
            uint32_t before = s_carry;
            for (uint32_t w = 0; w < warp; w++)
                before += s_warp[w];
            if (g < groups)
                gpre[g] = before + incl - c;
            __syncthreads();
            if (threadIdx.x == kRemapThreads - 1)
                s_carry = before + incl;
            __syncthreads();
        }
        if (threadIdx.x == 0)
            sigma[b] = s_carry;
        // ranks: number of present values at or below v (pseudo_adaptive.cpp:93-99: the marks summed in place); the
        // value of rank r is the block alphabet's entry r - 1 (every occurrence of a value writes the same entry)
        uint32_t* out = alphabet + uint64_t(b) * model_syms;
        auto rank_of = [&](uint32_t v) {
            const uint32_t w = v >> 5;
            const uint32_t r = gpre[w >> 5] + wpre[w] + uint32_t(__popc(bits[w] & (0xFFFFFFFFu >> (31u - (v & 31u)))));
            out[r - 1u] = v;
            return r;
        };
        uint4* ranks4 = reinterpret_cast<uint4*>(ranks + first);
        for (uint32_t i = threadIdx.x; i < len4; i += blockDim.x) {
            const uint4 q = __ldg(in4 + i);
            ranks4[i] = make_uint4(rank_of(min(q.x, max_value)), rank_of(min(q.y, max_value)), rank_of(min(q.z, max_value)), rank_of(min(q.w, max_value)));
        }
        if (threadIdx.x < (len & 3u)) {
            const uint32_t at = (len4 << 2) + threadIdx.x;
            ranks[first + at] = rank_of(min(__ldg(in + first + at), max_value));
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < len4; i += blockDim.x) { // leave the bitmap empty for the next block
            const uint4 q = __ldg(in4 + i);
            bits[min(q.x, max_value) >> 5] = 0u, bits[min(q.y, max_value) >> 5] = 0u;
            bits[min(q.z, max_value) >> 5] = 0u, bits[min(q.w, max_value) >> 5] = 0u;
        }
        if (threadIdx.x < (len & 3u))
            bits[min(__ldg(in + first + (len4 << 2) + threadIdx.x), max_value) >> 5] = 0u;
        __syncthreads();
    }
}

constexpr int kRemapAlphaThreads = 32;
constexpr int kRemapEncThreads = 256; // 8 warps = 8 blocks per CTA

// codeword of the centred minimal binary code as one LSB-first bit string (host/prelude.hpp put_centered)
__device__ __forceinline__ void remap_centered_code(uint64_t val, uint64_t u, uint64_t& code, uint32_t& nbits)
{
    code = 0;
    nbits = 0;
    if (u == 1)
        return;
    const uint32_t b = 64u - uint32_t(__clzll((long long)(u - 1u)));
    const uint64_t span = uint64_t(1) << b;
    const uint64_t longs = 2u * u - span;
    val += u - (longs >> 1);
    if (val > u)
        val -= u;
    const uint64_t shorts = span - u;
    if (val <= shorts) {
        code = val - 1u;
        nbits = b - 1u;
    } else {
        val += shorts;
        code = ((val - 1u) >> 1) | (((val - 1u) & 1u) << (b - 1u));
        nbits = b;
    }
}

// alphabet[b][0 .. sigma[b]) (sorted distinct values; summed in place here)  ->  [sigma][universe][interp] at
// alpha + b * alpha_stride, alpha_bytes[b].  One WARP per block: the recursion of the interpolative code is
// data-independent in shape -- the node that is coded p-th (pre-order) and the segment it splits follow from p and sigma
// alone, and its bounds are the list entries next to that segment -- so the lanes code 32 consecutive nodes at a time,
// a warp scan of the code lengths gives the bit positions, and the words are assembled in shared memory.
__global__ void __launch_bounds__(kRemapEncThreads) remap_alpha_kernel(uint32_t* __restrict__ alphabet, const uint32_t* __restrict__ sigma,
    uint32_t model_syms, uint32_t n_models, uint8_t* __restrict__ alpha, uint64_t alpha_stride, uint32_t* __restrict__ alpha_bytes,
    uint32_t* __restrict__ error)
{
    __shared__ uint32_t s_buf[kRemapEncThreads / 32][72];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t b = blockIdx.x * (kRemapEncThreads / 32) + warp;
    if (b >= n_models)
        return;
    uint32_t* v = alphabet + uint64_t(b) * model_syms;
    const uint32_t cnt_all = sigma[b];
    // running sums in place (pseudo_adaptive.cpp:101-103); a sum that would wrap 32 bits is outside the driver's domain
    uint64_t carry = 0;
    for (uint32_t base = 0; base < cnt_all; base += 32u) {
        const uint32_t j = base + lane;
        uint64_t x = j < cnt_all ? v[j] : 0u;
        for (int o = 1; o < 32; o <<= 1) {
            const uint64_t up = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= uint32_t(o))
                x += up;
        }
        x += carry;
        if (j < cnt_all)
            v[j] = uint32_t(min(x, uint64_t(0xFFFFFFFEull)));
        carry = __shfl_sync(0xffffffffu, x, 31);
    }
    if (carry > 0xFFFFFFFEull) {
        if (lane == 0)
            *error = 3u;
        carry = 0xFFFFFFFEull;
    }
    __syncwarp();
    const uint32_t universe = uint32_t(carry) + 1u;
    uint32_t* out32 = reinterpret_cast<uint32_t*>(alpha + uint64_t(b) * alpha_stride); // 16-byte aligned
    if (lane == 0) {
        out32[0] = cnt_all; // :105-107
        out32[1] = universe;
    }
    uint32_t* buf = s_buf[warp];
    uint64_t bitpos = 0; // bits written so far (behind the two header words)
    buf[lane] = 0u, buf[lane + 32u] = 0u;
    if (lane < 8u)
        buf[64u + lane] = 0u;
    __syncwarp();
    for (uint32_t base = 0; base < cnt_all; base += 32u) {
        const uint32_t p = base + lane;
        uint64_t code = 0;
        uint32_t nbits = 0;
        if (p < cnt_all) {
            uint32_t off = 0, cnt = cnt_all, q = p, h, left, right;
            for (;;) { // interp_put's recursion (interp.hpp:28-62), walked by position
                h = (cnt + 1u) >> 1, left = h - 1u, right = cnt - h;
                if (q == 0u)
                    break;
                if (q <= left) {
                    cnt = left;
                    q -= 1u;
                } else {
                    off += h;
                    cnt = right;
                    q -= 1u + left;
                }
            }
            const uint64_t mid = uint64_t(v[off + h - 1u]) + 1u;
            const uint64_t low = off == 0u ? 1u : uint64_t(v[off - 1u]) + 2u;
            const uint64_t high = off + cnt == cnt_all ? uint64_t(universe) + 1u : uint64_t(v[off + cnt]);
            remap_centered_code(mid - low - left + 1u, high - right - low - left + 1u, code, nbits);
        }
        uint32_t incl = nbits;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= uint32_t(o))
                incl += up;
        }
        const uint32_t chunk_bits = __shfl_sync(0xffffffffu, incl, 31);
        const uint64_t word0 = bitpos >> 5; // buf[0] is this word (partly filled)
        if (nbits) {
            const uint64_t at = bitpos + (incl - nbits);
            const uint32_t wi = uint32_t((at >> 5) - word0), sh = uint32_t(at & 31u);
            const uint64_t shifted = code << sh; // at most 32 bits shifted by at most 31: two words
            atomicOr(buf + wi, uint32_t(shifted));
            if (shifted >> 32)
                atomicOr(buf + wi + 1u, uint32_t(shifted >> 32));
        }
        __syncwarp();
        bitpos += chunk_bits;
        const uint32_t full = uint32_t((bitpos >> 5) - word0); // whole words ready (<= 33)
        for (uint32_t i = lane; i < full; i += 32u)
            out32[2u + word0 + i] = buf[i];
        __syncwarp();
        const uint32_t keep = buf[full]; // the partly filled word moves to the front
        __syncwarp();
        buf[lane] = 0u, buf[lane + 32u] = 0u;
        if (lane < 8u)
            buf[64u + lane] = 0u;
        __syncwarp();
        if (lane == 0)
            buf[0] = keep;
        __syncwarp();
    }
    uint32_t words = uint32_t(bitpos >> 5);
    if (bitpos & 31u) { // the last, partly filled word (bits.hpp: flushed as a whole word)
        if (lane == 0)
            out32[2u + words] = buf[0];
        words++;
    }
    const uint32_t bytes = 8u + 4u * words;
    if (lane == 0)
        alpha_bytes[b] = bytes;
    for (uint32_t z = bytes / 4u + lane; z < ((bytes + 15u) & ~15u) / 4u; z += 32u) // the compaction copies whole 16-byte chunks
        out32[z] = 0u;
}

// [sigma][universe][interp] of block b at src + offsets[b] (alpha_bytes[b] of them)  ->  the running sums the code carries, in
// alphabet[b][0 .. sigma), sigma[b] (0: not a valid alphabet).  One thread per block, the lockstep serial decoder of
// kernels.cuh (pa_interp_decode) on 64-bit bounds; remap_alpha_finish_kernel turns the sums back into values.
// (The first version -- a loop trip per node or per pop, the stack in local memory -- took 4.9 ms for 16 384 alphabets
// of ~5000 values, most of a remap decode.)
constexpr int kRemapDepth = 22; // pending right halves: log2(sigma) + 1, sigma <= 2^21
__global__ void __launch_bounds__(kRemapAlphaThreads) remap_alpha_decode_kernel(const uint8_t* __restrict__ src, const uint64_t* __restrict__ offsets,
    const uint32_t* __restrict__ alpha_bytes, uint64_t region_bytes, uint32_t model_syms, uint32_t n_models, uint32_t* __restrict__ alphabet,
    uint32_t* __restrict__ sigma)
{
    __shared__ uint32_t s_stack[kRemapDepth * 6 * kRemapAlphaThreads];
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_models)
        return;
    const bool inside = offsets[b] + alpha_bytes[b] <= region_bytes && (alpha_bytes[b] & 3u) == 0u && alpha_bytes[b] >= 8u; // the size table is untrusted
    const uint32_t* head = reinterpret_cast<const uint32_t*>(src + (inside ? offsets[b] : 0)); // (src and the offsets are 16-byte aligned)
    uint32_t cnt_all = inside ? __ldg(head) : 0u;
    const uint32_t universe = inside ? __ldg(head + 1) : 0u;
    bool bad = !inside || cnt_all == 0u || cnt_all > model_syms || cnt_all > (1u << (kRemapDepth - 1)) || universe == 0u;
    if (bad)
        cnt_all = 0u;
    PaWordReader r;
    r.start(src, (inside ? offsets[b] : 0) + 8u, region_bytes ? region_bytes - 1u : 0u);
    if (!pa_interp_decode<uint64_t>(r, cnt_all, uint64_t(1), uint64_t(universe) + 1u, ~uint64_t(0), alphabet + uint64_t(b) * model_syms,
            s_stack + threadIdx.x, kRemapAlphaThreads))
        bad = true;
    if (!bad && r.bits_used() > (uint64_t(alpha_bytes[b]) - 8u) * 8u)
        bad = true;
    sigma[b] = bad ? 0u : cnt_all;
}

// running sums back to the values, in place; they must climb.  One warp per block.
__global__ void __launch_bounds__(kRemapEncThreads) remap_alpha_finish_kernel(uint32_t* __restrict__ alphabet, uint32_t* __restrict__ sigma, uint32_t model_syms,
    uint32_t n_models, uint32_t* __restrict__ error)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t b = blockIdx.x * (kRemapEncThreads / 32) + (threadIdx.x >> 5);
    if (b >= n_models)
        return;
    uint32_t* v = alphabet + uint64_t(b) * model_syms;
    const uint32_t cnt_all = sigma[b];
    bool bad = cnt_all == 0u;
    uint32_t below = 0;
    for (uint32_t base = 0; base < cnt_all; base += 32u) {
        const uint32_t j = base + lane;
        const uint32_t here = j < cnt_all ? v[j] : 0xFFFFFFFFu;
        uint32_t prev = __shfl_up_sync(0xffffffffu, here, 1);
        if (lane == 0)
            prev = below;
        if (j < cnt_all) {
            if (j && here < prev)
                bad = true;
            v[j] = here - prev;
        }
        below = __shfl_sync(0xffffffffu, here, 31);
    }
    if (__any_sync(0xffffffffu, bad) && lane == 0) {
        sigma[b] = 0u;
        *error = 4u;
    }
}

// value = alphabet[block][rank - 1]
__global__ void __launch_bounds__(256) unremap_kernel(const uint32_t* __restrict__ ranks, uint64_t n, uint32_t model_syms,
    const uint32_t* __restrict__ alphabet, const uint32_t* __restrict__ sigma, uint32_t* __restrict__ out, uint32_t* __restrict__ error)
{
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x) {
        const uint64_t b = i / model_syms;
        const uint32_t r = ranks[i];
        if (r == 0u || r > sigma[b]) {
            *error = 5u;
            out[i] = 0u;
        } else {
            out[i] = __ldg(alphabet + b * model_syms + (r - 1u));
        }
    }
}

} // namespace b200ans
