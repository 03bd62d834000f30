// b200ans pseudo-adaptive coding kernels with the block's model in SHARED memory (sm_100a).
//
// Pseudo-adaptive mode (src/pseudo_adaptive.cpp:64-177) gives every block of `model_syms` symbols its own
// semi-static model; prelude(b) || payload(b) must be byte for byte what ans_X_compress() writes for block b, i.e.
// one reference-format stream with FOUR rANS chains per block.  The parallelism is therefore fixed -- 4 lanes per
// model, 8 models per warp, as in encode_kernel / decode_kernel of kernels.cuh -- and every step is one dependent
// chain  symbol -> table entry -> state update -> 4-lane scan of the bytes emitted/consumed.  The first generation
// kept the per-model tables (16 B per symbol / 8 B per frame slot) in global memory: 256 MB + 512 MB of tables for
// 16 384 models, far beyond L2, so every step waited for DRAM (ncu r01d: long-scoreboard stalls on the table load,
// decode kernel 65 % DRAM throughput for 4 % algorithmic roofline).
//
// Here each 4-lane group builds its model's tables in shared memory straight from the normalised frequencies
// (pa_nfreqs, u32 per bin, written by pa_fit_kernel resp. pa_prelude_decode_kernel) at the start of its task:
//   both:    cum[s] = slots before symbol s, u16, 2 B per symbol up to the largest symbol any model of the input
//            has: base = cum[s], freq = cum[s + 1] - cum[s];
//   encoder: the exact reciprocal the state update needs depends only on freq, so it comes from a per-CTA
//            table rcp[freq] for freq < kPaRcpN and is computed (IEEE divide + one fix-up) beyond that;
//   decoder: + a bucket index (symbol holding the first slot of every 2^bshift slots); slot -> symbol is a
//            bounded binary search between two bucket entries, and the unfolded value / raw byte count follow
//            arithmetically from the symbol (fold_unmap / msb_unmap), so no M-entry slot table exists at all
//            (ans_fold.hpp:179-207 builds one; the contents are a pure function of nfreqs, the layout is ours).
// The steps of a chain are serial and a warp issues in order, so the kernels are bound by latency x resident
// models (ncu, first version of these kernels: 6.3 cycles per issued instruction, 0.9 ms per round of tasks whatever the warp count); the
// tables are this small -- about 1.3 KB (encoder) / 1.9 KB (decoder) per model for a 527-symbol alphabet -- so that
// all 14 warps (112 models) an SM needs for 2^27 symbols in 8192-symbol blocks are resident at once.
// Both need log2 M <= 16 (cum fits 16 bits; frequencies are at most 65535 by the 16-bit guard of adjust_freqs,
// ans_util.hpp:139-146, so differences survive the wrap of cum = 65536); the host falls back to the global-table
// kernels otherwise.
#pragma once

#include "kernels.cuh"

namespace b200ans {

constexpr int kPaRcpN = 2048;    // reciprocals kept per CTA: frequencies below this never divide
constexpr int kPaCodeMaxWarps = 16;

// smallest double >= 1/f, f >= 1: trunc(x * rcp) == floor(x / f) for every x < 2^52 (kernels.cuh pa_reciprocal_up)
__device__ __forceinline__ double pa_rcp_up(uint32_t f)
{
    double r = 1.0 / double(f);
    if (fma(r, double(f), -1.0) < 0.0)
        r = __longlong_as_double(__double_as_longlong(r) + 1);
    return r;
}

// bytes of shared memory one 4-lane group needs
__host__ __device__ inline uint32_t pa_cum_bytes(uint32_t syms) { return ((syms + 1u) * 2u + 15u) & ~15u; }
__host__ __device__ inline uint32_t pa_enc_group_bytes(uint32_t syms) { return pa_cum_bytes(syms) + uint32_t(kEncRowBytes); }
__host__ __device__ inline uint32_t pa_dec_bucket_bytes(uint32_t bucket_log2) { return (((1u << bucket_log2) + 1u) * 2u + 15u) & ~15u; }
__host__ __device__ inline uint32_t pa_dec_group_bytes(uint32_t syms, uint32_t bucket_log2)
{
    return pa_cum_bytes(syms) + pa_dec_bucket_bytes(bucket_log2) + uint32_t(kDecRowBytes);
}

// The 4 lanes of a group build their model's prefix sums: cum[s] = nf[0] + .. + nf[s - 1] for s in [0, syms], syms a
// multiple of 16 (lane g takes a contiguous quarter).  BUCKETS: also bucket[b] = symbol that holds slot b << bshift
// for b < n_buckets, and bucket[n_buckets] = syms - 1 (upper bound of the search in the last bucket).
template <bool BUCKETS>
__device__ __forceinline__ void pa_build_cum(const uint32_t* __restrict__ nf_model, uint32_t syms, uint32_t g, uint16_t* cum,
    uint16_t* bucket, uint32_t bshift, uint32_t n_buckets)
{
    const uint32_t per = syms >> 2; // multiple of 4
    const uint4* nf = reinterpret_cast<const uint4*>(nf_model + g * per);
    uint32_t sum = 0;
    for (uint32_t i = 0; i < (per >> 2); i++) {
        const uint4 v = __ldg(nf + i);
        sum += v.x + v.y + v.z + v.w;
    }
    uint32_t incl = sum;
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, 1, 4);
    if (g >= 1u)
        incl += t;
    t = __shfl_up_sync(0xffffffffu, incl, 2, 4);
    if (g >= 2u)
        incl += t;
    uint32_t base = incl - sum;
    const uint32_t bround = (1u << bshift) - 1u;
    uint2* cv = reinterpret_cast<uint2*>(cum + g * per);
    for (uint32_t i = 0; i < (per >> 2); i++) {
        const uint4 v = __ldg(nf + i);
        const uint32_t fr[4] = { v.x, v.y, v.z, v.w };
        uint32_t c[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            c[j] = base;
            if (BUCKETS && fr[j]) {
                const uint32_t s = g * per + 4u * i + uint32_t(j);
                const uint32_t b1 = (base + fr[j] - 1u) >> bshift;
                for (uint32_t b = (base + bround) >> bshift; b <= b1; b++)
                    bucket[b] = uint16_t(s);
            }
            base += fr[j];
        }
        cv[i] = make_uint2((c[0] & 0xFFFFu) | (c[1] << 16), (c[2] & 0xFFFFu) | (c[3] << 16));
    }
    if (g == 3u) {
        cum[syms] = uint16_t(base); // = M (wraps to 0 for M = 65536: only differences are used)
        if (BUCKETS)
            bucket[n_buckets] = uint16_t(syms - 1u);
    }
}

// (pa_unmap: kernels.cuh)

// ------------------------------------------------------------------------------------------
// encoder: encode_kernel<METHOD, false, true> with the block's table in shared memory
// ------------------------------------------------------------------------------------------
template <int METHOD>
__global__ void __launch_bounds__(kPaCodeMaxWarps * 32) pa_smem_encode_kernel(const EncArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    double* s_rcp = reinterpret_cast<double*>(smem_raw);
    for (uint32_t i = threadIdx.x; i < uint32_t(kPaRcpN); i += blockDim.x)
        s_rcp[i] = i ? pa_rcp_up(i) : 0.0;
    __syncthreads();

    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps = blockDim.x >> 5;
    const uint32_t grp = lane >> 2;
    const uint32_t g = lane & 3u; // position inside a 4-symbol step; chain = 3 - g
    const uint32_t grp_base = lane & ~3u;
    const uint32_t syms = a.pa_syms; // multiple of 16
    uint8_t* gmem = smem_raw + size_t(kPaRcpN) * 8 + size_t(warp * 8u + grp) * pa_enc_group_bytes(syms);
    uint16_t* cum = reinterpret_cast<uint16_t*>(gmem);
    uint8_t* row = gmem + pa_cum_bytes(syms);
    const uint32_t f = a.fidelity;
    const uint32_t scan_mine = group_scan_mask(g, false);

    const uint32_t n_tasks = (a.n_blocks + 7u) >> 3;
    for (uint32_t task = blockIdx.x * warps + warp; task < n_tasks; task += gridDim.x * warps) {
        const uint32_t blk = task * 8 + grp;
        const bool have_block = blk < a.n_blocks;
        const uint32_t m = min(blk, a.n_blocks - 1u);
        const uint32_t log2_frame = a.pa_log2[m];

        // ---- the block's encoder table (ans_fold.hpp:84-91): prefix sums of nfreqs over the 4 lanes ----
        __syncwarp();
        pa_build_cum<false>(a.pa_nfreqs + uint64_t(m) * a.pa_nbins, syms, g, cum, nullptr, 0u, 0u);
        __syncwarp();

        const uint64_t lower = uint64_t(kAnsK) << log2_frame; // L = K * M (ans_fold.hpp:91)
        const uint64_t first = uint64_t(blk) * a.block_syms;
        const uint32_t nb = have_block ? uint32_t(min(uint64_t(a.block_syms), a.n - first)) : 0u;
        const uint32_t* src = a.in + first;
        const uint32_t tail = nb & 3u; // these go to chain 0 first (ans_fold.hpp:257-261)
        const uint32_t quads = nb >> 2;
        const uint32_t my_steps = quads + tail;
        const uint32_t max_steps = __reduce_max_sync(0xffffffffu, my_steps);

        uint64_t x = lower;   // initial_state() (ans_fold.hpp:113)
        bool bad = false;     // a symbol without a frequency in this block's model
        uint32_t fill = 0;    // bytes sitting in the row
        uint64_t flushed = 0; // bytes already drained to scratch
        uint8_t* dst = a.scratch + uint64_t(blk) * a.stride;

        auto fetch = [&](uint32_t s, uint32_t& v) -> bool {
            bool active;
            uint32_t idx;
            if (s < tail) {
                active = (g == 3u);
                idx = nb - 1u - s;
            } else {
                active = s < my_steps;
                idx = 4u * (quads - 1u - (s - tail)) + g;
            }
            v = active ? __ldg(src + idx) : 0u;
            return active;
        };

        uint32_t vnext[kEncTile];
        uint32_t anext = 0;
#pragma unroll
        for (int i = 0; i < kEncTile; i++)
            anext |= uint32_t(fetch(uint32_t(i), vnext[i])) << i;

        for (uint32_t s0 = 0; s0 < max_steps; s0 += kEncTile) {
            uint32_t vcur[kEncTile];
            const uint32_t acur = anext;
#pragma unroll
            for (int i = 0; i < kEncTile; i++)
                vcur[i] = vnext[i];
            anext = 0;
            if (s0 + kEncTile < max_steps) {
#pragma unroll
                for (int i = 0; i < kEncTile; i++)
                    anext |= uint32_t(fetch(s0 + kEncTile + uint32_t(i), vnext[i])) << i;
            }

            // Table entries one step ahead: the loads of step i + 1 are issued before the row stores of step i (which
            // the compiler must assume alias them), so their latency overlaps the scan and the state update.
            uint32_t k_n, freq_n, base_n;
            double rcp_n;
            auto lookup = [&](uint32_t v, bool active) {
                uint32_t sym = map_value<METHOD>(v, f, k_n);
                const bool outside = sym >= syms; // value beyond every model of this input: flag, stay in bounds
                sym = outside ? 0u : sym;
                base_n = cum[sym];
                freq_n = (uint32_t(cum[sym + 1u]) - base_n) & 0xFFFFu;
                rcp_n = freq_n < uint32_t(kPaRcpN) ? s_rcp[freq_n] : pa_rcp_up(freq_n);
                bad |= active && (outside || freq_n == 0u);
                k_n = active ? k_n : 0u;
            };
            lookup(vcur[0], acur & 1u);
#pragma unroll
            for (int i = 0; i < kEncTile; i++) {
                const bool active = (acur >> i) & 1u;
                const uint32_t v = vcur[i];
                const uint32_t k = k_n, freq = freq_n, base = base_n;
                const double rcp = rcp_n;
                if (i + 1 < kEncTile)
                    lookup(vcur[i + 1], (acur >> (i + 1)) & 1u);
                const bool renorm = active && (uint32_t(x >> 36) >= freq); // state >= freq * K * 2^32
                const uint32_t c = k + (renorm ? 4u : 0u);
                // reverse inclusive scan over the group: lanes with larger g emit first
                uint32_t incl, total;
                group_scan(k, renorm, grp_base, scan_mine, incl, total);
                const uint32_t w = uint32_t(x);
                if (renorm)
                    x >>= kRadixLog2;
                const uint64_t xn = ans_push(x, freq, base, rcp, log2_frame);
                uint8_t* p = row + fill + (incl - c);
                if (active) {
                    if (k > 0u)
                        p[0] = uint8_t(v);
                    if (k > 1u)
                        p[1] = uint8_t(v >> 8);
                    if (k > 2u)
                        p[2] = uint8_t(v >> 16);
                    if (renorm) {
                        p[k] = uint8_t(w);
                        p[k + 1] = uint8_t(w >> 8);
                        p[k + 2] = uint8_t(w >> 16);
                        p[k + 3] = uint8_t(w >> 24);
                    }
                    x = xn;
                }
                fill += total;
            }

            // drain whole 16-byte chunks, carry the remainder to the front of the row
            __syncwarp();
            {
                const uint32_t chunks = fill >> 4;
                const uint4* rv = reinterpret_cast<const uint4*>(row);
                uint4* gv = reinterpret_cast<uint4*>(dst + flushed);
                for (uint32_t ci = g; ci < chunks; ci += 4)
                    gv[ci] = rv[ci];
                uint4 carry = make_uint4(0, 0, 0, 0);
                if (g == 0u && chunks != 0u)
                    carry = rv[chunks];
                __syncwarp();
                if (g == 0u && chunks != 0u)
                    *reinterpret_cast<uint4*>(row) = carry;
                flushed += uint64_t(chunks) << 4;
                fill &= 15u;
            }
            __syncwarp();
        }

        // flush_state x4 in chain order 0..3 (ans_fold.hpp:115-120, 275-278), then the rest of the row
        if (have_block) {
            const uint64_t fin = x - lower;
            uint8_t* p = row + fill + 8u * (3u - g);
#pragma unroll
            for (int b = 0; b < 8; b++)
                p[b] = uint8_t(fin >> (8 * b));
        }
        fill += have_block ? 32u : 0u;
        for (uint32_t z = fill + g; z < ((fill + 15u) & ~15u); z += 4) // deterministic padding
            row[z] = 0;
        __syncwarp();
        {
            const uint32_t chunks = (fill + 15u) >> 4;
            const uint4* rv = reinterpret_cast<const uint4*>(row);
            uint4* gv = reinterpret_cast<uint4*>(dst + flushed);
            if (have_block)
                for (uint32_t ci = g; ci < chunks; ci += 4)
                    gv[ci] = rv[ci];
            if (have_block && g == 0u)
                a.block_bytes[blk] = uint32_t(flushed + fill);
        }
        if (bad)
            *a.error_flag = 1u;
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------
// decoder: decode_kernel<METHOD, DEC_FOLD_GLOBAL, true> without slot tables
// ------------------------------------------------------------------------------------------
template <int METHOD>
__global__ void __launch_bounds__(kPaCodeMaxWarps * 32) pa_smem_decode_kernel(const DecArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps = blockDim.x >> 5;
    const uint32_t grp = lane >> 2;
    const uint32_t g = lane & 3u;
    const uint32_t grp_base = lane & ~3u;
    const uint32_t scan_mine = group_scan_mask(g, true);
    const uint32_t syms = a.pa_syms; // multiple of 16, beyond the largest symbol of every model
    const uint32_t blog = a.pa_bucket_log2;
    uint8_t* gmem = smem_raw + size_t(warp * 8u + grp) * pa_dec_group_bytes(syms, blog);
    uint16_t* cum = reinterpret_cast<uint16_t*>(gmem);
    uint16_t* bucket = reinterpret_cast<uint16_t*>(gmem + pa_cum_bytes(syms));
    uint8_t* row = gmem + pa_cum_bytes(syms) + pa_dec_bucket_bytes(blog);
    const uint32_t f = a.fidelity;

    const uint32_t n_tasks = (a.n_blocks + 7u) >> 3;
    for (uint32_t task = blockIdx.x * warps + warp; task < n_tasks; task += gridDim.x * warps) {
        const uint32_t blk = task * 8 + grp;
        const bool have_block = blk < a.n_blocks;
        const uint32_t m = min(blk, a.n_blocks - 1u);
        const uint32_t log2_frame = a.pa_log2[m];
        const uint32_t bshift = log2_frame > blog ? log2_frame - blog : 0u;
        const uint32_t n_buckets = 1u << (log2_frame - bshift);

        // ---- cum[s] = slots before symbol s; bucket[b] = symbol that holds slot b << bshift ----
        __syncwarp();
        pa_build_cum<true>(a.pa_nfreqs + uint64_t(m) * a.pa_nbins, syms, g, cum, bucket, bshift, n_buckets);
        __syncwarp();

        const uint32_t frame_mask = (1u << log2_frame) - 1u;
        const uint64_t lower = uint64_t(kAnsK) << log2_frame;
        const uint64_t first = uint64_t(blk) * a.block_syms;
        const uint32_t nb = have_block ? uint32_t(min(uint64_t(a.block_syms), a.n - first)) : 0u;
        const uint32_t tail = nb & 3u;
        const uint32_t quads = nb >> 2;
        const uint32_t my_steps = quads + tail;
        const uint32_t max_steps = __reduce_max_sync(0xffffffffu, my_steps);
        uint32_t* out = a.out + first;

        // positions are relative to the 16-byte aligned address at or below the payload start
        const uint8_t* pay = a.payload + (have_block ? a.offsets[blk] : 0);
        const uint32_t delta = uint32_t(reinterpret_cast<uintptr_t>(pay) & 15u);
        const uint8_t* pay_al = pay - delta;
        const uint32_t size = have_block ? a.block_bytes[blk] : 32u;

        // init_state x4: chain 3 is the last 8 bytes ... chain 0 the first of the 32 (ans_fold.hpp:291-295)
        uint64_t x = lower;
        if (have_block) {
            const uint8_t* sp = pay + size - 32u + 8u * (3u - g);
            uint64_t v = 0;
#pragma unroll
            for (int b = 0; b < 8; b++)
                v |= uint64_t(sp[b]) << (8 * b);
            x = v + lower;
        }
        int32_t cursor = int32_t(delta + size - 32u); // one past the next byte to consume

        for (uint32_t s0 = 0; s0 < max_steps; s0 += kDecTile) {
            // stage [lo, lo + 288) around the cursor
            int32_t lo = cursor - (28 * kDecTile + 12);
            lo = lo < 0 ? 0 : (lo & ~15);
            __syncwarp();
            if (have_block) {
                const uint4* gsrc = reinterpret_cast<const uint4*>(pay_al + lo);
                uint4* sdst = reinterpret_cast<uint4*>(row + kDecPad);
                const int32_t need = (cursor + 4 - lo + 15) >> 4; // chunks that can be touched
#pragma unroll
                for (int ci = 0; ci < (kDecStageChunks + 3) / 4; ci++) {
                    const int32_t c = ci * 4 + int32_t(g);
                    if (c < need && c < kDecStageChunks)
                        sdst[c] = __ldg(gsrc + c);
                }
            }
            __syncwarp();
            const uint8_t* stage = row + kDecPad - lo; // stage[pos] is payload byte `pos`

#pragma unroll
            for (int i = 0; i < kDecTile; i++) {
                const uint32_t s = s0 + uint32_t(i);
                // quads first, all four lanes; then the tail on chain 0 (lane g == 3)
                const bool active = s < quads ? true : (s < my_steps && g == 3u);
                const uint32_t idx = s < quads ? 4u * s + g : s + 3u * quads; // 4*quads + (s - quads)
                const uint32_t slot = uint32_t(x) & frame_mask;
                // slot -> symbol: largest sym in [bucket[b], bucket[b + 1]] with cum[sym] <= slot
                const uint32_t b = slot >> bshift;
                // (measured in round 2 and dropped: the same search with a trip count fixed per task -- the widest bucket
                // of the warp's eight models -- and selects instead of the branch: 1.141 -> 1.164 ms, r02ay)
                uint32_t slo = bucket[b], shi = bucket[b + 1u];
                while (slo < shi) {
                    const uint32_t mid = (slo + shi + 1u) >> 1;
                    if (uint32_t(cum[mid]) <= slot)
                        slo = mid;
                    else
                        shi = mid - 1u;
                }
                const uint32_t c0 = cum[slo];
                const uint32_t freq = (uint32_t(cum[slo + 1u]) - c0) & 0xFFFFu;
                const uint32_t offset = slot - c0;
                uint32_t k;
                const uint32_t mapped = pa_unmap<METHOD>(slo, f, k);
                uint64_t nx = uint64_t(freq) * (x >> log2_frame) + offset;
                const bool renorm = active && nx < lower;
                k = active ? k : 0u;
                const uint32_t c = k + (renorm ? 4u : 0u);
                // forward inclusive scan: lane 0 of the group consumes first
                uint32_t incl, total;
                group_scan(k, renorm, grp_base, scan_mine, incl, total);
                const int32_t end = cursor - int32_t(incl - c); // my bytes are [end - c, end)
                // 8-byte window [end-8, end) from three aligned words
                const int32_t wa = end - 8;
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(stage + (wa & ~3));
                const uint32_t sh = (uint32_t(wa) & 3u) * 8u;
                const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2];
                const uint32_t win_lo = __funnelshift_r(w0, w1, sh);
                const uint32_t win_hi = __funnelshift_r(w1, w2, sh);
                if (renorm)
                    nx = (nx << kRadixLog2) | uint64_t(win_hi);
                const uint32_t exc_src = renorm ? win_lo : win_hi;
                const uint32_t exc = __funnelshift_rc(exc_src, 0u, 32u - 8u * k); // top k bytes; 0 when k == 0
                if (active) {
                    x = nx;
                    out[idx] = mapped + exc;
                }
                cursor -= int32_t(total);
            }
        }
    }
}

} // namespace b200ans
