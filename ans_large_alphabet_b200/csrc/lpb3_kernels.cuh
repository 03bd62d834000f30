// b200ans device code (sm_100a): lane-per-block encoder, third generation ("L3").
//
// ncu on the second-generation kernel (lpb_kernels.cuh) showed it bound by two things at once:
// the integer ALU pipe (46 ALU instructions per symbol at one warp instruction per 2 cycles and
// scheduler) and the shared-memory data pipe (21 wavefronts per 32 symbols).  This kernel keeps
// the decomposition -- one lane owns one block and carries the four interleaved rANS states of
// ans_fold.hpp:250-278 in registers, a warp works on 32 consecutive blocks -- and attacks both:
//
//  * input tiles (32 blocks x 8 symbols) arrive by TMA (cp.async.bulk.tensor.2d, 32-byte
//    swizzle, mbarrier completion): one instruction per warp and tile instead of 32 LDGSTS
//    lanes whose sectors each cost a shared-memory wavefront;
//  * the {freq, base, reciprocal} table is REPLICATED R times in shared memory, replica r
//    interleaved so that it only occupies the 16-byte bank group r: the 8 lanes of a quarter
//    warp read 8 different bank groups, so a random 128-bit lookup costs 4 wavefronts, not ~10;
//  * the state update runs on the FP64 and FMA pipes (exact reciprocal divide by DFMA, the
//    multiply by the frame size M as IMAD/IMAD.HI), the byte packer is a one-hot multiplier
//    (acc += piece * 2^fill as IMAD/IMAD.HI, the fill advances by a rotate), validity is a
//    sticky NaN accumulated by an FFMA -- what is left on the ALU pipe is ~16 instructions
//    per symbol.
//
// Byte-identity with the reference coder run on each block is unchanged: the emission order
// inside a block is exactly the serial order (ans_fold.hpp:100-112,257-278).
#pragma once

#include <cuda.h>

#include <type_traits>

#include "lpb_kernels.cuh"

namespace b200ans {

// Launch shapes.  TQ = quads per lane and input tile (rows of 16 * TQ bytes), STAGES = input tiles in
// flight per warp, DQ = quads between drains of the lane's output row, MAXW = warps per CTA (one CTA per SM).
// Measured on B200 (tools/micro/gather_bw.cu): the TMA gather of 32 rows x 64 B runs at 5.8 TB/s with 14
// warps per SM but drops to 3.9 TB/s with 28 (more streams in flight than the memory system likes), so the
// default shape is few, well-fed warps with registers to spare.
struct L3Cfg {
    static constexpr int TQ = 4, STAGES = 2, DQ = 2, MAXW = 16;
};
// (A 22-warp shape of the same encoder -- 92 registers per thread still fit -- was measured in round 2: parity green,
// but 2.11 ms against 1.705 ms for ANSfold-1 at 2^30 symbols, 2.31 against 2.01 for ANSfold-2, 2.04 against 1.98 for
// ANS: more TMA gather streams per SM than the memory system likes.  Removed.)
// output row: 12 carried bytes + 28 per quad + slack for the final partial word; an odd number of 16-byte chunks
template <class CFG> struct L3Geom {
    static constexpr int kTileBytes = 32 * CFG::TQ * 16;
    static constexpr int kOutRow = 112; // 28 carried bytes + 56 of two quads (+ the final partial word), 7 chunks
    static constexpr int kWarpBytes = CFG::STAGES * kTileBytes + 32 * kOutRow + 8 * CFG::STAGES;
    static_assert(CFG::DQ == 2 && 28 + 28 * CFG::DQ + 4 <= kOutRow && (kOutRow / 16) % 2 == 1, "output row too small");
    static_assert(CFG::TQ % CFG::DQ == 0 && (CFG::TQ == 2 || CFG::TQ == 4 || CFG::TQ == 8), "tile shape");
};

struct Enc3Args {
    uint32_t rep_log2;      // log2 of the number of table replicas (0..3)
    uint32_t table_alloc;   // entries per replica: covers every symbol the mapping can produce
    uint32_t n_full_blocks; // rows of the tensor map (blocks that hold block_syms symbols)
    uint32_t max_value;     // FOLD / MSB: largest input value whose symbol is < n_entries (RFOLD: largest reordered value)
    // RFOLD: perfect hash of the K = 2^(f+7) most frequent values (built by the host, b200ans.cu rfold_build_hash):
    // slot(v) = (v * mul + disp[(v * mul) >> bshift]) & (K - 1); v is one of them iff tags[slot] == v, and the table
    // entry of its id then sits AT the slot (the first K table entries are permuted accordingly)
    const uint32_t* rf_tags; // K values (0xFFFFFFFF = empty)
    const uint16_t* rf_disp; // 2^(32 - bshift) displacements
    uint32_t rf_mul;
    uint32_t rf_bshift;
    // PA (pseudo-adaptive mode: every block has its own model, EncArgs::pa_*): per-lane prefix-sum tables
    uint32_t pa_rcp_n;      // reciprocals kept per CTA: frequencies below this never divide
    uint32_t pa_tab_stride; // bytes between the lanes' tables (a multiple of 8)
};

// RFOLD: where the hash lives in shared memory
struct L3Rfold {
    uint32_t s_tags, s_disp, mul, bshift, kmask, K;
};

// ---- mbarrier / TMA (PTX ISA 8.x, sm_90+) ---------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "L3_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra L3_DONE;\n"
        "bra L3_WAIT;\n"
        "L3_DONE:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
// box (16 symbols x 32 blocks) at (column x, row y) of the [n_full_blocks][block_syms] view of the input
__device__ __forceinline__ void tma_load_tile(uint32_t dst, const CUtensorMap* map, int32_t x, int32_t y, uint32_t bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(x), "r"(y), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(pred));
    return pred != 0u;
}

__device__ __forceinline__ uint32_t bfind_u32(uint32_t x)
{
    uint32_t r;
    asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x));
    return r;
}
__device__ __forceinline__ uint32_t mad_lo(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t mad_hi(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }

// Word-granular byte packer.  `acc` holds the pending bytes (little-endian, fewer than 4),
// `am` = 2^(8 * pending bytes) as a one-hot multiplier: appending a piece is acc + piece * am
// (low half: the word to store, high half: what spills into the next word), the fill advances
// by rotating `am`, and a completed word shows as the rotate wrapping around.
struct L3Packer {
    uint32_t acc;
    uint32_t am;
    uint32_t wp; // shared-memory address of the next word of the lane's row
    __device__ __forceinline__ void put_bits(uint32_t piece, uint32_t nbits) // nbits in {0, 8, 16, 24}, piece < 2^nbits
    {
        const uint32_t t = mad_lo(piece, am, acc);
        const uint32_t am2 = __funnelshift_l(am, am, nbits);
        const bool full = am2 < am;
        acc = t;
        if (full) {
            sts32(wp, t);
            wp += 4u;
            acc = mad_hi(piece, am, 0u); // what spilled over the word
        }
        am = am2;
    }
    __device__ __forceinline__ void put_word(uint32_t w, bool on)
    {
        if (on) {
            sts32(wp, mad_lo(w, am, acc));
            wp += 4u;
            acc = mad_hi(w, am, 0u);
        }
    }
    __device__ __forceinline__ uint32_t pending_bytes() const { return bfind_u32(am) >> 3; }
};

// value -> shared-memory address of its table entry (replica chosen by `lane_tb`) and 8 * (raw low bytes)
// RFOLD (ans_reorder_fold.hpp:156-168: encode_symbol folds mapping[v]): `v` comes back as the reordered value the raw
// bytes are taken from -- 0 for one of the K most frequent values (no raw bytes, entry at its hash slot), v + K otherwise.
template <int METHOD, int FID>
__device__ __forceinline__ uint32_t l3_entry_addr(uint32_t& v, uint32_t lane_tb, uint32_t ent_stride, uint32_t k8_stride, uint32_t f,
    uint32_t n_entries, uint32_t& k8, const L3Rfold& rf)
{
    if (METHOD == METHOD_RFOLD) {
        const uint32_t h = v * rf.mul;
        uint32_t d, t;
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(d) : "r"(rf.s_disp + ((h >> rf.bshift) << 1)));
        const uint32_t s = (h + d) & rf.kmask;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t) : "r"(rf.s_tags + (s << 2)));
        const bool hot = t == v;
        v = hot ? 0u : v + rf.K;
        const uint32_t fl = bfind_u32(v | (1u << (f - 1u))) - (f - 1u);
        k8 = fl & 0x18u;
        const uint32_t idx = hot ? s : (v >> k8);
        return mad_lo(idx, ent_stride, mad_lo(k8, k8_stride, lane_tb));
    }
    if (METHOD == METHOD_INT) {
        k8 = 0;
        return mad_lo(min(v, n_entries), ent_stride, lane_tb); // entry n_entries is the "no such symbol" sentinel
    }
    uint32_t fl;
    if (METHOD == METHOD_MSB) {
        // k = (x > 2^8) + (x > 2^16) + (x > 2^24)  (ans_msb.hpp:41-73)  ==  highest bit of (x - 1) / 8
        fl = bfind_u32((max(v, 1u) - 1u) | 1u);
    } else if (FID == 1) {
        fl = bfind_u32(v | 1u); // thres = 2^8: k = highest bit / 8  (ans_fold.hpp:38-50)
    } else {
        fl = bfind_u32(v | (1u << (f - 1u))) - (f - 1u); // thres = 2^(f+7)
    }
    k8 = fl & 0x18u;
    const uint32_t xs = v >> k8;
    // symbol = xs + k * step  ->  address = xs * stride + k8 * (step * stride / 8) + replica base
    return mad_lo(xs, ent_stride, mad_lo(k8, k8_stride, lane_tb));
}

//
// PA = pseudo-adaptive mode (src/pseudo_adaptive.cpp:64-177: one model per block): the lane's block has its OWN model, so
// instead of the CTA's replicated table every lane keeps the prefix sums of its model's frequencies (u16 per symbol,
// built by the warp at the start of the task from EncArgs::pa_nfreqs) in its own stretch of shared memory; base and
// frequency are two 16-bit loads, the reciprocal comes from a per-CTA table indexed by the frequency.  Everything else
// -- TMA tiles, packer, state update, drains -- is the same code.  (The first pseudo-adaptive coders, pa_kernels.cuh,
// give each block four lanes, one per state: 150 instructions per 4-symbol step and a shuffle-scan on the chain.)
template <int METHOD, int FID, class CFG, bool PA = false>
__global__ void __launch_bounds__(CFG::MAXW * 32, 1) lpb3_encode_kernel(const EncArgs a, const Enc3Args g, const __grid_constant__ CUtensorMap tmap)
{
    constexpr int TQ = CFG::TQ, STAGES = CFG::STAGES, DQ = CFG::DQ;
    constexpr int kTileBytes = L3Geom<CFG>::kTileBytes;
    constexpr int kOutRow = L3Geom<CFG>::kOutRow;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // layout: [table: table_alloc << rep_log2 entries][input tiles: warps x STAGES, 1 KiB aligned][output rows][mbarriers]
    // (the table comes first: a symbol beyond the model then reads -- and flags -- whatever follows, inside the window)
    const uint32_t warps_per_cta = blockDim.x >> 5;
    const uint32_t lane = lane_id();
    const uint32_t warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0); // warp-uniform for the compiler too
    const uint32_t s_base = smem_u32(smem_raw);
    const uint32_t s_table = s_base;
    const uint32_t table_bytes = PA ? g.pa_rcp_n * 8u + warps_per_cta * 32u * g.pa_tab_stride : (g.table_alloc << g.rep_log2) * 16u;
    // RFOLD: [tags: K x u32][displacements: u16] between the table and the tiles
    L3Rfold rf {};
    uint32_t aux_bytes = 0;
    if (METHOD == METHOD_RFOLD) {
        rf.K = 1u << (a.fidelity + 7u);
        rf.kmask = rf.K - 1u;
        rf.mul = g.rf_mul;
        rf.bshift = g.rf_bshift;
        rf.s_tags = s_base + ((table_bytes + 1023u) & ~1023u);
        rf.s_disp = rf.s_tags + rf.K * 4u;
        aux_bytes = (rf.K * 4u + ((1u << (32u - g.rf_bshift)) << 1) + 1023u) & ~1023u;
        for (uint32_t i = threadIdx.x; i < rf.K; i += blockDim.x)
            sts32(rf.s_tags + i * 4u, __ldg(g.rf_tags + i));
        for (uint32_t i = threadIdx.x; i < (1u << (32u - g.rf_bshift)); i += blockDim.x)
            asm volatile("st.shared.u16 [%0], %1;" ::"r"(rf.s_disp + i * 2u), "h"(__ldg(g.rf_disp + i)) : "memory");
    }
    const uint32_t s_tiles = s_base + ((table_bytes + 1023u) & ~1023u) + aux_bytes;
    const uint32_t s_rows = s_tiles + warps_per_cta * uint32_t(STAGES * kTileBytes);
    const uint32_t s_bars = s_rows + warps_per_cta * 32u * kOutRow;

    // ---- replicated table: entry e of replica r at ((e << rep_log2) + r) * 16 ---------------
    if (PA) { // reciprocals by frequency; the lanes' tables follow per task
        for (uint32_t i = threadIdx.x; i < g.pa_rcp_n; i += blockDim.x) {
            const double r = i ? pa_reciprocal_up(i) : 0.0;
            asm volatile("st.shared.f64 [%0], %1;" ::"r"(s_table + i * 8u), "d"(r) : "memory");
        }
    } else {
        const uint32_t total = g.table_alloc << g.rep_log2;
        const uint4* src = reinterpret_cast<const uint4*>(a.table);
        for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) {
            const uint32_t e = i >> g.rep_log2;
            uint4 raw = make_uint4(0u, 0u, 0u, 0x7FF80000u); // NaN reciprocal marks "no such symbol"
            if (e < a.n_entries) {
                const uint4 t = __ldg(src + e);
                if (t.x != 0u)
                    raw = make_uint4(0u - t.x, t.y, t.z, t.w); // {-freq, base, reciprocal}
            }
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(s_table + i * 16u), "r"(raw.x), "r"(raw.y), "r"(raw.z), "r"(raw.w) : "memory");
        }
    }
    const uint32_t bar0 = s_bars + warp * uint32_t(8 * STAGES);
    if (lane == 0) {
#pragma unroll
        for (int st = 0; st < STAGES; st++)
            mbar_init(bar0 + 8u * st, 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    uint32_t log2_frame = a.log2_frame; // (PA: the lane's own, set per task)
    uint32_t frame = 1u << log2_frame;
    const uint32_t f = a.fidelity;
    uint32_t lower_hi = (log2_frame + 4u >= 32u) ? (1u << (log2_frame + 4u - 32u)) : 0u; // L = 16 M (ans_fold.hpp:91)
    uint32_t lower_lo = (log2_frame + 4u >= 32u) ? 0u : (1u << (log2_frame + 4u));
    uint32_t neg_bias = 0u - (0x43300000u << log2_frame); // cancels the double's exponent bits in q_hi * M
    const uint32_t pa_tabs = s_table + g.pa_rcp_n * 8u + warp * 32u * g.pa_tab_stride; // PA: the warp's 32 tables
    const uint16_t* my_cum = reinterpret_cast<const uint16_t*>(smem_raw + g.pa_rcp_n * 8u + (warp * 32u + lane) * g.pa_tab_stride);
    const double* pa_rcp = reinterpret_cast<const double*>(smem_raw);
    bool pa_bad = false;
    uint32_t dom_freq = 0; // PA: the largest frequency of the lane's model, and its reciprocal
    double dom_rcp = 0.0;
    const uint32_t ent_stride = 16u << g.rep_log2;
    const uint32_t step = (METHOD == METHOD_MSB) ? 256u : (255u << (f - 1u));
    const uint32_t k8_stride = (METHOD == METHOD_INT) ? 0u : (step * ent_stride) >> 3;
    const uint32_t lane_tb = s_table + ((lane & ((1u << g.rep_log2) - 1u)) << 4);
    const uint32_t B = a.block_syms;
    const uint32_t my_tiles = s_tiles + warp * uint32_t(STAGES * kTileBytes);
    const uint32_t my_row_in = lane * (16u * TQ);                 // row offset inside a tile
    // 32 / 64 / 128-byte swizzle: the 16-byte chunk index is XORed with address bits 7.. (Swizzle<log2 TQ, 4, 3>)
    const uint32_t swz = ((lane >> (TQ == 2 ? 2 : TQ == 4 ? 1 : 0)) & uint32_t(TQ - 1)) << 4;
    const uint32_t my_out = s_rows + (warp * 32u + lane) * kOutRow;
    const double two52 = 4503599627370496.0;
    uint32_t phases = 0; // bit st: parity of the phase the next wait on stage st's mbarrier is for

    const uint32_t n_tasks = (a.n_blocks + 31u) >> 5;
    for (uint32_t task = blockIdx.x * warps_per_cta + warp; task < n_tasks; task += gridDim.x * warps_per_cta) {
        const uint32_t blk0 = task * 32u;
        const uint32_t blk = blk0 + lane;
        const bool have_block = blk < a.n_blocks;
        uint8_t* dst = a.scratch + uint64_t(blk) * a.stride;
        // full task: every block of it that exists holds B symbols (B is a multiple of 8)
        const bool task_full = min(blk0 + 32u, a.n_blocks) <= g.n_full_blocks;

        uint32_t xlo[4], xhi[4];
#pragma unroll
        for (int c = 0; c < 4; c++) { // initial_state() (ans_fold.hpp:113)
            xlo[c] = lower_lo;
            xhi[c] = lower_hi;
        }
        L3Packer pk;
        pk.acc = 0u;
        pk.am = 1u;
        pk.wp = my_out;
        uint32_t flushed = 0;
        float badf = 0.f;
        uint32_t vmax = 0u;

        if (PA) {
            // the 32 models of the task: cum[s] = slots before symbol s for s in [0, syms], cum[syms + 1] = cum[syms]
            // (a symbol beyond the model reads frequency 0 there), 128 symbols per step of the warp
            const uint32_t syms = a.pa_syms; // multiple of 16
            __syncwarp();
            for (uint32_t j = 0; j < 32u; j++) {
                const uint32_t m = min(blk0 + j, a.n_blocks - 1u);
                const uint4* nf = reinterpret_cast<const uint4*>(a.pa_nfreqs + uint64_t(m) * a.pa_nbins);
                const uint32_t tab = pa_tabs + j * g.pa_tab_stride;
                uint32_t carry = 0, top = 0, low_big = 0xFFFFFFFFu; // largest frequency; smallest one beyond the reciprocal table
                for (uint32_t base = 0; base < syms; base += 128u) {
                    const uint32_t s = base + 4u * lane;
                    const uint4 q = s < syms ? __ldg(nf + (s >> 2)) : make_uint4(0u, 0u, 0u, 0u);
                    const uint32_t mine = q.x + q.y + q.z + q.w;
                    top = max(top, max(max(q.x, q.y), max(q.z, q.w)));
                    low_big = min(low_big, min(min(q.x >= g.pa_rcp_n ? q.x : 0xFFFFFFFFu, q.y >= g.pa_rcp_n ? q.y : 0xFFFFFFFFu),
                                           min(q.z >= g.pa_rcp_n ? q.z : 0xFFFFFFFFu, q.w >= g.pa_rcp_n ? q.w : 0xFFFFFFFFu)));
                    uint32_t incl = mine;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= uint32_t(o))
                            incl += up;
                    }
                    const uint32_t c0 = carry + incl - mine, c1 = c0 + q.x, c2 = c1 + q.y, c3 = c2 + q.z;
                    if (s < syms)
                        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(tab + 2u * s), "r"((c0 & 0xFFFFu) | (c1 << 16)), "r"((c2 & 0xFFFFu) | (c3 << 16)) : "memory");
                    carry += __shfl_sync(0xffffffffu, incl, 31);
                }
                if (lane == 0)
                    sts32(tab + 2u * syms, (carry & 0xFFFFu) | (carry << 16));
                top = __reduce_max_sync(0xffffffffu, top);
                low_big = __reduce_min_sync(0xffffffffu, low_big);
                if (lane == j)
                    dom_freq = top;
                if (low_big != 0xFFFFFFFFu && low_big != top && lane == 0)
                    atomicOr(a.error_flag, 2u); // not for this kernel
            }
            dom_rcp = pa_reciprocal_up(max(dom_freq, 1u));
            __syncwarp();
            log2_frame = a.pa_log2[min(blk, a.n_blocks - 1u)];
            frame = 1u << log2_frame;
            lower_hi = 0u; // log2 M <= 15 here
            lower_lo = 1u << (log2_frame + 4u);
            neg_bias = 0u - (0x43300000u << log2_frame);
#pragma unroll
            for (int c = 0; c < 4; c++) {
                xlo[c] = lower_lo;
                xhi[c] = lower_hi;
            }
        }
        // table entry of one value: {-freq, base, reciprocal}, 8 * (raw low bytes); v comes back as l3_entry_addr leaves it
        auto entry = [&](uint32_t& v, bool active, uint32_t& k8, uint32_t& nfr, uint32_t& base, double& rcp) {
            if (PA) {
                uint32_t fl = 0;
                if (METHOD == METHOD_INT)
                    fl = 0; // plain ANS (dense ranks): the symbol is the value
                else if (METHOD == METHOD_MSB)
                    fl = bfind_u32((max(v, 1u) - 1u) | 1u);
                else if (FID == 1)
                    fl = bfind_u32(v | 1u);
                else
                    fl = bfind_u32(v | (1u << (f - 1u))) - (f - 1u);
                k8 = fl & 0x18u;
                const uint32_t pa_step = (METHOD == METHOD_MSB) ? 256u : (METHOD == METHOD_INT) ? 0u : (255u << (f - 1u));
                const uint32_t sym = min(mad_lo(k8 >> 3, pa_step, v >> k8), a.pa_syms); // entry pa_syms: frequency 0
                // (plain loads, not asm volatile: the compiler is free to overlap the four lookups of a quad)
                const uint32_t c0 = my_cum[sym], c1 = my_cum[sym + 1u];
                const uint32_t fr = (c1 - c0) & 0xFFFFu;
                base = c0;
                nfr = 0u - fr;
                // reciprocal: by table; beyond it (frequencies above 4096: at most a few symbols of a model, nearly always
                // one) the lane's largest frequency has its own in a register.  Straight-line code: a divide in this
                // place, inlined or called, doubled the kernel's time (r02x / r02z2) -- a model with TWO different
                // frequencies beyond the table sets bit 1 of the error flag instead and the host codes the input again
                // with pa_kernels.cuh.
                rcp = fr >= g.pa_rcp_n ? dom_rcp : pa_rcp[min(fr, g.pa_rcp_n - 1u)];
                pa_bad |= active && fr == 0u;
            } else {
                const uint32_t addr = l3_entry_addr<METHOD, FID>(v, lane_tb, ent_stride, k8_stride, f, a.n_entries, k8, rf);
                const uint4 e = lds128(addr);
                nfr = e.x;
                base = e.y;
                rcp = __hiloint2double(int(e.w), int(e.z));
                if (active)
                    badf = fmaf(__uint_as_float(e.w), 0.f, badf); // NaN (no such symbol) sticks
            }
        };

        // Whole 32-byte sectors go to global memory (a 16-byte store is a partial-sector write, which the
        // memory system handles at less than half the speed); what is left (< 32 bytes) moves to the row's front.
        auto drain = [&]() {
            const uint32_t fill = pk.wp - my_out; // multiple of 4
            const uint32_t sectors = fill >> 5;   // <= 2: at most 28 carried + 56 new bytes
            uint8_t* gp = dst + flushed;
            if (sectors != 0u) {
                const uint4 c0 = lds128(my_out), c1 = lds128(my_out + 16u);
                const uint4 c2 = lds128(my_out + 32u), c3 = lds128(my_out + 48u);
                uint4 c4 = c2, c5 = c3;
                if (sectors > 1u) {
                    c4 = lds128(my_out + 64u);
                    c5 = lds128(my_out + 80u);
                }
                if (have_block) {
                    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(gp), "r"(c0.x), "r"(c0.y), "r"(c0.z), "r"(c0.w), "r"(c1.x),
                                 "r"(c1.y), "r"(c1.z), "r"(c1.w)
                                 : "memory");
                    if (sectors > 1u)
                        asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(gp + 32), "r"(c2.x), "r"(c2.y), "r"(c2.z), "r"(c2.w),
                                     "r"(c3.x), "r"(c3.y), "r"(c3.z), "r"(c3.w)
                                     : "memory");
                }
                const uint4 k0 = c4, k1 = c5;
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(my_out), "r"(k0.x), "r"(k0.y), "r"(k0.z), "r"(k0.w) : "memory");
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(my_out + 16u), "r"(k1.x), "r"(k1.y), "r"(k1.z), "r"(k1.w) : "memory");
            }
            flushed += sectors << 5;
            pk.wp = my_out + (fill & 31u);
        };

        // one quad: index 4t+3 (state 0) first ... index 4t (state 3) last (ans_fold.hpp:262-271)
        auto code_quad = [&](const uint4 v4, bool active) {
            uint32_t v[4] = { v4.w, v4.z, v4.y, v4.x };
            uint32_t k8[4], nfr[4], base[4], fr16[4];
            double rcp[4];
            bool rn[4];
            if (METHOD != METHOD_INT && METHOD != METHOD_RFOLD)
                vmax = max(vmax, max(max(v[0], v[1]), max(v[2], v[3])));
#pragma unroll
            for (int c = 0; c < 4; c++) {
                entry(v[c], active, k8[c], nfr[c], base[c], rcp[c]);
                fr16[c] = nfr[c] * 0xFFFFFFF0u; // freq * 16
            }
            if (METHOD == METHOD_RFOLD) // the reordered values (0 for the K most frequent ones)
                vmax = max(vmax, max(max(v[0], v[1]), max(v[2], v[3])));
#pragma unroll
            for (int c = 0; c < 4; c++) // state >= freq * K * 2^32 (ans_fold.hpp:105)
                rn[c] = active && (xhi[c] >= fr16[c]);
#pragma unroll
            for (int c = 0; c < 4; c++) {
                if (METHOD != METHOD_INT) { // raw low bytes first (ans_fold.hpp:59-63)
                    uint32_t mask;
                    asm("bmsk.clamp.b32 %0, 0, %1;" : "=r"(mask) : "r"(active ? k8[c] : 0u));
                    pk.put_bits(v[c] & mask, active ? k8[c] : 0u);
                }
                pk.put_word(xlo[c], rn[c]); // then the renormalisation word (ans_fold.hpp:106-109)
            }
#pragma unroll
            for (int c = 0; c < 4; c++) {
                // x -> (x / f) * M + x % f + base (ans_fold.hpp:111); x < 2^52 after the renormalisation
                const uint32_t y_lo = rn[c] ? xhi[c] : xlo[c];
                const uint32_t y_hi = rn[c] ? 0x43300000u : (xhi[c] | 0x43300000u);
                const double xd = __hiloint2double(int(y_hi), int(y_lo)) - two52;
                const double t = __fma_rz(xd, rcp[c], two52); // 2^52 + floor(x / f), see ans_push52
                const uint32_t q_lo = uint32_t(__double2loint(t));
                const uint32_t q_hr = uint32_t(__double2hiint(t)); // 0x43300000 + (q >> 32)
                const uint32_t r = mad_lo(q_lo, nfr[c], y_lo);     // x - q * f  (< f)
                // q * M + (r + base) as one 64-bit multiply-add; the addend's high word cancels the exponent bits
                // that q_hr * M leaves behind (zero for log2 M >= 12)
                uint64_t addend, prod;
                asm("mov.b64 %0, {%1, %2};" : "=l"(addend) : "r"(r + base[c]), "r"(neg_bias));
                asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(prod) : "r"(q_lo), "r"(frame), "l"(addend));
                const uint32_t nlo = uint32_t(prod);
                const uint32_t nhi = mad_lo(q_hr, frame, uint32_t(prod >> 32));
                if (active) {
                    xlo[c] = nlo;
                    xhi[c] = nhi;
                }
            }
        };

        if (task_full) {
            // tiles of 4 TQ symbols per block, walking backwards, STAGES - 1 tiles ahead of the coder
            const uint32_t tiles = B / (4u * TQ);
            auto issue = [&](uint32_t t, uint32_t st) { // tile t -> stage st
                if (elect_one()) {
                    mbar_expect_tx(bar0 + st * 8u, kTileBytes);
                    tma_load_tile(my_tiles + st * kTileBytes, &tmap, int32_t(B - 4u * TQ * (t + 1u)), int32_t(blk0), bar0 + st * 8u);
                }
            };
#pragma unroll
            for (int st = 0; st < STAGES - 1; st++)
                if (uint32_t(st) < tiles)
                    issue(uint32_t(st), uint32_t(st));
            uint32_t st = 0;              // stage of tile t
            uint32_t st_next = STAGES - 1; // stage of tile t + STAGES - 1
            for (uint32_t t = 0; t < tiles; t++) {
                __syncwarp(); // every lane has finished reading the stage that is refilled now
                if (t + uint32_t(STAGES - 1) < tiles)
                    issue(t + uint32_t(STAGES - 1), st_next);
                mbar_wait(bar0 + st * 8u, (phases >> st) & 1u);
                phases ^= 1u << st;
                const uint32_t row = my_tiles + st * kTileBytes + my_row_in;
#pragma unroll
                for (int j = TQ - 1; j >= 0; j -= 2) { // the last symbols of the tile are coded first
                    const uint4 qa = lds128(row + (uint32_t(j * 16) ^ swz));
                    const uint4 qb = lds128(row + (uint32_t((j - 1) * 16) ^ swz));
                    code_quad(qa, true);
                    code_quad(qb, true);
                    drain();
                }
                st = st + 1u == uint32_t(STAGES) ? 0u : st + 1u;
                st_next = st_next + 1u == uint32_t(STAGES) ? 0u : st_next + 1u;
            }
        } else {
            // the one task that holds the partial last block: predicated, loads straight from global memory
            const uint64_t first = uint64_t(blk) * B;
            const uint32_t nb = have_block ? uint32_t(min(uint64_t(B), a.n - first)) : 0u;
            const uint32_t tail = nb & 3u;
            const uint32_t quads = nb >> 2;
            const uint32_t max_quads = __reduce_max_sync(0xffffffffu, quads);
            if (__any_sync(0xffffffffu, tail != 0u)) { // tail symbols on state 0 (ans_fold.hpp:257-261)
                for (uint32_t s = 0; s < 3u; s++) {
                    const bool active = s < tail;
                    uint32_t v = active ? __ldg(a.in + first + (nb - 1u - s)) : 0u;
                    // run the symbol as a "quad" whose other three members are inactive: rotate it into slot 0
                    uint32_t k8;
                    uint4 e;
                    double e_rcp;
                    entry(v, active, k8, e.x, e.y, e_rcp);
                    e.z = uint32_t(__double2loint(e_rcp)), e.w = uint32_t(__double2hiint(e_rcp));
                    if (active)
                        vmax = max(vmax, v);
                    const uint32_t fr16 = e.x * 0xFFFFFFF0u;
                    const bool rn = active && (xhi[0] >= fr16);
                    if (METHOD != METHOD_INT) {
                        const uint32_t kk = active ? k8 : 0u;
                        pk.put_bits(v & ((1u << kk) - 1u), kk);
                    }
                    pk.put_word(xlo[0], rn);
                    const uint32_t y_lo = rn ? xhi[0] : xlo[0];
                    const uint32_t y_hi = rn ? 0x43300000u : (xhi[0] | 0x43300000u);
                    const double xd = __hiloint2double(int(y_hi), int(y_lo)) - two52;
                    const double t = __fma_rz(xd, __hiloint2double(int(e.w), int(e.z)), two52);
                    const uint32_t q_lo = uint32_t(__double2loint(t));
                    const uint32_t q_hr = uint32_t(__double2hiint(t));
                    const uint32_t r = mad_lo(q_lo, e.x, y_lo);
                    if (active) {
                        xlo[0] = mad_lo(q_lo, frame, r + e.y);
                        xhi[0] = mad_lo(q_hr, frame, mad_hi(q_lo, frame, neg_bias));
                    }
                }
                drain();
            }
            const uint4* in4 = reinterpret_cast<const uint4*>(a.in + first);
            const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
            // two quads of read-ahead in registers (consecutive quads share L1 lines)
            uint4 nxt0 = quads > 0u ? __ldg(in4 + (quads - 1u)) : zero4;
            uint4 nxt1 = quads > 1u ? __ldg(in4 + (quads - 2u)) : zero4;
            for (uint32_t q = 0; q < max_quads; q += 2u) {
                const uint4 qa = nxt0, qb = nxt1;
                nxt0 = (q + 2u < quads) ? __ldg(in4 + (quads - 3u - q)) : zero4;
                nxt1 = (q + 3u < quads) ? __ldg(in4 + (quads - 4u - q)) : zero4;
                code_quad(qa, q < quads);
                code_quad(qb, q + 1u < quads);
                drain();
            }
        }

        // flush_state x4 in chain order 0..3 (ans_fold.hpp:115-120, 275-278)
        {
#pragma unroll
            for (int c = 0; c < 4; c++) {
                // state - L as 64 bits
                const uint32_t lo = xlo[c] - lower_lo;
                const uint32_t hi = xhi[c] - lower_hi - (xlo[c] < lower_lo ? 1u : 0u);
                pk.put_word(lo, true);
                pk.put_word(hi, true);
            }
            uint32_t fill = pk.wp - my_out; // <= 28 + 32
            const uint32_t total = flushed + fill + pk.pending_bytes();
            // last partial word, then zero padding up to the next 16-byte boundary (deterministic container)
            sts32(pk.wp, pk.acc);
            fill += 4u;
            while (fill & 15u) {
                sts32(my_out + fill, 0u);
                fill += 4u;
            }
            const uint32_t chunks = (total - flushed + 15u) >> 4;
            uint4* gv = reinterpret_cast<uint4*>(dst + flushed);
#pragma unroll
            for (uint32_t ci = 0; ci < 4u; ci++)
                if (have_block && ci < chunks)
                    gv[ci] = lds128(my_out + ci * 16u);
            if (have_block)
                a.block_bytes[blk] = total;
        }
        bool bad = badf != badf || (PA && pa_bad);
        if (METHOD != METHOD_INT)
            bad = bad || vmax > g.max_value;
        if (bad && have_block)
            *a.error_flag = 1u;
        __syncwarp();
    }
}

} // namespace b200ans

namespace b200ans {

// ==========================================================================================
// L3 decoder
// ==========================================================================================
// Same decomposition as the encoder (lane = block, the four states of ans_fold.hpp:283-311 in
// registers).  What changed against lpb_kernels.cuh's decoder (47 ALU instructions per symbol):
//
//  * the payload is read as a stream of aligned 32-bit words going down, through a left-aligned
//    bit accumulator (`acc` holds the unread low bytes of the current word at its top): the
//    renormalisation word is acc | next >> fill, the k raw bytes are the top 8k bits -- no
//    per-symbol window of three words, no byte positions;
//  * the lane-private cp.async ring is refilled once per loop iteration by a warp-uniform loop
//    (usually one trip) instead of three predicated copies per quad;
//  * decoded quads go straight from registers to global memory as 16-byte (QI == 1) or 32-byte
//    (QI == 2: one full sector, STG.256) stores -- no staging rows, no __syncwarp in the loop.
template <int QI> struct L3DecGeom {
    static constexpr int kRing = QI == 1 ? 128 : 256;      // bytes of ring per lane (power of two)
    static constexpr int kDist = 56 * QI + 4;               // bytes below the cursor resident after a refill:
                                                            // this iteration's 28 QI, the next one's, word alignment
    static constexpr int kWarpBytes = 32 * kRing;
    static_assert(kDist + 15 + 16 <= kRing, "ring too small for the prefetch distance");
};

__device__ __forceinline__ void cp_async16_s(uint32_t saddr, const void* gptr)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(gptr) : "memory");
}

struct L3Reader {
    uint32_t acc;  // unread low bits of the current word (right-aligned, everything above is zero)
    uint32_t fill; // how many: 0, 8, 16, 24
    uint32_t mask; // 2^fill - 1
    int32_t rp;    // byte offset (from the 16-byte aligned payload base) of the next word to read
    uint32_t ring; // shared-memory address of the lane's ring (aligned to its size)
    template <int RING> __device__ __forceinline__ uint32_t next_addr() const { return (uint32_t(rp) & uint32_t(RING - 1)) | ring; }
    // renormalisation (ans_fold.hpp:221-224): state = state << 32 | the next 32 bits of the stream
    template <int RING> __device__ __forceinline__ void renorm(bool on, uint32_t& lo, uint32_t& hi)
    {
        if (on) {
            uint32_t t;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(t) : "r"(next_addr<RING>()));
            rp -= 4;
            hi = lo;
            lo = __funnelshift_r(t, acc, fill); // acc's `fill` bits on top of the new word's upper part
            acc = t & mask;
        }
    }
    // the next k bytes (0..3) of the stream as an integer (ans_fold.hpp:135-147)
    template <int RING> __device__ __forceinline__ uint32_t get_bytes(uint32_t k)
    {
        const int32_t s = int32_t(mad_lo(k, 0xFFFFFFF8u, fill)); // fill - 8 k
        uint32_t v = __funnelshift_r(acc, 0u, uint32_t(s));       // acc >> s when the bits are all in acc
        uint32_t lo = acc;
        if (s < 0) { // they run into the next word: (acc : word) >> (s + 32)
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(lo) : "r"(next_addr<RING>()));
            rp -= 4;
            v = __funnelshift_r(lo, acc, uint32_t(s));
        }
        fill = uint32_t(s) & 31u;
        asm("bmsk.clamp.b32 %0, 0, %1;" : "=r"(mask) : "r"(fill));
        acc = lo & mask;
        return v;
    }
};

// (Measured in round 2 and dropped: refilling the rings in 64-byte chunks by groups of four lanes, so that one LDGSTS
// touches 8 lines instead of 32 and brings whole sectors -- byte-exact, 1.955 -> 1.947 ms for ANSfold-1 at 2^30 symbols,
// 2.32 -> 2.47 ms on the rank/select tables: the copies are not what the LSU pipe spends its wavefronts on.)
// (Also measured in round 2 and dropped: the decoded symbols leaving through TMA -- a swizzled 1 KiB tile per warp in
// shared memory, two conflict-free STS.128 per lane and iteration, one cp.async.bulk.tensor store of 32 blocks x 8
// symbols -- instead of a 32-byte STG per lane (14 LSU wavefronts per warp and iteration).  Byte-exact, and the LSU
// wavefronts do go down, but the kernel does not get faster: ANSfold-1 1.956 -> 1.999 ms, ANS 2.18 -> 2.95 ms, the
// 2^18-slot model 2.274 -> 2.240 ms at 2^30 symbols.)
// (Measured in round 2 and dropped: a pseudo-adaptive flavour of this decoder -- one model per block, kept by its lane in
// shared memory as 16-bit prefix sums + a bucket index, slot -> symbol by a binary search with a trip count fixed per
// task, as the PA flavour of the encoder above keeps its tables.  Byte-exact, but 118 instructions per symbol against 39
// here, the search being most of them: 1.40 ms for 2^27 symbols in 16 384 blocks against 1.14 ms for the four-lanes-per-
// block decoder of pa_kernels.cuh; a finer bucket index no longer leaves room for four warps per SM.)
template <int METHOD, int TABLE, int QI, int MAXW>
__global__ void __launch_bounds__(MAXW * 32, 1) lpb3_decode_kernel(const DecArgs a)
{
    if (a.error_flag && *a.error_flag >= 3u) // the size checks in front of this launch failed (check_block_sizes_kernel)
        return;
    constexpr uint32_t ALIGN = 16u; // payload view and refill granularity
    constexpr int RING = L3DecGeom<QI>::kRing;
    constexpr int DIST = L3DecGeom<QI>::kDist;
    constexpr bool kInt = (METHOD == METHOD_INT);
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t frame = 1u << a.log2_frame;
    DecTables<METHOD, TABLE> tables;
    const size_t table_bytes = tables.load(a, smem_raw);
    uint8_t* s_rings = smem_raw + table_bytes;
    s_rings += (uint32_t(RING) - (smem_u32(s_rings) & uint32_t(RING - 1))) & uint32_t(RING - 1); // the host adds RING bytes of slack

    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps_per_cta = blockDim.x >> 5;
    const uint32_t log2_frame = a.log2_frame;
    const uint32_t frame_mask = frame - 1u;
    const uint32_t lower_shift = log2_frame + 4u; // L = 16 M (ans_fold.hpp:91); at most 32
    const uint32_t lower_lo = lower_shift >= 32u ? 0u : (1u << lower_shift);
    const uint32_t lower_hi = lower_shift >= 32u ? 1u : 0u;
    const uint32_t rfold_bias = (METHOD == METHOD_RFOLD) ? (1u << (a.fidelity + 7)) : 0u; // ans_reorder_fold.hpp:301
    const uint32_t B = a.block_syms;

    L3Reader rd;
    rd.ring = smem_u32(s_rings) + (warp * 32u + lane) * uint32_t(RING);

    const uint32_t n_tasks = (a.n_blocks + 31u) >> 5;
    for (uint32_t task = blockIdx.x * warps_per_cta + warp; task < n_tasks; task += gridDim.x * warps_per_cta) {
        const uint32_t blk = task * 32u + lane;
        const bool have_block = blk < a.n_blocks;
        const uint64_t first = uint64_t(blk) * B;
        const uint32_t nb = have_block ? uint32_t(min(uint64_t(B), a.n - first)) : 0u;
        const uint32_t tail = nb & 3u;
        const uint32_t quads = nb >> 2;
        const uint32_t max_quads = __reduce_max_sync(0xffffffffu, quads);
        const bool uniform = __all_sync(0xffffffffu, have_block && quads == max_quads);

        // ALIGN-byte aligned view of the payload; positions count bytes from pay16
        const uint8_t* pay = a.payload + (have_block ? a.offsets[blk] : 0);
        const uint32_t delta = uint32_t(reinterpret_cast<uintptr_t>(pay) & (ALIGN - 1u));
        const uint8_t* pay16 = pay - delta;
        const uint32_t size = have_block ? a.block_bytes[blk] : 32u;
        const int32_t p0 = int32_t(delta + size - 32u); // one past the last payload byte below the four states

        // init_state x4: state c sits at payload + size - 32 + 8c (ans_fold.hpp:291-295)
        uint32_t xlo[4] = { lower_lo, lower_lo, lower_lo, lower_lo }, xhi[4] = { lower_hi, lower_hi, lower_hi, lower_hi };
        rd.acc = 0u;
        rd.fill = (uint32_t(p0) & 3u) * 8u;
        rd.mask = (1u << rd.fill) - 1u;
        rd.rp = (p0 & ~3) - 4;
        if (have_block) {
            const uint32_t* words = reinterpret_cast<const uint32_t*>(pay16);
            const int32_t w = p0 >> 2;
            const uint32_t sh = rd.fill;
            uint32_t t[9];
#pragma unroll
            for (int i = 0; i < 9; i++)
                t[i] = (i < 8 || sh != 0u) ? __ldg(words + w + i) : 0u;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const uint32_t lo = __funnelshift_r(t[2 * c], t[2 * c + 1], sh);
                const uint32_t hi = __funnelshift_r(t[2 * c + 1], t[2 * c + 2], sh);
                const uint32_t slo = xlo[c] + lo;
                xhi[c] = xhi[c] + hi + (slo < lo ? 1u : 0u);
                xlo[c] = slo;
            }
            // the word that holds byte p0 - 1 .. : its low (p0 & 3) bytes are still unread
            rd.acc = t[0] & rd.mask;
        }

        // ring: chunks [fetched_lo, ...) are resident (or in flight); refill down to `DIST` bytes below the cursor
        int32_t fetched_lo = (rd.rp + 4 + int32_t(ALIGN) - 1) & ~int32_t(ALIGN - 1u);
        auto refill = [&]() {
            const int32_t want = max(rd.rp + 4 - DIST, 0) & ~int32_t(ALIGN - 1u);
            while (__any_sync(0xffffffffu, have_block && fetched_lo > want)) {
                if (have_block && fetched_lo > want) {
                    fetched_lo -= 16;
                    cp_async16_s(rd.ring + (uint32_t(fetched_lo) & uint32_t(RING - 1)), pay16 + fetched_lo);
                }
            }
            cp_async_commit();
        };
        refill();

        // one symbol per chain: slot lookup + state update (ans_fold.hpp:216-220); returns the renormalisation flag
        uint32_t nlo[4], nhi[4], raw[4], kk[4];
        bool rn[4];
        auto advance = [&](int g, bool active) {
            const int c = 3 - g; // index 4t+g is decoded by state 3-g (ans_fold.hpp:300-305)
            uint32_t freq, offset, k;
            tables.lookup_raw(xlo[c] & frame_mask, freq, offset, raw[g], k);
            kk[g] = active ? k : 0u;
            const uint32_t q_lo = __funnelshift_r(xlo[c], xhi[c], log2_frame);
            const uint32_t q_hi = xhi[c] >> log2_frame;
            const uint64_t p = uint64_t(freq) * q_lo + offset;
            nlo[g] = uint32_t(p);
            nhi[g] = uint32_t(p >> 32) + freq * q_hi;
            // state < L = 2^(log2 M + 4)
            rn[g] = active && (lower_shift >= 32u ? nhi[g] == 0u : (nhi[g] == 0u && nlo[g] < lower_lo));
        };
        auto finish = [&](int g, bool active) -> uint32_t {
            const int c = 3 - g;
            rd.renorm<RING>(rn[g], nlo[g], nhi[g]); // renormalisation word first (ans_fold.hpp:221-224)
            uint32_t value = raw[g];
            if (!kInt) // then k raw low bytes (ans_fold.hpp:135-147); raw = unfolded base | k << 30
                value = mad_lo(kk[g], 0xC0000000u, value + rd.get_bytes<RING>(kk[g]));
            if (active) {
                xlo[c] = nlo[g];
                xhi[c] = nhi[g];
            }
            return value - rfold_bias;
        };

        uint32_t* out = a.out + first;
        auto run = [&](auto full_tag) {
            constexpr bool FULL = decltype(full_tag)::value; // every lane has a block and all blocks have max_quads quads
            for (uint32_t q0 = 0; q0 < max_quads; q0 += QI) {
                refill();
                cp_async_wait<1>(); // everything this iteration can touch was committed an iteration ago
                uint32_t val[4 * QI];
#pragma unroll
                for (int j = 0; j < QI; j++) {
                    const bool active = FULL || (q0 + uint32_t(j)) < quads;
#pragma unroll
                    for (int g = 0; g < 4; g++)
                        advance(g, active);
#pragma unroll
                    for (int g = 0; g < 4; g++)
                        val[4 * j + g] = finish(g, active);
                }
                if (QI == 2) {
                    if (FULL || q0 + 1u < quads) {
                        asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(out + 4u * q0), "r"(val[0]), "r"(val[1]), "r"(val[2]),
                                     "r"(val[3]), "r"(val[4 * (QI - 1)]), "r"(val[4 * (QI - 1) + 1]), "r"(val[4 * (QI - 1) + 2]),
                                     "r"(val[4 * (QI - 1) + 3])
                                     : "memory");
                    } else if (q0 < quads) {
                        reinterpret_cast<uint4*>(out)[q0] = make_uint4(val[0], val[1], val[2], val[3]);
                    }
                } else {
                    if (FULL || q0 < quads)
                        reinterpret_cast<uint4*>(out)[q0] = make_uint4(val[0], val[1], val[2], val[3]);
                }
            }
        };
        if (uniform && (max_quads % QI) == 0u)
            run(std::true_type {});
        else
            run(std::false_type {});
        // tail symbols on state 0 (ans_fold.hpp:306-310); only the last block of the input has any
        if (__any_sync(0xffffffffu, tail != 0u)) {
            refill();
            cp_async_wait<0>();
            for (uint32_t s = 0; s < 3u; s++) {
                const bool active = s < tail;
                advance(3, active); // g = 3 <-> state 0
                const uint32_t v = finish(3, active);
                if (active)
                    out[4u * quads + s] = v;
            }
        }
        cp_async_wait<0>();
        // a well-formed block is consumed down to its first byte, exactly: the reader stands `fill / 8` bytes into the
        // word above rp
        if (have_block && a.error_flag && rd.rp + 4 + int32_t(rd.fill >> 3) != int32_t(delta))
            *a.error_flag = 2u;
    }
}

} // namespace b200ans
