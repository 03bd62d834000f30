// b200ans device code (sm_100a): "lane per block" (LPB) encoder / decoder -- the large-input path.
//
// kernels.cuh runs one block on 4 lanes (lane per rANS chain) and pays for it with three
// shuffles per symbol, byte-granular shared-memory stores and one 4-byte load per symbol.  When
// the input has at least a few thousand blocks there is enough parallelism to give every lane
// a whole block instead: the lane carries all four interleaved states of ans_fold.hpp:250-278 in
// registers, the byte cursor is a plain register (no scan), the emitted bytes are packed into
// 32-bit words in a register accumulator, and the four chains of a quad give the lane four
// independent dependency chains to overlap.  A warp works on 32 consecutive blocks.
//
// Byte-identity with the reference coder run on each block is unchanged: the emission order inside
// a block is exactly the serial order (tail symbols on state 0, then quads backwards with the
// symbol at index 4t+g on state 3-g; per symbol raw low bytes first, then the 32-bit
// renormalisation word: ans_fold.hpp:100-112,257-278).
#pragma once

#include "kernels.cuh"

namespace b200ans {

// ------------------------------------------------------------------------------------------
// exact state update
// ------------------------------------------------------------------------------------------
// x -> (x / f) * M + x % f + base  (ans_fold.hpp:111) for x < 2^52 (true whenever every f < 2^16,
// because x < f * 2^36 after the renormalisation test).  `rcp` is the smallest double >= 1/f.
// x is exact as a double; fma_rz(x, rcp, 2^52) = 2^52 + floor(x * rcp) because the sum lies in
// [2^52, 2^53) where doubles are the integers; and floor(x * rcp) = floor(x / f) because
// x * (rcp - 1/f) < 2^52 * 2^-52 / f = 1/f while frac(x / f) <= 1 - 1/f.  No fix-up needed.
__device__ __forceinline__ uint64_t ans_push52(uint32_t x_lo, uint32_t x_hi, uint32_t f, uint32_t base, double rcp,
    uint32_t log2_frame)
{
    const double two52 = 4503599627370496.0;
    const double xd = __hiloint2double(int(x_hi | 0x43300000u), int(x_lo)) - two52;
    const double t = __fma_rz(xd, rcp, two52);
    const uint32_t q_lo = uint32_t(__double2loint(t));
    const uint32_t q_hi = uint32_t(__double2hiint(t)) & 0xFFFFFu;
    const uint32_t r = x_lo - q_lo * f; // < f
    const uint64_t q = (uint64_t(q_hi) << 32) | q_lo;
    return (q << log2_frame) + uint64_t(r + base);
}

// ------------------------------------------------------------------------------------------
// K2-LPB: encoder
// ------------------------------------------------------------------------------------------
constexpr int kLpbEncWarps = 8;
constexpr int kLpbTileQuads = 4;                         // quads (of 4 symbols) per lane between drains
constexpr int kLpbInRow = kLpbTileQuads * 16 + 16;        // 80 B: 5 x 16 (odd) -> conflict-free 128-bit row reads
constexpr int kLpbOutRow = 176;                           // 12 carried + 28 * 16 symbols... see static_assert
static_assert(12 + 28 * kLpbTileQuads + 36 + 4 <= kLpbOutRow && (kLpbOutRow / 16) % 2 == 1, "output row too small");
constexpr int kLpbEncMaxSmemEntries = 6144;

// Word-granular emitter: bytes are collected little-endian in `acc` (sh = 8 * pending bytes,
// pending < 4) and complete 32-bit words go to the lane's row in shared memory.
struct LpbEmitter {
    uint32_t acc;
    uint32_t sh;  // 0, 8, 16, 24
    uint32_t wp;  // shared-memory byte address of the next word of the row
    __device__ __forceinline__ void put_bytes(uint32_t v, uint32_t k) // k in 0..3 low bytes of v
    {
        const uint32_t vm = v & ((1u << (8u * k)) - 1u);
        const uint32_t t = acc | (vm << sh);
        const uint32_t hi = __funnelshift_l(vm, 0u, sh); // vm >> (32 - sh), 0 when sh == 0
        const uint32_t nsh = sh + 8u * k;
        const bool full = nsh >= 32u;
        if (full) {
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(wp), "r"(t) : "memory");
            wp += 4u;
        }
        acc = full ? hi : t;
        sh = nsh & 31u;
    }
    __device__ __forceinline__ void put_word(uint32_t w, bool on)
    {
        const uint32_t t = acc | (w << sh);
        if (on) {
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(wp), "r"(t) : "memory");
            wp += 4u;
            acc = __funnelshift_l(w, 0u, sh); // w >> (32 - sh), 0 when sh == 0
        }
    }
};

template <int METHOD, bool TABLE_SMEM>
__global__ void __launch_bounds__(kLpbEncWarps * 32) lpb_encode_kernel(const EncArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    EncEntry* s_table = reinterpret_cast<EncEntry*>(smem_raw);
    uint8_t* s_rows = smem_raw + (TABLE_SMEM ? size_t(a.n_entries) * sizeof(EncEntry) : 0);
    if (TABLE_SMEM) {
        const uint4* src = reinterpret_cast<const uint4*>(a.table);
        uint4* dst = reinterpret_cast<uint4*>(s_table);
        for (uint32_t i = threadIdx.x; i < a.n_entries; i += blockDim.x)
            dst[i] = __ldg(src + i);
        __syncthreads();
    }
    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    uint8_t* in_rows = s_rows + size_t(warp) * 32 * (kLpbInRow + kLpbOutRow);
    uint8_t* out_rows = in_rows + 32 * kLpbInRow;
    uint8_t* my_in = in_rows + lane * kLpbInRow;
    uint8_t* my_out = out_rows + lane * kLpbOutRow;
    const uint32_t my_out_s = uint32_t(__cvta_generic_to_shared(my_out));
    const uint4* table4 = reinterpret_cast<const uint4*>(TABLE_SMEM ? s_table : a.table);
    const uint32_t log2_frame = a.log2_frame;
    const uint32_t f = a.fidelity;
    const bool fast = a.fast_div != 0u; // every frequency < 2^16
    const uint64_t lower = uint64_t(kAnsK) << log2_frame;
    const uint32_t B = a.block_syms;

    const uint32_t n_tasks = (a.n_blocks + 31u) >> 5;
    for (uint32_t task = blockIdx.x * kLpbEncWarps + warp; task < n_tasks; task += gridDim.x * kLpbEncWarps) {
        const uint32_t blk0 = task * 32u;
        const uint32_t blk = blk0 + lane;
        const bool have_block = blk < a.n_blocks;
        const uint64_t first = uint64_t(blk) * B;
        const uint32_t nb = have_block ? uint32_t(min(uint64_t(B), a.n - first)) : 0u;
        const uint32_t tail = nb & 3u;
        const uint32_t quads = nb >> 2;
        const uint32_t max_quads = __reduce_max_sync(0xffffffffu, quads);
        uint8_t* dst = a.scratch + uint64_t(blk) * a.stride;

        uint64_t x0 = lower, x1 = lower, x2 = lower, x3 = lower; // initial_state() (ans_fold.hpp:113)
        LpbEmitter em;
        em.acc = 0;
        em.sh = 0;
        em.wp = my_out_s;
        uint32_t flushed = 0;
        uint32_t bad = 0;

        auto encode_one = [&](uint32_t v, uint64_t& x, bool active) {
            if (METHOD == METHOD_RFOLD) {
                if (v >= a.n_values) { // value the model has never seen
                    bad |= uint32_t(active);
                    v = 0;
                }
                v = __ldg(a.mapping + v);
            }
            uint32_t k;
            uint32_t sym = map_value<METHOD>(v, f, k);
            if (sym >= a.n_entries) {
                bad |= uint32_t(active);
                sym = 0;
            }
            const uint4 raw = TABLE_SMEM ? table4[sym] : __ldg(table4 + sym);
            const uint32_t fr = raw.x, base = raw.y;
            const double rcp = __hiloint2double(int(raw.w), int(raw.z));
            bad |= uint32_t(active && fr == 0u);
            const uint32_t x_lo = uint32_t(x), x_hi = uint32_t(x >> 32);
            const bool renorm = active && ((x_hi >> 4) >= fr); // state >= freq * K * 2^32 (ans_fold.hpp:105)
            if (METHOD != METHOD_INT)
                em.put_bytes(v, active ? k : 0u); // raw low bytes first (ans_fold.hpp:59-63)
            em.put_word(x_lo, renorm);           // then the renormalisation word (ans_fold.hpp:106-109)
            const uint32_t y_lo = renorm ? x_hi : x_lo;
            const uint32_t y_hi = renorm ? 0u : x_hi;
            uint64_t nx;
            if (fast)
                nx = ans_push52(y_lo, y_hi, fr, base, rcp, log2_frame);
            else
                nx = ans_push((uint64_t(y_hi) << 32) | y_lo, fr, base, rcp, log2_frame);
            if (active)
                x = nx;
        };

        // tail symbols (only the very last block of the input has any): state 0 (ans_fold.hpp:257-261)
        if (__any_sync(0xffffffffu, tail != 0u)) {
            for (uint32_t s = 0; s < 3u; s++) {
                const bool active = s < tail;
                const uint32_t v = active ? __ldg(a.in + first + (nb - 1u - s)) : 0u;
                encode_one(v, x0, active);
            }
        }

        for (uint32_t q0 = 0; q0 < max_quads; q0 += kLpbTileQuads) {
            // ---- stage the next 16 symbols of each of the warp's 32 blocks (coalesced 64-byte pieces) ----
            __syncwarp();
#pragma unroll
            for (int it = 0; it < 4; it++) {
                const uint32_t row = uint32_t(it) * 8u + (lane >> 2);
                const uint32_t chunk = lane & 3u;
                const uint32_t rblk = blk0 + row;
                if (rblk < a.n_blocks) {
                    const uint64_t rfirst = uint64_t(rblk) * B;
                    const uint32_t rnb = uint32_t(min(uint64_t(B), a.n - rfirst));
                    const uint32_t rquads = rnb >> 2;
                    // quad index (ascending in memory) held by this chunk
                    const int64_t qi = int64_t(rquads) - int64_t(q0) - int64_t(kLpbTileQuads) + int64_t(chunk);
                    if (qi >= 0) {
                        const uint4 val = __ldg(reinterpret_cast<const uint4*>(a.in + rfirst) + qi);
                        *reinterpret_cast<uint4*>(in_rows + row * kLpbInRow + chunk * 16u) = val;
                    }
                }
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < kLpbTileQuads; j++) {
                const bool active = (q0 + uint32_t(j)) < quads;
                const uint4 v4 = *reinterpret_cast<const uint4*>(my_in + (kLpbTileQuads - 1 - j) * 16);
                // walking backwards: index 4t+3 (state 0) first ... index 4t (state 3) last (ans_fold.hpp:262-271)
                encode_one(v4.w, x0, active);
                encode_one(v4.z, x1, active);
                encode_one(v4.y, x2, active);
                encode_one(v4.x, x3, active);
            }
            // ---- drain whole 16-byte chunks of the lane's row, carry the rest to the front ----
            {
                const uint32_t fill = em.wp - my_out_s; // bytes in the row (multiple of 4)
                const uint32_t chunks = fill >> 4;
                const uint32_t max_chunks = __reduce_max_sync(0xffffffffu, chunks);
                const uint4* rv = reinterpret_cast<const uint4*>(my_out);
                uint4* gv = reinterpret_cast<uint4*>(dst + flushed);
                for (uint32_t ci = 0; ci < max_chunks; ci++)
                    if (ci < chunks)
                        gv[ci] = rv[ci];
                if (chunks != 0u) {
                    const uint4 carry = rv[chunks];
                    *reinterpret_cast<uint4*>(my_out) = carry;
                }
                flushed += chunks << 4;
                em.wp = my_out_s + (fill & 15u);
            }
        }

        // flush_state x4 in chain order 0..3 (ans_fold.hpp:115-120, 275-278)
        {
            const uint64_t s0 = x0 - lower, s1 = x1 - lower, s2 = x2 - lower, s3 = x3 - lower;
            em.put_word(uint32_t(s0), have_block), em.put_word(uint32_t(s0 >> 32), have_block);
            em.put_word(uint32_t(s1), have_block), em.put_word(uint32_t(s1 >> 32), have_block);
            em.put_word(uint32_t(s2), have_block), em.put_word(uint32_t(s2 >> 32), have_block);
            em.put_word(uint32_t(s3), have_block), em.put_word(uint32_t(s3 >> 32), have_block);
            uint32_t fill = em.wp - my_out_s;
            const uint32_t total = flushed + fill + (em.sh >> 3);
            // last partial word, then zero padding up to the next 16-byte boundary (deterministic container)
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(em.wp), "r"(em.acc) : "memory");
            fill += 4u;
            while (fill & 15u) {
                asm volatile("st.shared.u32 [%0], %1;" ::"r"(my_out_s + fill), "r"(0u) : "memory");
                fill += 4u;
            }
            const uint32_t chunks = (total - flushed + 15u) >> 4;
            const uint32_t max_chunks = __reduce_max_sync(0xffffffffu, have_block ? chunks : 0u);
            const uint4* rv = reinterpret_cast<const uint4*>(my_out);
            uint4* gv = reinterpret_cast<uint4*>(dst + flushed);
            for (uint32_t ci = 0; ci < max_chunks; ci++)
                if (have_block && ci < chunks)
                    gv[ci] = rv[ci];
            if (have_block)
                a.block_bytes[blk] = total;
        }
        if (bad)
            *a.error_flag = 1u;
    }
}

// ------------------------------------------------------------------------------------------
// K3-LPB: decoder
// ------------------------------------------------------------------------------------------
// The lane walks its block's payload downwards from the four states.  The compressed bytes are
// staged in a lane-private 128-byte ring in shared memory, filled by asynchronous 16-byte copies
// (cp.async / LDGSTS) that run kLpbPrefetch bytes ahead of the cursor: one quad (<= 28 bytes)
// ahead is always complete, so the decode loop never waits on HBM.  Per symbol the 8 bytes
// below the cursor are assembled from three aligned ring words; the four chains of a quad are
// independent up to the cursor (a plain register add), which gives the lane four table lookups
// to overlap.  Decoded symbols are staged per warp and written out in coalesced 64-byte pieces.
constexpr int kLpbDecWarps = 8;
constexpr int kLpbDecOutRow = kLpbTileQuads * 16 + 16; // 80 B
constexpr int kLpbRing = 128;                           // bytes of ring per lane (power of two, 128-aligned rows)
constexpr int kLpbPrefetch = 80;                        // bytes below the cursor kept in flight: >= 2 * 28 + 15
static_assert(kLpbPrefetch + 15 + 16 <= kLpbRing, "ring too small for the prefetch distance");

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ uint32_t lds32(uint32_t smem_addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_addr) : "memory");
    return v;
}

template <int METHOD, int TABLE>
__global__ void __launch_bounds__(kLpbDecWarps * 32) lpb_decode_kernel(const DecArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    constexpr bool kInt = (METHOD == METHOD_INT);
    const uint32_t frame = 1u << a.log2_frame;
    DecTables<METHOD, TABLE> tables;
    const size_t table_bytes = tables.load(a, smem_raw);
    // rings must be 128-byte aligned (ring | offset addressing); the host adds 128 bytes of slack
    uint8_t* s_rows = smem_raw + table_bytes;
    s_rows += (128u - (uint32_t(__cvta_generic_to_shared(s_rows)) & 127u)) & 127u;

    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    // per warp: 32 rings of 128 B (128-aligned), then 32 output rows
    uint8_t* warp_base = s_rows + size_t(warp) * 32 * (kLpbRing + kLpbDecOutRow);
    const uint32_t ring = uint32_t(__cvta_generic_to_shared(warp_base)) + lane * kLpbRing;
    uint8_t* out_rows = warp_base + 32 * kLpbRing;
    uint8_t* my_out = out_rows + lane * kLpbDecOutRow;
    const uint32_t log2_frame = a.log2_frame;
    const uint32_t frame_mask = frame - 1u;
    const uint64_t lower = uint64_t(kAnsK) << log2_frame;
    const uint32_t rfold_bias = (METHOD == METHOD_RFOLD) ? (1u << (a.fidelity + 7)) : 0u; // ans_reorder_fold.hpp:301
    const uint32_t B = a.block_syms;

    const uint32_t n_tasks = (a.n_blocks + 31u) >> 5;
    for (uint32_t task = blockIdx.x * kLpbDecWarps + warp; task < n_tasks; task += gridDim.x * kLpbDecWarps) {
        const uint32_t blk0 = task * 32u;
        const uint32_t blk = blk0 + lane;
        const bool have_block = blk < a.n_blocks;
        const uint64_t first = uint64_t(blk) * B;
        const uint32_t nb = have_block ? uint32_t(min(uint64_t(B), a.n - first)) : 0u;
        const uint32_t tail = nb & 3u;
        const uint32_t quads = nb >> 2;
        const uint32_t max_quads = __reduce_max_sync(0xffffffffu, quads);

        // 16-byte aligned view of the payload; `pos` counts bytes from pay16, one past the next byte to consume
        const uint8_t* pay = a.payload + (have_block ? a.offsets[blk] : 0);
        const uint32_t delta = uint32_t(reinterpret_cast<uintptr_t>(pay) & 15u);
        const uint8_t* pay16 = pay - delta;
        const uint32_t size = have_block ? a.block_bytes[blk] : 32u;
        int32_t pos = int32_t(delta + size - 32u);

        // init_state x4: state c sits at payload + size - 32 + 8c (ans_fold.hpp:291-295)
        uint64_t x0 = lower, x1 = lower, x2 = lower, x3 = lower;
        if (have_block) {
            const uint32_t* words = reinterpret_cast<const uint32_t*>(pay16);
            const int32_t w = pos >> 2;
            const uint32_t sh = (uint32_t(pos) & 3u) * 8u;
            uint32_t t[9];
#pragma unroll
            for (int i = 0; i < 9; i++)
                t[i] = (i < 8 || sh != 0u) ? __ldg(words + w + i) : 0u;
            uint32_t u[8];
#pragma unroll
            for (int i = 0; i < 8; i++)
                u[i] = __funnelshift_r(t[i], t[i + 1], sh);
            x0 += (uint64_t(u[1]) << 32) | u[0];
            x1 += (uint64_t(u[3]) << 32) | u[2];
            x2 += (uint64_t(u[5]) << 32) | u[4];
            x3 += (uint64_t(u[7]) << 32) | u[6];
        }

        // ring fill: everything from the chunk that holds byte pos-1 down to pos - kLpbPrefetch
        int32_t fetched_lo = (pos + 15) & ~15; // chunks [fetched_lo, top) are in the ring
        auto prefetch = [&](int max_chunks) {
            const int32_t want = max(pos - kLpbPrefetch, 0) & ~15;
#pragma unroll
            for (int i = 0; i < max_chunks; i++) {
                if (have_block && fetched_lo > want) {
                    fetched_lo -= 16;
                    cp_async16(ring + (uint32_t(fetched_lo) & uint32_t(kLpbRing - 1)), pay16 + fetched_lo);
                }
            }
            cp_async_commit();
        };
        prefetch((kLpbPrefetch + 15 + 15) / 16 + 1);

        auto decode_one = [&](uint64_t& x, bool active) -> uint32_t {
            const uint32_t slot = uint32_t(x) & frame_mask;
            uint32_t freq, offset, mapped, k;
            tables.lookup(slot, freq, offset, mapped, k);
            uint64_t nx = uint64_t(freq) * (x >> log2_frame) + offset;
            const bool renorm = active && nx < lower;
            // bytes [pos-8, pos) from the three aligned ring words that cover them
            const uint32_t p1 = uint32_t(pos - 1);
            const uint32_t w0 = lds32(ring | (p1 & uint32_t(kLpbRing - 4)));
            const uint32_t w1 = lds32(ring | ((p1 - 4u) & uint32_t(kLpbRing - 4)));
            const uint32_t sh = (p1 & 3u) * 8u + 8u; // 8, 16, 24, 32
            const uint32_t win_hi = __funnelshift_rc(w1, w0, sh); // the 4 bytes just below the cursor
            if (renorm) // renormalisation word first (ans_fold.hpp:221-224)
                nx = (nx << kRadixLog2) | uint64_t(win_hi);
            uint32_t value = mapped;
            uint32_t c = renorm ? 4u : 0u;
            if (!kInt) { // then k raw low bytes (ans_fold.hpp:135-147)
                const uint32_t w2 = lds32(ring | ((p1 - 8u) & uint32_t(kLpbRing - 4)));
                const uint32_t win_lo = __funnelshift_rc(w2, w1, sh);
                const uint32_t exc_src = renorm ? win_lo : win_hi;
                value += __funnelshift_rc(exc_src, 0u, 32u - 8u * k);
                c += k;
            }
            if (active) {
                x = nx;
                pos -= int32_t(c);
            }
            return value - rfold_bias;
        };

        for (uint32_t q0 = 0; q0 < max_quads; q0 += kLpbTileQuads) {
            __syncwarp();
#pragma unroll
            for (int j = 0; j < kLpbTileQuads; j++) {
                const bool active = (q0 + uint32_t(j)) < quads;
                // keep the ring kLpbPrefetch bytes ahead; everything this quad can touch was committed a quad ago
                prefetch(3);
                cp_async_wait<1>();
                uint4 o;
                // index 4t+g is decoded by state 3-g (ans_fold.hpp:300-305)
                o.x = decode_one(x3, active);
                o.y = decode_one(x2, active);
                o.z = decode_one(x1, active);
                o.w = decode_one(x0, active);
                *reinterpret_cast<uint4*>(my_out + j * 16) = o;
            }
            __syncwarp();
            // coalesced write-out: 8 rows x 64 bytes per instruction
#pragma unroll
            for (int it = 0; it < 4; it++) {
                const uint32_t row = uint32_t(it) * 8u + (lane >> 2);
                const uint32_t chunk = lane & 3u;
                const uint32_t rblk = blk0 + row;
                if (rblk < a.n_blocks) {
                    const uint64_t rfirst = uint64_t(rblk) * B;
                    const uint32_t rnb = uint32_t(min(uint64_t(B), a.n - rfirst));
                    const uint32_t qi = q0 + chunk;
                    if (qi < (rnb >> 2)) {
                        const uint4 val = *reinterpret_cast<const uint4*>(out_rows + row * kLpbDecOutRow + chunk * 16u);
                        reinterpret_cast<uint4*>(a.out + rfirst)[qi] = val;
                    }
                }
            }
        }
        // tail symbols on state 0 (ans_fold.hpp:306-310); only the last block of the input has any
        if (__any_sync(0xffffffffu, tail != 0u)) {
            prefetch(3);
            cp_async_wait<0>();
            for (uint32_t s = 0; s < 3u; s++) {
                const bool active = s < tail;
                const uint32_t v = decode_one(x0, active);
                if (active)
                    a.out[first + 4u * quads + s] = v;
            }
        }
        cp_async_wait<0>();
    }
}

} // namespace b200ans
