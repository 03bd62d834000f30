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
// asynchronous 16-byte global -> shared copies (LDGSTS)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem_ptr, const void* gptr)
{
    const uint32_t s = uint32_t(__cvta_generic_to_shared(smem_ptr));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr int kLpbMaxWarps = 28;   // warps per CTA; one CTA per SM (the host picks the count: one balanced wave of tasks)
constexpr int kLpbTileQuads = 4;   // quads (of 4 symbols) per lane and output tile of the decoder
constexpr int kLpbInQuads = 2;     // quads per lane and input tile of the encoder

// ------------------------------------------------------------------------------------------
// K2-LPB: encoder
// ------------------------------------------------------------------------------------------
// Per warp: two input buffers of 32 rows x 32 B (the next tile is in flight while the current one
// is coded) and 32 output rows.  A quad is coded in four phases so that the lane's four chains
// overlap: (1) fold + table lookups of the 4 symbols, (2) the 4 renormalisation tests, (3) the
// serial emission into the word accumulator, (4) the 4 exact state updates.
constexpr int kLpbInRow = kLpbInQuads * 16 + 16;   // 48 B: 3 x 16 (odd) -> conflict-free 128-bit row reads
constexpr int kLpbDrainQuads = kLpbInQuads;        // the output row is drained after every input tile
constexpr int kLpbOutRow = 112;                    // 12 carried + 28 * 2 quads + 32 states + 4 partial <= 104
static_assert(12 + 28 * kLpbDrainQuads + 36 <= kLpbOutRow && (kLpbOutRow / 16) % 2 == 1, "output row too small");
constexpr int kLpbEncWarpBytes = 32 * (2 * kLpbInRow + kLpbOutRow);
constexpr int kLpbEncMaxSmemEntries = 4096;

// Word-granular emitter: bytes are collected little-endian in `acc` (sh = 8 * pending bytes,
// pending < 4) and complete 32-bit words go to the lane's row in shared memory.
struct LpbEmitter {
    uint32_t acc;
    uint32_t sh;   // 0, 8, 16, 24
    uint32_t* wp;  // next word of the row
    __device__ __forceinline__ void put_bytes(uint32_t v, uint32_t k) // k in 0..3 low bytes of v
    {
        const uint32_t vm = v & ((1u << (8u * k)) - 1u);
        const uint32_t t = acc | (vm << sh);
        const uint32_t hi = __funnelshift_l(vm, 0u, sh); // vm >> (32 - sh), 0 when sh == 0
        const uint32_t nsh = sh + 8u * k;
        const bool full = nsh >= 32u;
        if (full)
            *wp++ = t;
        acc = full ? hi : t;
        sh = nsh & 31u;
    }
    __device__ __forceinline__ void put_word(uint32_t w, bool on)
    {
        const uint32_t t = acc | (w << sh);
        if (on) {
            *wp++ = t;
            acc = __funnelshift_l(w, 0u, sh); // w >> (32 - sh), 0 when sh == 0
        }
    }
};

template <int METHOD, bool TABLE_SMEM, bool FAST>
__global__ void __launch_bounds__(kLpbMaxWarps * 32, 1) lpb_encode_kernel(const EncArgs a)
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
    const uint32_t warps_per_cta = blockDim.x >> 5;
    uint8_t* in_rows = s_rows + size_t(warp) * kLpbEncWarpBytes; // 2 buffers of 32 rows
    uint8_t* out_rows = in_rows + 2 * 32 * kLpbInRow;
    uint8_t* my_out = out_rows + lane * kLpbOutRow;
    const uint4* table4 = reinterpret_cast<const uint4*>(TABLE_SMEM ? s_table : a.table);
    const uint32_t log2_frame = a.log2_frame;
    const uint32_t f = a.fidelity;
    const uint64_t lower = uint64_t(kAnsK) << log2_frame;
    const uint32_t B = a.block_syms;

    const uint32_t n_tasks = (a.n_blocks + 31u) >> 5;
    for (uint32_t task = blockIdx.x * warps_per_cta + warp; task < n_tasks; task += gridDim.x * warps_per_cta) {
        const uint32_t blk0 = task * 32u;
        const uint32_t blk = blk0 + lane;
        const bool have_block = blk < a.n_blocks;
        const uint64_t first = uint64_t(blk) * B;
        const uint32_t nb = have_block ? uint32_t(min(uint64_t(B), a.n - first)) : 0u;
        const uint32_t tail = nb & 3u;
        const uint32_t quads = nb >> 2;
        const uint32_t max_quads = __reduce_max_sync(0xffffffffu, quads);
        uint8_t* dst = a.scratch + uint64_t(blk) * a.stride;

        uint64_t xs[4] = { lower, lower, lower, lower }; // initial_state() (ans_fold.hpp:113)
        LpbEmitter em;
        em.acc = 0;
        em.sh = 0;
        em.wp = reinterpret_cast<uint32_t*>(my_out);
        uint32_t flushed = 0;
        uint32_t bad = 0;

        // symbol -> {k raw bytes, freq, base, reciprocal}
        auto model = [&](uint32_t& v, bool active, uint32_t& k, uint32_t& fr, uint32_t& base, double& rcp) {
            if (METHOD == METHOD_RFOLD) {
                if (v >= a.n_values) { // value the model has never seen
                    bad |= uint32_t(active);
                    v = 0;
                }
                v = __ldg(a.mapping + v);
            }
            uint32_t sym = map_value<METHOD>(v, f, k);
            if (sym >= a.n_entries || (METHOD != METHOD_INT && v >= (1u << 30))) { // (ans_fold.hpp:198-200: values < 2^30)
                bad |= uint32_t(active);
                sym = 0;
            }
            const uint4 raw = TABLE_SMEM ? table4[sym] : __ldg(table4 + sym);
            fr = raw.x;
            base = raw.y;
            rcp = __hiloint2double(int(raw.w), int(raw.z));
            bad |= uint32_t(active && fr == 0u);
        };
        auto push = [&](uint64_t x, bool renorm, uint32_t fr, uint32_t base, double rcp) -> uint64_t {
            const uint32_t x_lo = uint32_t(x), x_hi = uint32_t(x >> 32);
            const uint32_t y_lo = renorm ? x_hi : x_lo;
            const uint32_t y_hi = renorm ? 0u : x_hi;
            if (FAST)
                return ans_push52(y_lo, y_hi, fr, base, rcp, log2_frame);
            return ans_push((uint64_t(y_hi) << 32) | y_lo, fr, base, rcp, log2_frame);
        };
        auto drain = [&](uint32_t max_chunks_hint) {
            const uint32_t fill = uint32_t(reinterpret_cast<uint8_t*>(em.wp) - my_out); // multiple of 4
            const uint32_t chunks = fill >> 4;
            const uint4* rv = reinterpret_cast<const uint4*>(my_out);
            uint4* gv = reinterpret_cast<uint4*>(dst + flushed);
#pragma unroll
            for (uint32_t ci = 0; ci < 5u; ci++)
                if (ci < max_chunks_hint && ci < chunks)
                    gv[ci] = rv[ci];
            if (chunks != 0u) {
                const uint4 carry = rv[chunks];
                *reinterpret_cast<uint4*>(my_out) = carry;
            }
            flushed += chunks << 4;
            em.wp = reinterpret_cast<uint32_t*>(my_out + (fill & 15u));
        };

        // tail symbols (only the very last block of the input has any): state 0 (ans_fold.hpp:257-261)
        if (__any_sync(0xffffffffu, tail != 0u)) {
            for (uint32_t s = 0; s < 3u; s++) {
                const bool active = s < tail;
                uint32_t v = active ? __ldg(a.in + first + (nb - 1u - s)) : 0u;
                uint32_t k, fr, base;
                double rcp;
                model(v, active, k, fr, base, rcp);
                const bool renorm = active && ((uint32_t(xs[0] >> 32) >> 4) >= fr);
                if (METHOD != METHOD_INT)
                    em.put_bytes(v, active ? k : 0u);
                em.put_word(uint32_t(xs[0]), renorm);
                const uint64_t nx = push(xs[0], renorm, fr, base, rcp);
                if (active)
                    xs[0] = nx;
            }
            drain(2);
        }

        // stage the 8 symbols per block of tile q0 into buffer `buf`: coalesced 32-byte pieces, asynchronous
        auto stage = [&](uint32_t q0, uint32_t buf) {
            if (q0 < max_quads) {
#pragma unroll
                for (int it = 0; it < 2; it++) {
                    const uint32_t row = uint32_t(it) * 16u + (lane >> 1);
                    const uint32_t chunk = lane & 1u;
                    const uint32_t rblk = blk0 + row;
                    if (rblk < a.n_blocks) {
                        const uint64_t rfirst = uint64_t(rblk) * B;
                        const uint32_t rnb = uint32_t(min(uint64_t(B), a.n - rfirst));
                        // quad index (ascending in memory) held by this chunk
                        const int32_t qi = int32_t(rnb >> 2) - int32_t(q0) - kLpbInQuads + int32_t(chunk);
                        if (qi >= 0)
                            cp_async16(in_rows + (buf * 32u + row) * kLpbInRow + chunk * 16u,
                                reinterpret_cast<const uint4*>(a.in + rfirst) + qi);
                    }
                }
            }
            cp_async_commit();
        };

        stage(0, 0);
        uint32_t buf = 0;
        for (uint32_t q0 = 0; q0 < max_quads; q0 += kLpbInQuads, buf ^= 1u) {
            stage(q0 + kLpbInQuads, buf ^ 1u);
            cp_async_wait<1>(); // this tile has landed (for the copies this lane issued) ...
            __syncwarp();       // ... and for everybody else's
            const uint8_t* my_in = in_rows + (buf * 32u + lane) * kLpbInRow;
#pragma unroll
            for (int j = 0; j < kLpbInQuads; j++) {
                const bool active = (q0 + uint32_t(j)) < quads;
                const uint4 v4 = *reinterpret_cast<const uint4*>(my_in + (kLpbInQuads - 1 - j) * 16);
                // walking backwards: index 4t+3 (state 0) first ... index 4t (state 3) last (ans_fold.hpp:262-271)
                uint32_t v[4] = { v4.w, v4.z, v4.y, v4.x };
                uint32_t k[4], fr[4], base[4];
                double rcp[4];
                bool rn[4];
#pragma unroll
                for (int c = 0; c < 4; c++)
                    model(v[c], active, k[c], fr[c], base[c], rcp[c]);
#pragma unroll
                for (int c = 0; c < 4; c++) // state >= freq * K * 2^32 (ans_fold.hpp:105)
                    rn[c] = active && ((uint32_t(xs[c] >> 32) >> 4) >= fr[c]);
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (METHOD != METHOD_INT)
                        em.put_bytes(v[c], active ? k[c] : 0u); // raw low bytes first (ans_fold.hpp:59-63)
                    em.put_word(uint32_t(xs[c]), rn[c]);        // then the renormalisation word (ans_fold.hpp:106-109)
                }
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const uint64_t nx = push(xs[c], rn[c], fr[c], base[c], rcp[c]);
                    if (active)
                        xs[c] = nx;
                }
            }
            drain(5);
            __syncwarp(); // every lane is done with this buffer before the next stage overwrites it
        }
        cp_async_wait<0>();

        // flush_state x4 in chain order 0..3 (ans_fold.hpp:115-120, 275-278)
        {
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const uint64_t st = xs[c] - lower;
                em.put_word(uint32_t(st), have_block);
                em.put_word(uint32_t(st >> 32), have_block);
            }
            uint32_t fill = uint32_t(reinterpret_cast<uint8_t*>(em.wp) - my_out);
            const uint32_t total = flushed + fill + (em.sh >> 3);
            // last partial word, then zero padding up to the next 16-byte boundary (deterministic container)
            *em.wp = em.acc;
            fill += 4u;
            while (fill & 15u) {
                *reinterpret_cast<uint32_t*>(my_out + fill) = 0u;
                fill += 4u;
            }
            const uint32_t chunks = (total - flushed + 15u) >> 4;
            const uint4* rv = reinterpret_cast<const uint4*>(my_out);
            uint4* gv = reinterpret_cast<uint4*>(dst + flushed);
#pragma unroll
            for (uint32_t ci = 0; ci < uint32_t(kLpbOutRow / 16); ci++)
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
// (cp.async / LDGSTS) that run kLpbPrefetch bytes ahead of the cursor: what a quad (<= 28 bytes)
// can touch was committed one quad earlier, so the decode loop does not wait on HBM.  A quad is
// decoded in three phases so that the lane's four chains overlap: (1) the 4 slot lookups and
// state updates, which tell how many bytes each symbol consumes, (2) the 4 cursor positions
// (register adds), (3) per symbol the 8 bytes below its cursor, assembled from three aligned ring
// words.  Decoded symbols are staged per warp and written out in coalesced 64-byte pieces.
constexpr int kLpbDecOutRow = kLpbTileQuads * 16 + 16; // 80 B
constexpr int kLpbRing = 128;                           // bytes of ring per lane (power of two, 128-aligned rows)
constexpr int kLpbPrefetch = 80;                        // bytes below the cursor kept in flight: >= 2 * 28 + 15
static_assert(kLpbPrefetch + 15 + 16 <= kLpbRing, "ring too small for the prefetch distance");
constexpr int kLpbDecWarpBytes = 32 * (kLpbRing + kLpbDecOutRow);

template <int METHOD, int TABLE>
__global__ void __launch_bounds__(kLpbMaxWarps * 32, 1) lpb_decode_kernel(const DecArgs a)
{
    if (a.error_flag && *a.error_flag >= 3u) // the size checks in front of this launch failed (check_block_sizes_kernel)
        return;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    constexpr bool kInt = (METHOD == METHOD_INT);
    const uint32_t frame = 1u << a.log2_frame;
    DecTables<METHOD, TABLE> tables;
    const size_t table_bytes = tables.load(a, smem_raw);
    // rings must be 128-byte aligned (ring + (offset & 127) addressing); the host adds 128 bytes of slack
    uint8_t* s_rows = smem_raw + table_bytes;
    s_rows += (128u - (uint32_t(__cvta_generic_to_shared(s_rows)) & 127u)) & 127u;

    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t warps_per_cta = blockDim.x >> 5;
    // per warp: 32 rings of 128 B, then 32 output rows
    uint8_t* warp_base = s_rows + size_t(warp) * kLpbDecWarpBytes;
    uint8_t* ring = warp_base + lane * kLpbRing;
    uint8_t* out_rows = warp_base + 32 * kLpbRing;
    uint8_t* my_out = out_rows + lane * kLpbDecOutRow;
    const uint32_t log2_frame = a.log2_frame;
    const uint32_t frame_mask = frame - 1u;
    const uint64_t lower = uint64_t(kAnsK) << log2_frame;
    const uint32_t rfold_bias = (METHOD == METHOD_RFOLD) ? (1u << (a.fidelity + 7)) : 0u; // ans_reorder_fold.hpp:301
    const uint32_t B = a.block_syms;

    const uint32_t n_tasks = (a.n_blocks + 31u) >> 5;
    for (uint32_t task = blockIdx.x * warps_per_cta + warp; task < n_tasks; task += gridDim.x * warps_per_cta) {
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
        uint64_t xs[4] = { lower, lower, lower, lower };
        if (have_block) {
            const uint32_t* words = reinterpret_cast<const uint32_t*>(pay16);
            const int32_t w = pos >> 2;
            const uint32_t sh = (uint32_t(pos) & 3u) * 8u;
            uint32_t t[9];
#pragma unroll
            for (int i = 0; i < 9; i++)
                t[i] = (i < 8 || sh != 0u) ? __ldg(words + w + i) : 0u;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const uint32_t lo = __funnelshift_r(t[2 * c], t[2 * c + 1], sh);
                const uint32_t hi = __funnelshift_r(t[2 * c + 1], t[2 * c + 2], sh);
                xs[c] += (uint64_t(hi) << 32) | lo;
            }
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

        // Word window: t0 = the aligned ring word that holds byte pos-1, t1 and t2 the two words below it.
        auto ring_word = [&](int32_t word_index) -> uint32_t {
            return *reinterpret_cast<const uint32_t*>(ring + ((uint32_t(word_index) << 2) & uint32_t(kLpbRing - 4)));
        };
        // phase 1 of a symbol: slot lookup + state update; returns the bytes it consumes
        auto advance = [&](uint64_t x, bool active, uint64_t& nx, bool& renorm, uint32_t& mapped, uint32_t& k) -> uint32_t {
            uint32_t freq, offset;
            tables.lookup(uint32_t(x) & frame_mask, freq, offset, mapped, k);
            nx = uint64_t(freq) * (x >> log2_frame) + offset;
            renorm = active && nx < lower;
            return active ? k + (renorm ? 4u : 0u) : 0u;
        };
        // phase 3: renormalisation word first (ans_fold.hpp:221-224), then k raw low bytes (ans_fold.hpp:135-147),
        // both taken from the 8 bytes below cursor p, which the three cached words t0 t1 t2 cover
        auto finish = [&](int32_t p, uint32_t t0, uint32_t t1, uint32_t t2, uint64_t& nx, bool renorm, uint32_t mapped,
                          uint32_t k) -> uint32_t {
            const uint32_t sh = (uint32_t(p - 1) & 3u) * 8u + 8u; // 8, 16, 24, 32
            const uint32_t win_hi = __funnelshift_rc(t1, t0, sh); // the 4 bytes just below the cursor
            if (renorm)
                nx = (nx << kRadixLog2) | uint64_t(win_hi);
            uint32_t value = mapped;
            if (!kInt) {
                const uint32_t win_lo = __funnelshift_rc(t2, t1, sh); // the 4 bytes below those
                const uint32_t exc_src = renorm ? win_lo : win_hi;
                value += __funnelshift_rc(exc_src, 0u, 32u - 8u * k);
            }
            return value - rfold_bias;
        };
        cp_async_wait<0>();
        int32_t wi = (pos - 1) >> 2; // word index of byte pos-1 (pos >= 1: at least the states' first byte is below... )
        uint32_t t0 = ring_word(wi), t1 = ring_word(wi - 1), t2 = ring_word(wi - 2);

        for (uint32_t q0 = 0; q0 < max_quads; q0 += kLpbTileQuads) {
            __syncwarp();
#pragma unroll
            for (int j = 0; j < kLpbTileQuads; j++) {
                const bool active = (q0 + uint32_t(j)) < quads;
                // keep the ring kLpbPrefetch bytes ahead; everything this quad can touch was committed a quad ago
                prefetch(3);
                cp_async_wait<1>();
                uint64_t nx[4];
                bool rn[4];
                uint32_t mp[4], k[4], c[4], val[4];
                int32_t p[5], w[5];
                // index 4t+g is decoded by state 3-g (ans_fold.hpp:300-305): g = 0 reads the stream first
#pragma unroll
                for (int g = 0; g < 4; g++)
                    c[g] = advance(xs[3 - g], active, nx[g], rn[g], mp[g], k[g]);
                p[0] = pos;
                w[0] = wi;
#pragma unroll
                for (int g = 0; g < 4; g++) {
                    p[g + 1] = p[g] - int32_t(c[g]);
                    w[g + 1] = (p[g + 1] - 1) >> 2;
                }
                // every word the four windows will need beyond the cached three: a symbol moves the
                // cursor across at most two word boundaries (7 bytes), so two predicated loads each
                uint32_t ra[4], rb[4];
#pragma unroll
                for (int g = 0; g < 4; g++) {
                    ra[g] = (w[g + 1] < w[g]) ? ring_word(w[g] - 3) : 0u;
                    rb[g] = (!kInt && w[g + 1] < w[g] - 1) ? ring_word(w[g] - 4) : 0u;
                }
#pragma unroll
                for (int g = 0; g < 4; g++) {
                    val[g] = finish(p[g], t0, t1, t2, nx[g], rn[g], mp[g], k[g]);
                    if (active)
                        xs[3 - g] = nx[g];
                    if (w[g + 1] < w[g]) {
                        t0 = t1;
                        t1 = t2;
                        t2 = ra[g];
                    }
                    if (!kInt && w[g + 1] < w[g] - 1) {
                        t0 = t1;
                        t1 = t2;
                        t2 = rb[g];
                    }
                }
                pos = p[4];
                wi = w[4];
                *reinterpret_cast<uint4*>(my_out + j * 16) = make_uint4(val[0], val[1], val[2], val[3]);
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
                uint64_t nx;
                bool rn;
                uint32_t mp, k;
                const uint32_t c = advance(xs[0], active, nx, rn, mp, k);
                t0 = ring_word((pos - 1) >> 2), t1 = ring_word(((pos - 1) >> 2) - 1), t2 = ring_word(((pos - 1) >> 2) - 2);
                const uint32_t v = finish(pos, t0, t1, t2, nx, rn, mp, k);
                pos -= int32_t(c);
                if (active) {
                    xs[0] = nx;
                    a.out[first + 4u * quads + s] = v;
                }
            }
        }
        cp_async_wait<0>();
    }
}

} // namespace b200ans
