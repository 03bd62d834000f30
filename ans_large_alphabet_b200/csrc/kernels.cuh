// b200ans device code (sm_100a): histogram, block-parallel rANS encoder / decoder, size scan,
// payload compaction.
//
// Parallel decomposition ("lane per chain"): the reference coder keeps four interleaved 64-bit
// rANS states per stream (ans_fold.hpp:250-278).  Here the input is cut into independent
// blocks; one *group* of 4 lanes owns one block -- lane g carries chain 3-g, i.e. the symbols
// at block-relative index 4t+g -- and a warp runs 8 blocks side by side.  The four chains are
// arithmetically independent; they meet only at the shared byte cursor, which a 4-lane
// (segmented) scan of "bytes emitted this step" reproduces exactly, so each block's payload is
// byte-identical to the reference coder run on that block.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../host/fold.hpp"

namespace b200ans {

// ------------------------------------------------------------------------------------------
// Tables
// ------------------------------------------------------------------------------------------

// Encoder entry, one per (mapped) symbol.  Replaces enc_entry_fold/_int/_msb/_reorder_fold
// {freq, base, sym_upper_bound} (ans_fold.hpp:30-34 ...): the upper bound is recomputed as
// freq << 36, and `rcp` is the smallest double >= 1 / freq (b200ans.cu reciprocal_up): ans_push52 (lpb_kernels.cuh)
// gets floor(state / freq) exactly from one DFMA.RZ with it; ans_push below multiplies towards zero and fixes up.
struct __align__(16) EncEntry {
    uint32_t freq;
    uint32_t base;
    double rcp;
};

// Decoder slot entry for the folded coders.  Same content as dec_entry_fold
// (ans_fold.hpp:128-132): freq, offset-in-symbol, unfolded base value | nbytes << 30.
struct __align__(8) DecEntry {
    uint16_t freq;
    uint16_t offset;
    uint32_t mapped; // for rfold: most_frequent[s] + 2^(f+7) resp. unfolded value
};

// Plain ANS (ans_int) replaces the reference's M-entry slot tables of 8 / 12 bytes per slot
// (ans_int.hpp:100-110,129-161) by a rank/select index over the frame -- the table *contents*
// are a pure function of nfreqs, so the layout is ours to choose:
//   rec : per 32 slots {bits, cnt}: bit s set iff slot s is the first slot of a symbol with freq > 0,
//         cnt = number of set bits in all earlier words
//   ent : per present symbol, in slot order: {base, freq}
//   sym : per present symbol its id -- only needed when some id below max_sym is absent
// slot -> rank = cnt + popc(bits & ((2 << (s & 31)) - 1)) - 1 -> ent[rank]: two dependent 8-byte loads.
// M / 4 + 8 * sigma bytes: a 2^17-slot model with 1 900 symbols takes 47 KiB and sits in
// one SM's shared memory, where the reference layout (1 MiB) would live in L2.
struct __align__(8) RankRec {
    uint32_t bits;
    uint32_t cnt;
};
struct __align__(8) RankEntry {
    uint32_t base;
    uint32_t freq;
};

struct EncArgs {
    const uint32_t* in;      // symbols
    uint64_t n;              // number of symbols
    uint32_t block_syms;     // symbols per block (multiple of 4), or n for single-block mode
    uint32_t n_blocks;
    const EncEntry* table;   // global copy of the encoder table
    uint32_t n_entries;      // entries in table (alphabet bins for folded methods, max_sym+1 for INT)
    uint32_t log2_frame;     // log2 M
    uint32_t fidelity;
    const uint32_t* mapping; // RFOLD: value -> reordered value (ans_reorder_fold.hpp:92-106)
    uint32_t n_values;       // RFOLD: entries in mapping
    uint8_t* scratch;        // block b's payload goes to scratch + b * stride
    uint64_t stride;
    uint32_t* block_bytes;   // out: payload bytes of block b
    uint32_t* error_flag;    // set to non-zero if a symbol had no model entry
    uint32_t fast_div;       // every frequency < 2^16: states stay below 2^52 (lpb_kernels.cuh ans_push52)
    // pseudo-adaptive mode: block b has its own model -- table + b * pa_table_stride, log2 M = pa_log2[b]
    const uint8_t* pa_log2;
    uint64_t pa_table_stride; // entries
    // pa_kernels.cuh (tables built in shared memory): normalised frequencies of model b at pa_nfreqs + b * pa_nbins
    const uint32_t* pa_nfreqs;
    uint32_t pa_nbins;
    uint32_t pa_syms;         // table entries per model: beyond the largest symbol of the input, multiple of 16
};

struct DecArgs {
    const uint8_t* payload;      // base of the payload area
    const uint64_t* offsets;     // byte offset of block b's payload from `payload`
    const uint32_t* block_bytes; // payload bytes of block b
    uint32_t* out;
    uint64_t n;
    uint32_t block_syms;
    uint32_t n_blocks;
    uint32_t log2_frame;
    uint32_t fidelity;
    const DecEntry* slots;     // folded methods: M entries
    const RankRec* rank_rec;   // INT: M / 32 records
    const uint32_t* rank_sym;  // INT: n_present symbol ids (unused when dense)
    const RankEntry* rank_ent; // INT: n_present entries
    uint32_t n_present;
    uint32_t dense;            // INT: every id 0..max_sym is present, so the rank is the symbol
    // pseudo-adaptive mode: block b decodes with slots + b * pa_slot_stride and log2 M = pa_log2[b]
    const uint8_t* pa_log2;
    uint64_t pa_slot_stride; // entries
    // pa_kernels.cuh (no slot tables): normalised frequencies of model b at pa_nfreqs + b * pa_nbins
    const uint32_t* pa_nfreqs;
    uint32_t pa_nbins;
    uint32_t pa_syms;         // table entries per model: beyond the largest symbol of every model, multiple of 16
    uint32_t pa_bucket_log2;  // the bucket index has 2^pa_bucket_log2 entries per model
    uint32_t* error_flag;     // set to 2 when a block's payload is not consumed exactly (corrupt container); may be null
};

enum dec_table_kind : int { DEC_FOLD_SMEM = 0, DEC_FOLD_GLOBAL = 1, DEC_INT_SMEM = 2, DEC_INT_GLOBAL = 3, DEC_INT_SMEM_PURE = 4, DEC_FOLD_SMEM2 = 5 };

// DEC_FOLD_SMEM2: the M-entry slot table TWICE in shared memory, copy r in the odd / even 8-byte units (entry `slot` of
// copy r at unit 2 * slot + r), lanes with an even lane id reading copy 0 and the others copy 1.  A random 8-byte lookup by
// a half-warp then spreads 8 lanes over 8 bank pairs twice instead of 16 lanes over 16 -- an expected maximum of 2.4
// lanes per pair instead of 3.1, 4.8 shared-memory wavefronts per lookup instead of 6.2 -- with the same instructions (the
// copy is chosen by the lane's base pointer, the index scales by 16 bytes instead of 8).

// DEC_INT_SMEM_PURE: the same index, but a 32-slot group that lies inside ONE symbol (no symbol starts at its slots
// 1..31) carries that symbol's answer in its own record -- x = base | rank<11:0> << 20, y = 1 << 31 | rank<22:12> << 20 |
// freq -- so only the lanes whose slot falls into a mixed group take the second lookup (geometric(0.01) at M = 2^17:
// 11 % of them).  The decoder of a large input is bound by shared-memory wavefronts, and a lookup by 3 lanes costs one
// where a lookup by 32 costs six.  Needs M <= 2^19 and fewer than 2^23 symbols (the host checks).
constexpr uint32_t kRankPureFlag = 0x80000000u;

// Decoder tables as the kernels see them: global pointers, or a copy in shared memory.
// The rank/select layouts (DEC_INT_*) also serve the FOLDED coders when their frame is too large for an M-entry slot
// table in shared memory (M = 2^18 for a 2^32-symbol input: adjust_freqs keeps doubling until the largest frequency
// nears 16 bits): `sym` then holds, per present symbol, the slot table's packed word (unfolded base value | k << 30).
template <int METHOD, int TABLE> struct DecTables {
    static constexpr bool kPure = (TABLE == DEC_INT_SMEM_PURE);
    static constexpr bool kInt = (TABLE == DEC_INT_SMEM || TABLE == DEC_INT_GLOBAL || kPure); // rank/select lookup
    static constexpr bool kPacked = kInt && METHOD != METHOD_INT;                             // sym[] = value | k << 30
    static constexpr bool kTwice = (TABLE == DEC_FOLD_SMEM2);
    static constexpr bool kSmem = (TABLE == DEC_FOLD_SMEM || TABLE == DEC_INT_SMEM || kPure || kTwice);
    const DecEntry* slots;
    const RankRec* rec;
    const uint32_t* sym;
    const RankEntry* ent;
    uint32_t dense;

    __host__ __device__ static size_t smem_bytes(uint32_t frame, uint32_t n_present, uint32_t dense)
    {
        if (!kSmem)
            return 0;
        if (kInt) {
            const size_t words = (size_t(frame) + 31) / 32;
            return ((words * 8 + 15) & ~size_t(15)) + ((size_t(n_present) * 8 + 15) & ~size_t(15))
                + (dense ? 0 : ((size_t(n_present) * 4 + 15) & ~size_t(15)));
        }
        return (size_t(frame) * sizeof(DecEntry) * (kTwice ? 2 : 1) + 15) & ~size_t(15);
    }

    // cooperative copy into shared memory (all threads of the CTA); returns the bytes used
    __device__ __forceinline__ size_t load(const DecArgs& a, uint8_t* smem)
    {
        slots = a.slots;
        rec = a.rank_rec;
        sym = a.rank_sym;
        ent = a.rank_ent;
        dense = a.dense;
        if (!kSmem)
            return 0;
        const uint32_t frame = 1u << a.log2_frame;
        if (kInt) {
            const size_t words = (size_t(frame) + 31) / 32;
            const size_t rec_b = (words * 8 + 15) & ~size_t(15);
            const size_t ent_b = (size_t(a.n_present) * 8 + 15) & ~size_t(15);
            const uint2* s0 = reinterpret_cast<const uint2*>(a.rank_rec);
            uint2* d0 = reinterpret_cast<uint2*>(smem);
            for (size_t i = threadIdx.x; i < words; i += blockDim.x)
                d0[i] = __ldg(s0 + i);
            const uint2* s1 = reinterpret_cast<const uint2*>(a.rank_ent);
            uint2* d1 = reinterpret_cast<uint2*>(smem + rec_b);
            for (size_t i = threadIdx.x; i < a.n_present; i += blockDim.x)
                d1[i] = __ldg(s1 + i);
            rec = reinterpret_cast<const RankRec*>(smem);
            ent = reinterpret_cast<const RankEntry*>(smem + rec_b);
            if (kPacked || !a.dense) {
                uint32_t* d2 = reinterpret_cast<uint32_t*>(smem + rec_b + ent_b);
                for (size_t i = threadIdx.x; i < a.n_present; i += blockDim.x)
                    d2[i] = __ldg(a.rank_sym + i);
                sym = d2;
            }
        } else {
            const uint2* src = reinterpret_cast<const uint2*>(a.slots);
            uint2* dst = reinterpret_cast<uint2*>(smem);
            if (kTwice) {
                for (size_t i = threadIdx.x; i < frame; i += blockDim.x) {
                    const uint2 e = __ldg(src + i);
                    *reinterpret_cast<uint4*>(dst + 2 * i) = make_uint4(e.x, e.y, e.x, e.y);
                }
                slots = reinterpret_cast<const DecEntry*>(smem) + (threadIdx.x & 1u);
            } else {
                for (size_t i = threadIdx.x; i < frame; i += blockDim.x)
                    dst[i] = __ldg(src + i);
                slots = reinterpret_cast<const DecEntry*>(smem);
            }
        }
        __syncthreads();
        return smem_bytes(frame, a.n_present, a.dense);
    }

    // as lookup(), but the folded coders get the packed word (unfolded base value | k << 30) back
    __device__ __forceinline__ void lookup_raw(uint32_t slot, uint32_t& freq, uint32_t& offset, uint32_t& raw, uint32_t& k) const
    {
        if (kInt) {
            lookup(slot, freq, offset, raw, k);
            if (kPacked)
                raw |= k << 30;
        } else {
            const uint2 e = kTwice ? *reinterpret_cast<const uint2*>(slots + 2u * slot)
                : kSmem            ? *reinterpret_cast<const uint2*>(slots + slot)
                                   : __ldg(reinterpret_cast<const uint2*>(slots + slot));
            freq = e.x & 0xFFFFu;
            offset = e.x >> 16;
            raw = e.y;
            k = e.y >> 30;
        }
    }

    // slot -> {freq, offset inside the symbol's slot range, unfolded base value / symbol id, raw byte count}
    __device__ __forceinline__ void lookup(uint32_t slot, uint32_t& freq, uint32_t& offset, uint32_t& mapped, uint32_t& k) const
    {
        if (kInt) {
            const uint32_t w = slot >> 5;
            const uint2 rc = kSmem ? *reinterpret_cast<const uint2*>(rec + w) : __ldg(reinterpret_cast<const uint2*>(rec + w));
            uint32_t r = rc.y + uint32_t(__popc(rc.x & (0xFFFFFFFFu >> (31u - (slot & 31u))))) - 1u;
            uint2 e;
            if (kPure) {
                const bool pure = int32_t(rc.y) < 0;
                if (pure)
                    r = (rc.x >> 20) | ((rc.y >> 8) & 0x7FF000u);
                e = make_uint2(rc.x & 0xFFFFFu, rc.y & 0xFFFFFu);
                if (!pure)
                    e = *reinterpret_cast<const uint2*>(ent + r);
            } else {
                e = kSmem ? *reinterpret_cast<const uint2*>(ent + r) : __ldg(reinterpret_cast<const uint2*>(ent + r));
            }
            freq = e.y;
            offset = slot - e.x;
            mapped = r;
            if (kPacked || !dense)
                mapped = kSmem ? sym[r] : __ldg(sym + r);
            k = 0;
            if (kPacked) {
                k = mapped >> 30;
                mapped &= 0x3FFFFFFFu;
            }
        } else {
            const uint2 e = kTwice ? *reinterpret_cast<const uint2*>(slots + 2u * slot)
                : kSmem            ? *reinterpret_cast<const uint2*>(slots + slot)
                                   : __ldg(reinterpret_cast<const uint2*>(slots + slot));
            freq = e.x & 0xFFFFu;
            offset = e.x >> 16;
            mapped = e.y & 0x3FFFFFFFu;
            k = e.y >> 30;
        }
    }
};

// K3a, the decoder's table fill (ans_fold.hpp:192-204, ans_msb.hpp:189-201, ans_reorder_fold.hpp:262-276): the host sends
// one {first slot, freq, packed word} per present symbol, a thread per frame slot finds its symbol by binary search.
struct SlotSym {
    uint32_t base;   // first slot of the symbol
    uint32_t freq;
    uint32_t packed; // unfolded base value | raw bytes << 30 (rfold: most_frequent[s] + 2^(f+7) for the reordered values)
};
__global__ void __launch_bounds__(256) dec_slots_fill_kernel(const SlotSym* __restrict__ syms, uint32_t n_present, uint32_t frame,
    DecEntry* __restrict__ slots)
{
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < frame; slot += gridDim.x * blockDim.x) {
        uint32_t lo = 0, hi = n_present; // syms[lo].base <= slot < syms[hi].base (syms[n_present].base = frame)
        while (hi - lo > 1u) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(&syms[mid].base) <= slot)
                lo = mid;
            else
                hi = mid;
        }
        const uint32_t base = __ldg(&syms[lo].base), freq = __ldg(&syms[lo].freq);
        slots[slot] = DecEntry { uint16_t(freq), uint16_t(slot - base), __ldg(&syms[lo].packed) };
    }
}

// ------------------------------------------------------------------------------------------
// Small helpers
// ------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

template <int METHOD> __device__ __forceinline__ uint32_t map_value(uint32_t v, uint32_t f, uint32_t& k)
{
    if (METHOD == METHOD_INT) {
        k = 0;
        return v;
    } else if (METHOD == METHOD_MSB) {
        return msb_map(v, k);
    } else {
        return fold_map(v, f, k);
    }
}

// Inclusive scan and total of c = k + 4 * r (k = 0..3 raw bytes, r = renormalisation) over the 4 lanes of a group,
// from three ballots instead of three dependent shuffles: the group's nibbles are spread to one bit per unit of
// weight (k bit 0 once, k bit 1 twice, r four times), masked with the lanes that count, and counted.
// `mine` = that 28-bit spread of {lanes of my group at or before me in emission order} (a per-lane constant).
__device__ __forceinline__ uint32_t group_scan_mask(uint32_t g, bool forward)
{
    const uint32_t lanes = forward ? (0xFu >> (3u - g)) : ((0xFu << g) & 0xFu); // forward: g' <= g; reverse: g' >= g
    return lanes * 0x1111111u;
}
__device__ __forceinline__ void group_scan(uint32_t k, bool r, uint32_t gshift, uint32_t mine, uint32_t& incl, uint32_t& total)
{
    const uint32_t b0 = __ballot_sync(0xffffffffu, (k & 1u) != 0u);
    const uint32_t b1 = __ballot_sync(0xffffffffu, (k & 2u) != 0u);
    const uint32_t br = __ballot_sync(0xffffffffu, r);
    const uint32_t v = ((b0 >> gshift) & 0xFu) + ((b1 >> gshift) & 0xFu) * 0x110u + ((br >> gshift) & 0xFu) * 0x1111000u;
    incl = uint32_t(__popc(v & mine));
    total = uint32_t(__popc(v));
}

// state -> (state / freq) * M + state % freq + base, exactly (ans_fold.hpp:111).
// Precondition: state < freq * 2^36 (always true after the renormalisation check) so the
// quotient is below 2^36 and the double product errs by less than 2^-13.
__device__ __forceinline__ uint64_t ans_push(uint64_t x, uint32_t freq, uint32_t base, double rcp, uint32_t log2_frame)
{
    const double qd = __dmul_rz(__ull2double_rz(x), rcp);
    uint64_t q = __double2ull_rz(qd);
    uint32_t r = uint32_t(x) - uint32_t(q) * freq; // true remainder or remainder + freq; both < 2^32
    if (r >= freq) {
        r -= freq;
        q += 1;
    }
    return (q << log2_frame) + uint64_t(r + base);
}

// ------------------------------------------------------------------------------------------
// K1: histogram (+ running maximum of the raw values)
// ------------------------------------------------------------------------------------------
// Replaces the counting loops of ans_fold.hpp:74-78, ans_msb.hpp:81-85, ans_int.hpp:42-49 and
// ans_reorder_fold.hpp:75-84,110-115.
//
// Shared-memory counting with lane-striped replicas: a CTA keeps R copies of the (folded)
// alphabet and lane l counts into copy l % R, each copy shifted by one bank.  Lanes of a warp
// that hold the SAME symbol therefore aggregate over R different banks instead of serialising
// on one address (a Zipf head symbol owns 2..12 lanes of every warp), and lanes with different
// symbols conflict no more than random bank traffic does.  (A __match_any_sync leader election
// was measured first: 54 issue cycles per warp-level count, 10 % of the HBM roofline.)
// Each CTA then adds its non-zero bins, summed over the replicas, to the global 64-bit bins.
//
// HEAD > 0 (raw alphabets that do not fit): values below HEAD are counted in shared memory as
// above (the hot head of a skewed distribution), the rest goes straight to global atomics.

constexpr int kHistThreads = 512;
constexpr uint32_t kHistSmemWords = 12288; // 48 KiB of counters per CTA
constexpr uint32_t kHistHeadBins = 8192;   // privatised head of a large raw alphabet

struct HistArgs {
    const uint32_t* in;
    uint64_t n;
    uint32_t fidelity;
    const uint32_t* mapping; // RFOLD: value -> reordered value
    uint32_t n_values;
    unsigned long long* bins;
    uint32_t nbins;
    uint32_t smem_bins; // bins counted in shared memory (== nbins, or the head of a raw alphabet)
    uint32_t replicas;  // power of two
    uint32_t stride;    // words between replicas (odd multiple: +1 bank per replica)
    uint32_t* max_value;
    unsigned int* bins32; // BINS32: 32-bit global bins (raw alphabets of inputs below 2^32 symbols: half the L2 footprint)
};

__device__ __forceinline__ void red_shared_inc(uint32_t addr)
{
    asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(addr) : "memory");
}

// ALL_SMEM: every bin has a shared-memory counter (all folded alphabets up to fidelity 4, small raw alphabets).
// Then counting is branch-free: bin address straight from the bit position of the value (as in lpb3_kernels.cuh
// l3_entry_addr), one `red.shared` on a 32-bit shared address.
// BINS32 (raw alphabets beyond the shared-memory head, ANSrfold's pre-pass): the counters that take the global atomics
// are 32 bits wide -- 64 MB instead of 128 MB for config 3's 2^24 values, so they stay in the 126 MB L2 while the
// input streams past them (loads marked evict-first); with 64-bit bins every cold count is a DRAM read-modify-write.
template <int METHOD, bool ALL_SMEM, bool BINS32 = false>
__global__ void __launch_bounds__(kHistThreads) histogram_kernel(const HistArgs a)
{
    extern __shared__ uint32_t s_bins[];
    const uint32_t total_words = a.replicas * a.stride;
    for (uint32_t i = threadIdx.x; i < total_words; i += blockDim.x)
        s_bins[i] = 0;
    __syncthreads();
    uint32_t* mine = s_bins + (lane_id() & (a.replicas - 1u)) * a.stride;
    const uint32_t mine_s = uint32_t(__cvta_generic_to_shared(mine));
    const uint32_t smem_bins = a.smem_bins, nbins = a.nbins, f = a.fidelity;
    const uint32_t step = METHOD == METHOD_MSB ? 256u : (255u << (f - 1u)); // symbols per extra raw byte
    const uint32_t fbit = METHOD == METHOD_MSB ? 1u : (1u << (f - 1u));
    uint32_t local_max = 0;
    auto count = [&](uint32_t v) {
        uint32_t mv = v;
        if (METHOD == METHOD_RFOLD)
            mv = v < a.n_values ? __ldg(a.mapping + v) : 0u;
        if (ALL_SMEM) {
            if (METHOD == METHOD_INT) {
                if (mv < nbins)
                    red_shared_inc(mine_s + (mv << 2));
            } else {
                // k = raw low bytes = (highest bit above the threshold) / 8, at most 3: the bin is always in range
                uint32_t fl;
                if (METHOD == METHOD_MSB)
                    fl = 31u - __clz((max(mv, 1u) - 1u) | 1u);
                else
                    fl = 31u - __clz(mv | fbit) - (f - 1u);
                const uint32_t k8 = fl & 0x18u;
                // 4 * (xs + k * step) = (8 * xs + k8 * step) / 2
                red_shared_inc(mine_s + ((((mv >> k8) << 3) + k8 * step) >> 1));
            }
        } else {
            uint32_t k;
            const uint32_t bin = map_value<METHOD>(mv, f, k);
            if (bin < smem_bins)
                atomicAdd(mine + bin, 1u);
            else if (bin < nbins) {
                if (BINS32)
                    atomicAdd(a.bins32 + bin, 1u);
                else
                    atomicAdd(a.bins + bin, 1ull);
            }
        }
        local_max = max(local_max, v);
    };

    // 128-bit loads over the 16-byte aligned body, two in flight per thread; scalar tail
    const uint64_t n4 = a.n >> 2;
    const uint4* in4 = reinterpret_cast<const uint4*>(a.in);
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    for (; i + stride < n4; i += 2 * stride) {
        const uint4 q0 = BINS32 ? __ldcs(in4 + i) : __ldg(in4 + i);
        const uint4 q1 = BINS32 ? __ldcs(in4 + i + stride) : __ldg(in4 + i + stride);
        count(q0.x), count(q0.y), count(q0.z), count(q0.w);
        count(q1.x), count(q1.y), count(q1.z), count(q1.w);
    }
    if (i < n4) {
        const uint4 q0 = __ldg(in4 + i);
        count(q0.x), count(q0.y), count(q0.z), count(q0.w);
    }
    if (blockIdx.x == 0 && threadIdx.x < 3) {
        const uint64_t t = (n4 << 2) + threadIdx.x;
        if (t < a.n)
            count(a.in[t]);
    }
    // maximum of the raw values
    for (int o = 16; o > 0; o >>= 1)
        local_max = max(local_max, __shfl_xor_sync(0xffffffffu, local_max, o));
    if (lane_id() == 0 && local_max != 0)
        atomicMax(a.max_value, local_max);
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < smem_bins; b += blockDim.x) {
        uint32_t c = 0;
        for (uint32_t r = 0; r < a.replicas; r++)
            c += s_bins[r * a.stride + b];
        if (c) {
            if (BINS32)
                atomicAdd(a.bins32 + b, c);
            else
                atomicAdd(a.bins + b, (unsigned long long)c);
        }
    }
}

// 32-bit counters -> 64-bit bins (added: the caller zeroes)
__global__ void __launch_bounds__(256) widen_bins_kernel(const unsigned int* __restrict__ in, unsigned long long* __restrict__ out, uint32_t n)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        out[i] = in[i];
}

// ------------------------------------------------------------------------------------------
// K2: block encoder
// ------------------------------------------------------------------------------------------
// One warp = 8 blocks x 4 chains.  Per step every lane takes one symbol:
//   fold it (k raw bytes), look up {freq, base, rcp}, decide renormalisation, and announce
//   c = k + 4*renorm bytes.  A reverse 4-lane scan (chain 0 emits first, ans_fold.hpp:262-271)
//   gives the lane its byte offset behind the block's cursor; it stores its raw bytes, then its
//   32-bit word (ans_fold.hpp:102-110) into the block's shared-memory row, and updates its state.
// Every kEncTile steps the rows are drained to global memory in aligned 16-byte chunks.

constexpr int kEncWarps = 8;           // warps per CTA
constexpr int kEncTile = 8;            // steps between drains
constexpr int kEncRowBytes = 272;      // 15 carried + 28 * kEncTile + 32 (states) <= 271; 272 = 17 * 16 spreads banks
constexpr int kEncMaxSmemEntries = 6144; // 96 KiB of EncEntry

template <int METHOD, bool TABLE_SMEM, bool PA = false>
__global__ void __launch_bounds__(kEncWarps * 32) encode_kernel(const EncArgs a)
{
    static_assert(!(PA && TABLE_SMEM), "per-block models live in global memory");
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
    const uint32_t grp = lane >> 2;
    const uint32_t g = lane & 3u;           // position inside a 4-symbol step; chain = 3 - g
    const uint32_t grp_base = lane & ~3u;
    const uint32_t scan_mine = group_scan_mask(g, false);
    uint8_t* row = s_rows + (size_t(warp) * 8 + grp) * kEncRowBytes;
    const EncEntry* table0 = TABLE_SMEM ? s_table : a.table;
    const uint32_t f = a.fidelity;

    const uint32_t n_tasks = (a.n_blocks + 7u) >> 3;
    for (uint32_t task = blockIdx.x * kEncWarps + warp; task < n_tasks; task += gridDim.x * kEncWarps) {
        const uint32_t blk = task * 8 + grp;
        const bool have_block = blk < a.n_blocks;
        // the model: one for all blocks, or (pseudo-adaptive) the block's own
        const EncEntry* table = table0;
        uint32_t log2_frame = a.log2_frame;
        if (PA) {
            const uint32_t m = min(blk, a.n_blocks - 1u);
            table = a.table + uint64_t(m) * a.pa_table_stride;
            log2_frame = a.pa_log2[m];
        }
        const uint64_t lower = uint64_t(kAnsK) << log2_frame; // L = K * M (ans_fold.hpp:91)
        const uint64_t first = uint64_t(blk) * a.block_syms;
        const uint32_t nb = have_block ? uint32_t(min(uint64_t(a.block_syms), a.n - first)) : 0u;
        const uint32_t* src = a.in + first;
        const uint32_t tail = nb & 3u;      // these go to chain 0 first (ans_fold.hpp:257-261)
        const uint32_t quads = nb >> 2;
        const uint32_t my_steps = quads + tail;
        const uint32_t max_steps = __reduce_max_sync(0xffffffffu, my_steps);

        uint64_t x = lower;    // initial_state() (ans_fold.hpp:113)
        uint32_t fill = 0;     // bytes sitting in the row
        uint64_t flushed = 0;  // bytes already drained to scratch
        uint8_t* dst = a.scratch + uint64_t(blk) * a.stride;

        // symbol for (step s, this lane) or "inactive"
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
            v = 0;
            if (active) {
                v = __ldg(src + idx);
                if (METHOD == METHOD_RFOLD) {
                    if (v >= a.n_values) { // value the model has never seen
                        *a.error_flag = 1u;
                        v = 0;
                    }
                    v = __ldg(a.mapping + v);
                }
            }
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

#pragma unroll
            for (int i = 0; i < kEncTile; i++) {
                const bool active = (acur >> i) & 1u;
                const uint32_t v = vcur[i];
                uint32_t k;
                uint32_t sym = map_value<METHOD>(v, f, k);
                // value outside the model, or (folded coders) beyond 2^30 - 1, where the decoder's packed word has no
                // room for it (ans_fold.hpp:198-200): flag, stay in bounds
                if (sym >= a.n_entries || (METHOD != METHOD_INT && v >= (1u << 30))) {
                    if (active)
                        *a.error_flag = 1u;
                    sym = 0;
                }
                const EncEntry e = table[sym];
                const bool renorm = active && (uint32_t(x >> 36) >= e.freq); // state >= freq * K * 2^32
                const uint32_t c = active ? k + (renorm ? 4u : 0u) : 0u;
                // reverse inclusive scan over the group: lanes with larger g emit first
                uint32_t incl, total;
                group_scan(active ? k : 0u, renorm, grp_base, scan_mine, incl, total);
                uint8_t* p = row + fill + (incl - c);
                if (active) {
                    if (k > 0u)
                        p[0] = uint8_t(v);
                    if (k > 1u)
                        p[1] = uint8_t(v >> 8);
                    if (k > 2u)
                        p[2] = uint8_t(v >> 16);
                    if (renorm) {
                        const uint32_t w = uint32_t(x);
                        p[k] = uint8_t(w);
                        p[k + 1] = uint8_t(w >> 8);
                        p[k + 2] = uint8_t(w >> 16);
                        p[k + 3] = uint8_t(w >> 24);
                        x >>= kRadixLog2;
                    }
                    if (e.freq == 0u)
                        *a.error_flag = 1u;
                    x = ans_push(x, e.freq, e.base, e.rcp, log2_frame);
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
        // zero the padding up to the next 16-byte boundary so the container is deterministic
        for (uint32_t z = fill + g; z < ((fill + 15u) & ~15u); z += 4)
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
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------
// K3: block decoder
// ------------------------------------------------------------------------------------------
// Mirror image: lane g of a group decodes the symbols at block index 4t+g with chain 3-g
// (ans_fold.hpp:300-305).  Each step: slot lookup, state update, renormalisation test, forward
// 4-lane scan of bytes consumed (renorm word first, then k raw bytes: ans_fold.hpp:221-226),
// then an unaligned 8-byte window below the lane's position is assembled from the block's
// staged bytes in shared memory.  Every kDecTile steps the stage is refilled with the next
// 28*kDecTile (worst case) bytes below the cursor, in aligned 16-byte chunks.

constexpr int kDecWarps = 8;
constexpr int kDecTile = 8;
constexpr int kDecPad = 16;        // bytes in front of the staged window (window under-read lands here)
constexpr int kDecStageChunks = 18; // 16-byte chunks staged per refill: covers 28*8 + 12 + 4 + alignment
constexpr int kDecRowBytes = kDecPad + kDecStageChunks * 16; // 304 = 19 * 16

template <int METHOD, int TABLE, bool PA = false>
__global__ void __launch_bounds__(kDecWarps * 32) decode_kernel(const DecArgs a)
{
    if (a.error_flag && *a.error_flag >= 3u) // the size checks in front of this launch failed (check_block_sizes_kernel)
        return;
    static_assert(!PA || TABLE == DEC_FOLD_GLOBAL, "per-block models: folded slot tables in global memory");
    extern __shared__ __align__(16) uint8_t smem_raw[];
    DecTables<METHOD, TABLE> tables;
    uint8_t* s_rows = smem_raw + tables.load(a, smem_raw);

    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t grp = lane >> 2;
    const uint32_t g = lane & 3u;
    const uint32_t grp_base = lane & ~3u;
    const uint32_t scan_mine = group_scan_mask(g, true);
    uint8_t* row = s_rows + (size_t(warp) * 8 + grp) * kDecRowBytes;
    const uint32_t rfold_bias = (METHOD == METHOD_RFOLD) ? (1u << (a.fidelity + 7)) : 0u; // ans_reorder_fold.hpp:301

    const uint32_t n_tasks = (a.n_blocks + 7u) >> 3;
    for (uint32_t task = blockIdx.x * kDecWarps + warp; task < n_tasks; task += gridDim.x * kDecWarps) {
        const uint32_t blk = task * 8 + grp;
        const bool have_block = blk < a.n_blocks;
        uint32_t log2_frame = a.log2_frame;
        if (PA) {
            const uint32_t m = min(blk, a.n_blocks - 1u);
            tables.slots = a.slots + uint64_t(m) * a.pa_slot_stride;
            log2_frame = a.pa_log2[m];
        }
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
        bool underrun = false;

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
                uint32_t freq, offset, mapped, k;
                tables.lookup(slot, freq, offset, mapped, k);
                uint64_t nx = uint64_t(freq) * (x >> log2_frame) + offset;
                const bool renorm = active && nx < lower;
                const uint32_t c = active ? k + (renorm ? 4u : 0u) : 0u;
                // forward inclusive scan: lane 0 of the group consumes first
                uint32_t incl, total;
                group_scan(active ? k : 0u, renorm, grp_base, scan_mine, incl, total);
                // (a corrupt payload can ask for more bytes than the block has: positions stop at 0, the row's front
                // padding takes the window's under-read, the block is flagged below)
                const int32_t end = max(cursor - int32_t(incl - c), 0); // my bytes are [end - c, end)
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
                    out[idx] = mapped + exc - rfold_bias;
                }
                underrun |= cursor < int32_t(total);
                cursor = max(cursor - int32_t(total), 0);
            }
        }
        // a well-formed block is consumed down to its first byte, exactly
        if (have_block && a.error_flag && (underrun || cursor != int32_t(delta)))
            *a.error_flag = 2u;
    }
}

// ------------------------------------------------------------------------------------------
// K3 for plain ANS (ans_int.hpp:257-306), word-aligned payloads: the same 4-lanes-per-block decomposition as
// decode_kernel, stripped to what a stream without raw bytes needs.  A block's payload is then a sequence of aligned
// 32-bit renormalisation words (every container block starts 16-byte aligned), so a step is: rank/select lookup,
// state update, ONE ballot for the group's renormalisation flags (the lane's word is the popc(flags below me)-th
// below the cursor), one aligned word from the staged window.  About half the instructions of the generic kernel
// (no 3-word window + funnel shifts, no raw-byte counts); with 4x the warps of the lane-per-block decoder it is the
// faster one while the input has fewer than ~112 blocks per SM (C2: 24 415 blocks of 4096 symbols).
// ------------------------------------------------------------------------------------------
constexpr int kIntDecTile = 16; // steps per staged window: 16 steps x 4 lanes x 4 bytes = 256 bytes + alignment
// (Measured and dropped: a 128-word ring per block refilled one tile ahead with cp.async -- the extra 26 KB per CTA
// cost a resident CTA per SM and the per-lane 16-byte copies more than the hidden latency gave back: C2 0.41 -> 0.43 ms.)

template <int TABLE>
__global__ void __launch_bounds__(kDecWarps * 32) int_decode_kernel(const DecArgs a)
{
    if (a.error_flag && *a.error_flag >= 3u) // the size checks in front of this launch failed (check_block_sizes_kernel)
        return;
    static_assert(kDecPad + (kIntDecTile + 1) * 16 <= kDecRowBytes, "staged window does not fit the row");
    extern __shared__ __align__(16) uint8_t smem_raw[];
    DecTables<METHOD_INT, TABLE> tables;
    uint8_t* s_rows = smem_raw + tables.load(a, smem_raw);

    const uint32_t lane = lane_id();
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t grp = lane >> 2;
    const uint32_t g = lane & 3u;
    const uint32_t grp_base = lane & ~3u;
    const uint32_t below = (1u << g) - 1u; // lanes of my group that consume before me (ans_int.hpp:273-289 order)
    uint8_t* row = s_rows + (size_t(warp) * 8 + grp) * kDecRowBytes;
    const uint32_t log2_frame = a.log2_frame;
    const uint32_t frame_mask = (1u << log2_frame) - 1u;
    const uint64_t lower = uint64_t(kAnsK) << log2_frame;

    const uint32_t n_tasks = (a.n_blocks + 7u) >> 3;
    for (uint32_t task = blockIdx.x * kDecWarps + warp; task < n_tasks; task += gridDim.x * kDecWarps) {
        const uint32_t blk = task * 8 + grp;
        const bool have_block = blk < a.n_blocks;
        const uint64_t first = uint64_t(blk) * a.block_syms;
        const uint32_t nb = have_block ? uint32_t(min(uint64_t(a.block_syms), a.n - first)) : 0u;
        const uint32_t tail = nb & 3u;
        const uint32_t quads = nb >> 2;
        const uint32_t my_steps = quads + tail;
        const uint32_t max_steps = __reduce_max_sync(0xffffffffu, my_steps);
        uint32_t* out = a.out + first;

        // word positions are relative to the 16-byte aligned address at or below the payload start
        const uint8_t* pay = a.payload + (have_block ? a.offsets[blk] : 0);
        const uint32_t delta = uint32_t(reinterpret_cast<uintptr_t>(pay) & 15u);
        const uint4* pay_al = reinterpret_cast<const uint4*>(pay - delta);
        const uint32_t size = have_block ? a.block_bytes[blk] : 32u;

        // init_state x4: chain 3 is the last 8 bytes ... chain 0 the first of the 32 (ans_int.hpp:262-270)
        uint64_t x = lower;
        if (have_block) {
            const uint32_t* sp = reinterpret_cast<const uint32_t*>(pay + size - 32u + 8u * (3u - g));
            x = (uint64_t(__ldg(sp)) | (uint64_t(__ldg(sp + 1)) << 32)) + lower;
        }
        int32_t wcur = int32_t((delta + size - 32u) >> 2); // one past the next word to consume
        bool underrun = false;

        for (uint32_t s0 = 0; s0 < max_steps; s0 += kIntDecTile) {
            // stage the words [lo, wcur): at most 4 * kIntDecTile are consumed before the next refill
            int32_t lo = wcur - 4 * kIntDecTile;
            lo = lo < 0 ? 0 : (lo & ~3);
            __syncwarp();
            if (have_block) {
                const uint4* gsrc = pay_al + (lo >> 2);
                uint4* sdst = reinterpret_cast<uint4*>(row + kDecPad);
                const int32_t need = (wcur - lo + 3) >> 2; // chunks that can be touched: <= kIntDecTile + 1
#pragma unroll
                for (int ci = 0; ci < (kIntDecTile + 1 + 3) / 4; ci++) {
                    const int32_t c = ci * 4 + int32_t(g);
                    if (c < need)
                        sdst[c] = __ldg(gsrc + c);
                }
            }
            __syncwarp();
            const uint32_t* stage = reinterpret_cast<const uint32_t*>(row + kDecPad) - lo; // stage[w] is payload word w

#pragma unroll
            for (int i = 0; i < kIntDecTile; i++) {
                const uint32_t s = s0 + uint32_t(i);
                // quads first, all four lanes; then the tail on chain 0 (lane g == 3)
                const bool active = s < quads ? true : (s < my_steps && g == 3u);
                const uint32_t idx = s < quads ? 4u * s + g : s + 3u * quads;
                const uint32_t slot = uint32_t(x) & frame_mask;
                uint32_t freq, offset, sym, k;
                tables.lookup(slot, freq, offset, sym, k);
                uint64_t nx = uint64_t(freq) * (x >> log2_frame) + offset;
                const bool renorm = active && nx < lower;
                const uint32_t flags = (__ballot_sync(0xffffffffu, renorm) >> grp_base) & 0xFu;
                // my word is the popc(flags below me)-th below the cursor (unread when I do not renormalise: the row
                // has kDecPad bytes in front, so the address stays inside it)
                const uint32_t w = stage[wcur - 1 - int32_t(__popc(flags & below))];
                if (renorm)
                    nx = (nx << kRadixLog2) | uint64_t(w);
                if (active) {
                    x = nx;
                    out[idx] = sym;
                }
                underrun |= wcur < int32_t(__popc(flags));
                wcur = max(wcur - int32_t(__popc(flags)), 0); // corrupt payload: stay inside the row
            }
        }
        if (have_block && a.error_flag && (underrun || wcur != int32_t(delta >> 2)))
            *a.error_flag = 2u;
    }
}

// ------------------------------------------------------------------------------------------
// Size scan + payload compaction
// ------------------------------------------------------------------------------------------

// offsets[b] = sum_{b' < b} roundup16(bytes[b']); total[0] = sum of all; total[1] = unpadded sum.
// One CTA, three phases: every warp sums a contiguous run of 32-element rows (coalesced), the 32 warp totals are
// scanned, then every warp scans its rows again with a running carry and writes the offsets (coalesced).
__global__ void __launch_bounds__(1024) scan_block_sizes_kernel(const uint32_t* __restrict__ bytes, uint32_t n_blocks,
    uint64_t* __restrict__ offsets, uint64_t* __restrict__ totals, uint64_t* __restrict__ header_total)
{
    __shared__ uint64_t s_pad[32];
    __shared__ uint64_t s_raw[32];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t rows = (n_blocks + 31u) >> 5;
    const uint32_t rows_per_warp = (rows + 31u) >> 5;
    const uint32_t r0 = min(rows, warp * rows_per_warp), r1 = min(rows, r0 + rows_per_warp);
    uint64_t pad = 0, raw = 0;
#pragma unroll 4
    for (uint32_t r = r0; r < r1; r++) {
        const uint32_t b = (r << 5) + lane;
        const uint32_t v = b < n_blocks ? __ldg(bytes + b) : 0u;
        pad += (uint64_t(v) + 15u) & ~uint64_t(15);
        raw += v;
    }
    for (int o = 16; o > 0; o >>= 1) {
        pad += __shfl_xor_sync(0xffffffffu, pad, o);
        raw += __shfl_xor_sync(0xffffffffu, raw, o);
    }
    if (lane == 0) {
        s_pad[warp] = pad;
        s_raw[warp] = raw;
    }
    __syncthreads();
    const uint64_t wp = s_pad[lane], wr = s_raw[lane];
    uint64_t carry = lane < warp ? wp : 0, all_pad = wp, all_raw = wr;
    for (int o = 16; o > 0; o >>= 1) {
        carry += __shfl_xor_sync(0xffffffffu, carry, o);
        all_pad += __shfl_xor_sync(0xffffffffu, all_pad, o);
        all_raw += __shfl_xor_sync(0xffffffffu, all_raw, o);
    }
    if (threadIdx.x == 0) {
        totals[0] = all_pad;
        totals[1] = all_raw;
        if (header_total) // the container header's payload_bytes field
            *header_total = all_pad;
    }
#pragma unroll 4
    for (uint32_t r = r0; r < r1; r++) {
        const uint32_t b = (r << 5) + lane;
        const uint32_t v = b < n_blocks ? __ldg(bytes + b) : 0u;
        const uint64_t p = (uint64_t(v) + 15u) & ~uint64_t(15);
        uint64_t incl = p;
        for (int o = 1; o < 32; o <<= 1) {
            const uint64_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= uint32_t(o))
                incl += t;
        }
        if (b < n_blocks)
            offsets[b] = carry + incl - p;
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
}

// Untrusted size table: every block must hold its four 64-bit states (>= 32 bytes) and, where `align_mask` says so
// (plain ANS: whole 32-bit words), be aligned; sets *flag = 3 otherwise.
// `totals` / `expected` (optional): the size scan's sum must be what the container's header says (flag 4).  The block
// decoders return at once when the flag is set, so neither verdict needs a host round trip in front of the decode.
__global__ void __launch_bounds__(256) check_block_sizes_kernel(const uint32_t* __restrict__ bytes, uint32_t n_blocks, uint32_t align_mask,
    uint32_t* __restrict__ flag, const uint64_t* __restrict__ totals = nullptr, uint64_t expected = 0)
{
    if (totals && blockIdx.x == 0 && threadIdx.x == 0 && totals[0] != expected)
        atomicMax(flag, 4u);
    bool bad = false;
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < n_blocks; b += gridDim.x * blockDim.x) {
        const uint32_t v = __ldg(bytes + b);
        bad |= v < 32u || (v & align_mask) != 0u || v > 0x7FFFFF00u;
    }
    if (__any_sync(0xffffffffu, bad) && lane_id() == 0)
        atomicMax(flag, 3u);
}

// one warp per block: scratch + b*stride  ->  payload + offsets[b]; also fills the container's size table.
// Four 16-byte chunks per lane in flight (a block is ~0.9 chunks per lane and pass otherwise: latency-bound).
__global__ void __launch_bounds__(256) compact_kernel(const uint8_t* __restrict__ scratch, uint64_t stride,
    const uint32_t* __restrict__ bytes, const uint64_t* __restrict__ offsets, uint32_t n_blocks,
    uint8_t* __restrict__ payload, uint32_t* __restrict__ size_table, uint64_t payload_capacity)
{
    const uint32_t warps_per_cta = blockDim.x >> 5;
    const uint32_t lane = lane_id();
    for (uint32_t b = blockIdx.x * warps_per_cta + (threadIdx.x >> 5); b < n_blocks; b += gridDim.x * warps_per_cta) {
        const uint32_t sz = bytes[b];
        const uint32_t chunks = (sz + 15u) >> 4;
        const uint64_t at = offsets[b];
        if (at + (uint64_t(chunks) << 4) > payload_capacity)
            continue; // destination too small: the host reports it from the scan's total
        const uint4* src = reinterpret_cast<const uint4*>(scratch + uint64_t(b) * stride);
        uint4* dst = reinterpret_cast<uint4*>(payload + at);
        uint32_t c = lane;
        for (; c + 96u < chunks; c += 128u) {
            const uint4 v0 = __ldcs(src + c), v1 = __ldcs(src + c + 32u), v2 = __ldcs(src + c + 64u), v3 = __ldcs(src + c + 96u);
            dst[c] = v0;
            dst[c + 32u] = v1;
            dst[c + 64u] = v2;
            dst[c + 96u] = v3;
        }
        for (; c < chunks; c += 32u)
            dst[c] = __ldcs(src + c);
        if (lane == 0 && size_table)
            size_table[b] = sz;
    }
}

// ------------------------------------------------------------------------------------------
// Pseudo-adaptive mode (src/pseudo_adaptive.cpp:64-177): one model per block, no shared histogram
// ------------------------------------------------------------------------------------------

// shared-memory byte offset of a value's bin (folded / msb alphabets), as in histogram_kernel<.., true>
// (plain ANS -- dense ranks behind the driver's remap: the bin is the value, clamped to the last bin; the host
// rejects the input by its maximum then)
template <int METHOD> __device__ __forceinline__ uint32_t pa_bin_offset(uint32_t v, uint32_t f, uint32_t last_bin = 0)
{
    if (METHOD == METHOD_INT)
        return min(v, last_bin) << 2;
    const uint32_t step = METHOD == METHOD_MSB ? 256u : (255u << (f - 1u));
    uint32_t fl;
    if (METHOD == METHOD_MSB)
        fl = 31u - __clz((max(v, 1u) - 1u) | 1u);
    else
        fl = 31u - __clz(v | (1u << (f - 1u))) - (f - 1u);
    const uint32_t k8 = fl & 0x18u;
    return (((v >> k8) << 3) + k8 * step) >> 1;
}

// One CTA per model block: the block's histogram over the folded alphabet (u32 bins, plain stores -- every bin of
// every model is written, no memset needed) and the running maximum of the raw values (domain check).
constexpr int kPaThreads = 256;
template <int METHOD>
__global__ void __launch_bounds__(kPaThreads) pa_histogram_kernel(const uint32_t* __restrict__ in, uint64_t n, uint32_t model_syms,
    uint32_t n_models, uint32_t fidelity, uint32_t nbins, uint32_t* __restrict__ bins, uint32_t* __restrict__ max_value)
{
    extern __shared__ uint32_t s_bins[];
    constexpr uint32_t R = 4; // lane-striped replicas, one bank of skew each
    const uint32_t stride = nbins | 1u;
    const uint32_t mine_s = uint32_t(__cvta_generic_to_shared(s_bins + (lane_id() & (R - 1u)) * stride));
    uint32_t local_max = 0;
    for (uint32_t m = blockIdx.x; m < n_models; m += gridDim.x) {
        for (uint32_t i = threadIdx.x; i < R * stride; i += blockDim.x)
            s_bins[i] = 0;
        __syncthreads();
        const uint64_t first = uint64_t(m) * model_syms;
        const uint32_t len = uint32_t(min(uint64_t(model_syms), n - first));
        const uint4* in4 = reinterpret_cast<const uint4*>(in + first); // model_syms is a multiple of 128: aligned
        for (uint32_t i = threadIdx.x; i < (len >> 2); i += blockDim.x) {
            const uint4 q = __ldg(in4 + i);
            red_shared_inc(mine_s + pa_bin_offset<METHOD>(q.x, fidelity, nbins - 1u));
            red_shared_inc(mine_s + pa_bin_offset<METHOD>(q.y, fidelity, nbins - 1u));
            red_shared_inc(mine_s + pa_bin_offset<METHOD>(q.z, fidelity, nbins - 1u));
            red_shared_inc(mine_s + pa_bin_offset<METHOD>(q.w, fidelity, nbins - 1u));
            local_max = max(local_max, max(max(q.x, q.y), max(q.z, q.w)));
        }
        if (threadIdx.x < (len & 3u)) {
            const uint32_t v = in[first + (len & ~3u) + threadIdx.x];
            red_shared_inc(mine_s + pa_bin_offset<METHOD>(v, fidelity, nbins - 1u));
            local_max = max(local_max, v);
        }
        __syncthreads();
        for (uint32_t b = threadIdx.x; b < nbins; b += blockDim.x)
            bins[uint64_t(m) * nbins + b] = s_bins[b] + s_bins[stride + b] + s_bins[2 * stride + b] + s_bins[3 * stride + b];
        __syncthreads();
    }
    for (int o = 16; o > 0; o >>= 1)
        local_max = max(local_max, __shfl_xor_sync(0xffffffffu, local_max, o));
    if (lane_id() == 0 && local_max != 0)
        atomicMax(max_value, local_max);
}

// exclusive prefix sum of a model's nfreqs into shared memory: s_base[0 .. nbins], s_base[nbins] = M
// (2048 bins per pass of the CTA: each thread owns 8 consecutive bins; s_warp: kPaThreads / 32 + 1 words)
__device__ __forceinline__ void pa_prefix(const uint32_t* __restrict__ nfreqs, uint32_t nbins, uint32_t* s_base, uint32_t* s_warp)
{
    uint32_t before = 0; // bins of the passes already done
    for (uint32_t pass0 = 0; pass0 < nbins; pass0 += kPaThreads * 8u) {
        const uint32_t b0 = pass0 + threadIdx.x * 8u;
        uint32_t v[8], sum = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            v[j] = b0 + j < nbins ? __ldg(nfreqs + b0 + j) : 0u;
            sum += v[j];
        }
        uint32_t incl = sum;
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane_id() >= uint32_t(o))
                incl += t;
        }
        __syncthreads(); // (s_warp of the pass before has been read)
        if (lane_id() == 31)
            s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        uint32_t carry = before, all = 0;
        for (uint32_t w = 0; w < kPaThreads / 32; w++) {
            carry += w < (threadIdx.x >> 5) ? s_warp[w] : 0u;
            all += s_warp[w];
        }
        uint32_t run = carry + incl - sum;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (b0 + j <= nbins)
                s_base[b0 + j] = run;
            run += v[j];
        }
        before += all;
    }
    if (threadIdx.x == 0)
        s_base[nbins] = before;
    __syncthreads();
}

// smallest double >= 1/f (b200ans.cu reciprocal_up, same IEEE operations)
__device__ __forceinline__ double pa_reciprocal_up(uint32_t f)
{
    double r = 1.0 / double(f);
    if (fma(r, double(f), -1.0) < 0.0)
        r = __longlong_as_double(__double_as_longlong(r) + 1);
    return r;
}

// encoder tables of all models from their normalised frequencies (ans_fold.hpp:84-91): one CTA per model
__global__ void __launch_bounds__(kPaThreads) pa_enc_tables_kernel(const uint32_t* __restrict__ nfreqs, uint32_t nbins, uint32_t n_models,
    EncEntry* __restrict__ tables)
{
    extern __shared__ uint32_t s_base[]; // nbins + 8 words
    __shared__ uint32_t s_warp[kPaThreads / 32];
    for (uint32_t m = blockIdx.x; m < n_models; m += gridDim.x) {
        const uint32_t* nf = nfreqs + uint64_t(m) * nbins;
        pa_prefix(nf, nbins, s_base, s_warp);
        for (uint32_t b = threadIdx.x; b < nbins; b += blockDim.x) {
            const uint32_t fr = __ldg(nf + b);
            EncEntry e;
            e.freq = fr;
            e.base = s_base[b];
            e.rcp = fr ? pa_reciprocal_up(fr) : 0.0;
            tables[uint64_t(m) * nbins + b] = e;
        }
        __syncthreads();
    }
}

// decoder slot tables of all models (ans_fold.hpp:179-207, ans_msb.hpp:179-204): one CTA per model, one thread per
// slot, the slot's symbol by binary search over the prefix sums
template <int METHOD>
__global__ void __launch_bounds__(kPaThreads) pa_dec_tables_kernel(const uint32_t* __restrict__ nfreqs, uint32_t nbins, uint32_t n_models,
    uint32_t fidelity, const uint8_t* __restrict__ log2_frame, DecEntry* __restrict__ slots, uint64_t slot_stride)
{
    extern __shared__ uint32_t s_base[]; // nbins + 8 words
    __shared__ uint32_t s_warp[kPaThreads / 32];
    for (uint32_t m = blockIdx.x; m < n_models; m += gridDim.x) {
        const uint32_t* nf = nfreqs + uint64_t(m) * nbins;
        pa_prefix(nf, nbins, s_base, s_warp);
        const uint32_t frame = 1u << log2_frame[m];
        DecEntry* out = slots + uint64_t(m) * slot_stride;
        for (uint32_t slot = threadIdx.x; slot < frame; slot += blockDim.x) {
            // largest s with base[s] <= slot (symbols of frequency 0 share a base with their successor: take the last)
            uint32_t lo = 0, hi = nbins; // base[lo] <= slot < base[hi] = M
            while (hi - lo > 1u) {
                const uint32_t mid = (lo + hi) >> 1;
                if (s_base[mid] <= slot)
                    lo = mid;
                else
                    hi = mid;
            }
            const uint32_t fr = s_base[lo + 1] - s_base[lo];
            uint32_t k = 0, value;
            if (METHOD == METHOD_INT)
                value = lo;
            else if (METHOD == METHOD_MSB)
                value = msb_unmap(lo, k);
            else
                value = fold_unmap(lo, fidelity, k);
            out[slot] = DecEntry { uint16_t(fr), uint16_t(slot - s_base[lo]), value + (k << 30) };
        }
        __syncthreads();
    }
}


// symbol -> unfolded base value and raw byte count, arithmetically (pseudo-adaptive decoders: no slot table exists)
template <int METHOD> __device__ __forceinline__ uint32_t pa_unmap(uint32_t s, uint32_t f, uint32_t& k)
{
    if (METHOD == METHOD_INT) {
        k = 0;
        return s;
    }
    if (METHOD == METHOD_MSB)
        return msb_unmap(s, k);
    // fold_unmap (ans_fold.hpp:150-175) without the divide: k <= 3 because values stay below 2^30
    const uint32_t thres = 1u << (f + 7);
    const uint32_t step = 255u << (f - 1);
    k = uint32_t(s >= thres) + uint32_t(s >= thres + step) + uint32_t(s >= thres + 2u * step);
    return (s - step * k) << (8u * k);
}

// Device restatement of adjust_freqs (ans_util.hpp:100-157) for the per-block models, in two kernels.
//   pa_sort_kernel, one CTA per model: symbols sorted by (count, symbol) ascending -- bitonic sort of 64-bit keys in
//      shared memory -- written to global memory with the model's entropy limit H * (1 + H_approx / 1000).
//   pa_fit_kernel, one thread per (model, frame): thread j of a model's 8-lane group fits frame M0 << (8r + j) with
//      the reference's serial chain ratio = M / rest (IEEE divide), share = (uint32)fma(ratio, count, 0.5),
//      M -= share, rest -= count -- the same IEEE operations in the same order as the host code, so shares, the
//      "fit failed" flag and the largest share are bit-identical -- and accumulates the cross entropy of its fit on
//      the way.  The group then walks its eight results in the reference's order (skip failed fits; stop at the
//      16-bit guard with the last good fit; stop when XH < limit; else remember the fit as last good) and the
//      chosen frame is fitted once more, this time storing the shares.  All models are in flight at once, which is
//      what hides the latency of the serial chains.
// Entropy and cross entropy are sums of p * log2(q): the device's log2 and summation order differ from glibc's in
// the last bits, so a model whose decision margin |XH - limit| is below kPaMargin (a thousand times the worst-case
// difference and more) is not decided here: status 2 sends it to the host's adjust_freqs.  Everything else is exact.
constexpr int kPaNormThreads = 64;
constexpr int kPaFitThreads = 128;
constexpr double kPaMargin = 1e-9;

struct PaMeta {
    double limit;       // H * (1 + H_approx / 1000)
    double inv_total;   // 1 / symbols in the block (cross entropy only; the fits divide exactly)
    uint64_t total;     // symbols in the block
    uint32_t first;     // index of the first non-zero key
    uint32_t frame0_lg; // log2 of the first frame tried: sigma rounded up to a power of two
    uint32_t sigma;
    uint32_t pad;
};

// STORE: write the shares to `shares` (the winning fit); otherwise accumulate the cross entropy
// log2 of the integers below 2^16 (the shares of every fit that can win), computed once per context by the same device
// log2 the fits would call: a load instead of ~40 FP64 instructions per symbol and frame, same values
constexpr uint32_t kPaLog2Tab = 65536;
__global__ void __launch_bounds__(256) pa_log2_table_kernel(double* __restrict__ tab)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < kPaLog2Tab)
        tab[i] = log2(double(i));
}

template <bool STORE>
__device__ __forceinline__ void pa_fit(const uint64_t* __restrict__ keys, uint32_t first, uint32_t n_keys, uint32_t lgM, uint64_t rest,
    double inv_total, uint32_t* __restrict__ shares, bool& failed, uint32_t& top, double& xh, const double* __restrict__ log2_tab = nullptr)
{
    int64_t M = int64_t(1) << lgM;
    const double lgm = double(lgM);
    double acc = 0.0;
    uint32_t mx = 0;
    // the keys come in batches of four, one batch ahead of the chain M -> ratio -> share -> M: a global load per step on
    // that chain was most of its latency (ncu r02x: 1.0k cycles per step with 14 warps per SM)
    constexpr uint32_t kAhead = 4;
    uint64_t nxt[kAhead];
#pragma unroll
    for (uint32_t i = 0; i < kAhead; i++)
        nxt[i] = first + i < n_keys ? __ldg(keys + first + i) : 0u;
    // ... and the log2 of a share (a gather from a 512 KiB table) is consumed one batch later: a warp issues in order, so
    // the cross-entropy term waiting for its load would hold up the next step of the chain
    double w_prev[kAhead], l_prev[kAhead];
#pragma unroll
    for (uint32_t i = 0; i < kAhead; i++)
        w_prev[i] = 0.0, l_prev[i] = 0.0;
    bool done = false;
    for (uint32_t r = first; r < n_keys && !done; r += kAhead) {
        uint64_t cur[kAhead];
#pragma unroll
        for (uint32_t i = 0; i < kAhead; i++) {
            cur[i] = nxt[i];
            nxt[i] = r + kAhead + i < n_keys ? __ldg(keys + r + kAhead + i) : 0u;
        }
#pragma unroll
        for (uint32_t i = 0; i < kAhead; i++) {
            if (done || r + i >= n_keys)
                break;
            const uint64_t key = cur[i];
            const uint32_t sym = uint32_t(key) & 0xFFFFu;
            const uint64_t cnt = key >> 16;
            const double ratio = double(M) / double(rest);
            uint32_t share = uint32_t(fma(ratio, double(cnt), 0.5));
            if (share == 0)
                share = 1;
            if (STORE) {
                shares[sym] = share;
            } else {
                acc = fma(w_prev[i], l_prev[i] - lgm, acc); // the term of four steps ago (0 x anything at the start)
                w_prev[i] = double(cnt) * inv_total;
                l_prev[i] = share < kPaLog2Tab ? __ldg(log2_tab + share) : log2(double(share));
            }
            mx = max(mx, share);
            M -= int64_t(share);
            rest -= cnt;
            done = M < 0;
        }
    }
    if (!STORE) {
        // the pending terms, oldest first.  The batch may have ended anywhere: position i of the last (partial) batch holds
        // a newer term than position i + 1 .. of the batch before, so the order is by age, not by index
        const uint32_t steps = n_keys > first ? n_keys - first : 0u;
        const uint32_t start = steps % kAhead; // oldest pending position (when the loop ran to the end)
#pragma unroll
        for (uint32_t i = 0; i < kAhead; i++) {
            const uint32_t at = (start + i) % kAhead;
            double w = 0.0, l = 0.0;
#pragma unroll
            for (uint32_t z = 0; z < kAhead; z++)
                if (z == at)
                    w = w_prev[z], l = l_prev[z];
            acc = fma(w, l - lgm, acc);
        }
    }
    failed = M != 0;
    top = mx;
    xh = -acc;
}

__global__ void __launch_bounds__(kPaNormThreads) pa_sort_bitonic_kernel(const uint32_t* __restrict__ counts, uint32_t nbins, uint32_t n_models,
    uint32_t h_approx, uint64_t* __restrict__ gkeys, PaMeta* __restrict__ meta, uint32_t* __restrict__ nfreqs)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint32_t n_keys = 1;
    while (n_keys < nbins)
        n_keys <<= 1;
    uint64_t* keys = reinterpret_cast<uint64_t*>(smem_raw);
    __shared__ double s_h[kPaNormThreads];
    __shared__ uint32_t s_cnt[kPaNormThreads];
    __shared__ uint64_t s_sum[kPaNormThreads];

    for (uint32_t model = blockIdx.x; model < n_models; model += gridDim.x) {
        const uint32_t* row = counts + uint64_t(model) * nbins;
        uint32_t* out = nfreqs + uint64_t(model) * nbins;
        // Only the symbols that occur are sorted: thread t compacts the non-zero bins of its contiguous run to the
        // back of a power-of-two window of n_sort >= sigma keys (zeros in front), 2.4x less bitonic work for a
        // 500-symbol block in a 1024-bin alphabet.  Keys (count << 16 | symbol) are unique: the order is the same.
        const uint32_t per = n_keys / blockDim.x; // n_keys is a multiple of the CTA size (1024 or 2048 over 64)
        const uint32_t b0 = threadIdx.x * per;
        uint32_t sigma = 0;
        uint64_t total = 0;
        for (uint32_t b = b0; b < b0 + per; b++) {
            const uint32_t cnt = b < nbins ? __ldg(row + b) : 0u;
            sigma += cnt != 0u;
            total += cnt;
            if (b < nbins)
                out[b] = 0; // the winning fit only writes the symbols that occur
        }
        s_cnt[threadIdx.x] = sigma;
        s_sum[threadIdx.x] = total;
        __syncthreads();
        uint32_t before = 0;
        sigma = 0;
        total = 0;
        for (int t = 0; t < kPaNormThreads; t++) {
            before += t < int(threadIdx.x) ? s_cnt[t] : 0u;
            sigma += s_cnt[t];
            total += s_sum[t];
        }
        uint32_t n_sort = 2;
        while (n_sort < sigma)
            n_sort <<= 1;
        const uint32_t pad = n_sort - sigma;
        for (uint32_t r = threadIdx.x; r < pad; r += blockDim.x)
            keys[r] = 0;
        {
            uint32_t at = pad + before;
            for (uint32_t b = b0; b < b0 + per; b++) {
                const uint32_t cnt = b < nbins ? __ldg(row + b) : 0u;
                if (cnt)
                    keys[at++] = (uint64_t(cnt) << 16) | b;
            }
        }
        // bitonic sort, ascending
        for (uint32_t k = 2; k <= n_sort; k <<= 1)
            for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                __syncthreads();
                for (uint32_t t = threadIdx.x; t < (n_sort >> 1); t += blockDim.x) {
                    const uint32_t i = ((t & ~(j - 1u)) << 1) | (t & (j - 1u)); // index with bit j clear
                    const uint32_t l = i | j;
                    const uint64_t a = keys[i], b = keys[l];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) {
                        keys[i] = b;
                        keys[l] = a;
                    }
                }
            }
        __syncthreads();
        const uint32_t first = n_keys - sigma;
        const double totald = double(total);
        uint64_t* gk = gkeys + uint64_t(model) * n_keys + (n_keys - n_sort); // the fits read [first, n_keys) only
        double acc = 0.0; // entropy: fixed-shape reduction
        for (uint32_t r = threadIdx.x; r < n_sort; r += blockDim.x) {
            const uint64_t key = keys[r];
            gk[r] = key;
            if (r >= pad) {
                const double p = double(key >> 16) / totald;
                acc = fma(p, log2(p), acc);
            }
        }
        s_h[threadIdx.x] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double H = 0.0;
            for (int t = 0; t < kPaNormThreads; t++)
                H += s_h[t];
            H = -H;
            uint32_t lg = 0; // sigma if it is a power of two, else next_power_of_two: 2^(floor(log2 sigma) + 1)
            while ((1u << lg) < sigma)
                lg++;
            PaMeta mt;
            mt.limit = H * (1.0 + double(h_approx) / 1000.0);
            mt.inv_total = 1.0 / totald;
            mt.total = total;
            mt.first = first;
            mt.frame0_lg = lg;
            mt.sigma = sigma;
            mt.pad = 0;
            meta[model] = mt;
        }
        __syncthreads();
    }
}

// The same result -- keys (count << 16 | symbol) of the symbols that occur, ascending, at gkeys[model][first .. n_keys),
// the entropy limit, the zeroed frequency row -- by a COUNTING sort, one warp per model (round 2; the bitonic sort above
// was 0.6 ms of a 2.8 ms encode for 16 384 models: 244 M warp instructions, bound by shared-memory latency).  A block of
// model_syms symbols has few distinct counts: counts below kPaSortSmall are histogrammed in shared memory, scanned, and
// the symbols scattered in ascending symbol order (a warp-wide match on the count keeps equal counts in symbol order);
// the at most model_syms / kPaSortSmall symbols with larger counts are ranked among themselves by comparison.
constexpr int kPaSortWarps = 8;
constexpr uint32_t kPaSortSmall = 2048; // counts below this go through the histogram
constexpr uint32_t kPaSortBig = 64;     // room for the others: model_syms <= kPaSortSmall * kPaSortBig
__host__ __device__ inline size_t pa_sort_smem_bytes() { return size_t(kPaSortWarps) * (size_t(kPaSortSmall) * 4 + size_t(kPaSortBig) * 8); }

__global__ void __launch_bounds__(kPaSortWarps * 32) pa_sort_kernel(const uint32_t* __restrict__ counts, uint32_t nbins, uint32_t n_keys, uint32_t n_models,
    uint32_t h_approx, uint64_t* __restrict__ gkeys, PaMeta* __restrict__ meta, uint32_t* __restrict__ nfreqs, const double* __restrict__ log2_tab)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem_raw) + size_t(warp) * kPaSortSmall;
    uint64_t* bigk = reinterpret_cast<uint64_t*>(smem_raw + size_t(kPaSortWarps) * kPaSortSmall * 4) + size_t(warp) * kPaSortBig;
    for (uint32_t c = lane; c < kPaSortSmall; c += 32u)
        hist[c] = 0u;
    __syncwarp();
    for (uint32_t model = blockIdx.x * kPaSortWarps + warp; model < n_models; model += gridDim.x * kPaSortWarps) {
        const uint32_t* row = counts + uint64_t(model) * nbins;
        uint32_t* out = nfreqs + uint64_t(model) * nbins;
        // pass 1: how many symbols occur, how many have each small count, the large ones aside
        uint32_t sigma = 0, n_big = 0, max_small = 0, used_bins = 0; // used_bins: behind the last chunk that holds a count
        uint64_t total = 0;
        for (uint32_t base = 0; base < nbins; base += 32u) {
            const uint32_t b = base + lane;
            const uint32_t cnt = b < nbins ? __ldg(row + b) : 0u;
            if (b < nbins)
                out[b] = 0; // the winning fit only writes the symbols that occur
            const bool small = cnt != 0u && cnt < kPaSortSmall, big = cnt >= kPaSortSmall;
            const uint32_t present = __ballot_sync(0xffffffffu, cnt != 0u);
            sigma += uint32_t(__popc(present));
            used_bins = present ? base + 32u : used_bins;
            total += cnt;
            if (small) {
                atomicAdd(hist + cnt, 1u);
                max_small = max(max_small, cnt);
            }
            const uint32_t bm = __ballot_sync(0xffffffffu, big);
            if (big) {
                const uint32_t slot = n_big + uint32_t(__popc(bm & lt));
                if (slot < kPaSortBig)
                    bigk[slot] = (uint64_t(cnt) << 16) | b;
            }
            n_big += uint32_t(__popc(bm));
        }
        for (int o = 16; o > 0; o >>= 1) {
            total += __shfl_xor_sync(0xffffffffu, total, o);
            max_small = max(max_small, __shfl_xor_sync(0xffffffffu, max_small, o));
        }
        n_big = min(n_big, kPaSortBig); // (the host launches this kernel only when more cannot occur)
        __syncwarp();
        // exclusive scan of the histogram: hist[c] = keys with a smaller count
        uint32_t n_small = 0;
        for (uint32_t base = 0; base <= max_small; base += 32u) {
            const uint32_t c = base + lane;
            const uint32_t x = c <= max_small ? hist[c] : 0u;
            uint32_t incl = x;
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= uint32_t(o))
                    incl += up;
            }
            if (c <= max_small)
                hist[c] = n_small + incl - x;
            n_small += __shfl_sync(0xffffffffu, incl, 31);
        }
        __syncwarp();
        // pass 2: scatter in ascending symbol order; sum of count * log2(count) for the entropy
        const uint32_t first = n_keys - sigma;
        uint64_t* gk = gkeys + uint64_t(model) * n_keys + first;
        double acc = 0.0;
        for (uint32_t base = 0; base < used_bins; base += 32u) {
            const uint32_t b = base + lane;
            const uint32_t cnt = b < nbins ? __ldg(row + b) : 0u;
            const bool small = cnt != 0u && cnt < kPaSortSmall;
            const uint32_t peers = __match_any_sync(0xffffffffu, small ? cnt : 0u);
            const uint32_t at = small ? hist[cnt] : 0u;
            __syncwarp();
            if (small && (peers & lt) == 0u)
                hist[cnt] = at + uint32_t(__popc(peers));
            __syncwarp();
            if (small)
                gk[at + uint32_t(__popc(peers & lt))] = (uint64_t(cnt) << 16) | b;
            if (cnt != 0u)
                acc = fma(double(cnt), cnt < kPaLog2Tab ? __ldg(log2_tab + cnt) : log2(double(cnt)), acc);
        }
        // the large counts: rank among themselves
        for (uint32_t i = lane; i < n_big; i += 32u) {
            const uint64_t key = bigk[i];
            uint32_t rank = 0;
            for (uint32_t j = 0; j < n_big; j++)
                rank += bigk[j] < key;
            gk[n_small + rank] = key;
        }
        for (uint32_t c = lane; c <= max_small; c += 32u) // leave the histogram as it was found
            hist[c] = 0u;
        for (int o = 16; o > 0; o >>= 1)
            acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) {
            const double totald = double(total);
            const double H = total ? log2(totald) - acc / totald : 0.0; // -sum p log2 p, p = count / total
            uint32_t lg = 0; // sigma if it is a power of two, else next_power_of_two: 2^(floor(log2 sigma) + 1)
            while ((1u << lg) < sigma)
                lg++;
            PaMeta mt;
            mt.limit = H * (1.0 + double(h_approx) / 1000.0);
            mt.inv_total = 1.0 / totald;
            mt.total = total;
            mt.first = first;
            mt.frame0_lg = lg;
            mt.sigma = sigma;
            mt.pad = 0;
            meta[model] = mt;
        }
        __syncwarp();
    }
}

// status: 0 = decided here, 1 = unusable (the reference would return an empty model), 2 = undecided (host decides)
// Four lanes per model fit four frames at a time (round 2: eight at a time left five of eight chains unused on typical
// blocks, whose frame is M0 .. 8 M0); the shares of the winning frame are written by pa_fit_store_kernel, one THREAD per
// model (inside this kernel that pass ran with one lane in eight active).
constexpr int kPaFitGroup = 4;
__global__ void __launch_bounds__(kPaFitThreads) pa_fit_kernel(const uint64_t* __restrict__ gkeys, uint32_t n_keys, const PaMeta* __restrict__ meta,
    uint32_t n_models, double margin, uint8_t* __restrict__ log2_frame, uint8_t* __restrict__ status, uint32_t* __restrict__ max_log2,
    const double* __restrict__ log2_tab, bool require_u16)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t j = t & uint32_t(kPaFitGroup - 1);
    const bool live = (t / kPaFitGroup) < n_models;
    const uint32_t model = live ? (t / kPaFitGroup) : n_models - 1u; // spare lanes shadow the last model: shuffles stay full-warp
    const PaMeta mt = meta[model];
    const uint64_t* keys = gkeys + uint64_t(model) * n_keys;
    int st = mt.sigma == 0 ? 1 : -1, choice = -1, last_good = -1;
    for (int round = 0; round < 32 / kPaFitGroup; round++) {
        if (!__any_sync(0xffffffffu, st < 0))
            break;
        bool failed = true;
        uint32_t top = 0;
        double xh = 0.0;
        if (st < 0)
            pa_fit<false>(keys, mt.first, n_keys, mt.frame0_lg + uint32_t(kPaFitGroup * round) + j, mt.total, mt.inv_total, nullptr, failed, top, xh,
                log2_tab);
        const uint32_t code = failed ? 0xFFFFFFFFu : top;
#pragma unroll
        for (int jj = 0; jj < kPaFitGroup; jj++) { // every lane of the group walks the results in the reference's order
            const uint32_t cj = __shfl_sync(0xffffffffu, code, jj, kPaFitGroup);
            const double xj = __shfl_sync(0xffffffffu, xh, jj, kPaFitGroup);
            if (st >= 0 || cj == 0xFFFFFFFFu)
                continue;
            if (cj >= 65535u) { // the 16-bit guard: back to the last good fit (possibly none)
                if (!require_u16) { // plain ANS has no guard (ans_int.hpp:50) and these coders no wider tables: the host decides
                    st = 2;
                    continue;
                }
                choice = last_good;
                st = last_good < 0 ? 1 : 0;
                continue;
            }
            const double d = xj - mt.limit;
            if (fabs(d) <= margin) {
                st = 2;
            } else if (d < 0.0) {
                choice = kPaFitGroup * round + jj;
                st = 0;
            } else {
                last_good = kPaFitGroup * round + jj;
            }
        }
    }
    if (st < 0)
        st = 2; // no decision in 32 doublings: let the host look at it
    if (live && j == 0u) {
        const uint32_t lg = st == 0 ? mt.frame0_lg + uint32_t(choice) : 0u;
        status[model] = uint8_t(st);
        log2_frame[model] = uint8_t(lg);
        atomicMax(max_log2, lg); // the coder picks its table layout by the largest frame
    }
}

__global__ void __launch_bounds__(kPaFitThreads) pa_fit_store_kernel(const uint64_t* __restrict__ gkeys, uint32_t n_keys, const PaMeta* __restrict__ meta,
    uint32_t nbins, uint32_t n_models, const uint8_t* __restrict__ log2_frame, const uint8_t* __restrict__ status, uint32_t* __restrict__ nfreqs)
{
    const uint32_t model = blockIdx.x * blockDim.x + threadIdx.x;
    if (model >= n_models || status[model] != 0)
        return;
    const PaMeta mt = meta[model];
    bool failed;
    uint32_t top;
    double xh;
    pa_fit<true>(gkeys + uint64_t(model) * n_keys, mt.first, n_keys, log2_frame[model], mt.total, mt.inv_total, nfreqs + uint64_t(model) * nbins, failed, top, xh);
}


// ------------------------------------------------------------------------------------------
// Preludes on the device (pseudo-adaptive mode): the reference-format prelude of every model,
//   vbyte(max_sym) . u8(log2 M) . binary interpolative code of the prefix sums of (nfreq + 1)
// (ans_util.hpp:25-63, interp.hpp:28-118, bits.hpp:146-218; same bits as host/prelude.hpp).  The recursion of the
// interpolative code is data-independent in shape: the encoder walks it by position (a warp per model, 32 nodes at a
// time), the decoder on an explicit stack (a thread per model: where a node starts depends on every node before it).
// ------------------------------------------------------------------------------------------
constexpr int kPaPreThreads = 64;
constexpr int kPaPreEncWarps = 8; // models per CTA of the prelude encoder

// codeword of the centred minimal binary code as one LSB-first bit string (host/prelude.hpp put_centered), u < 2^31
__device__ __forceinline__ void pa_centered_code(uint32_t val, uint32_t u, uint32_t& code, uint32_t& nbits)
{
    code = 0;
    nbits = 0;
    if (u == 1)
        return;
    const uint32_t b = 32u - __clz(u - 1u);
    const uint32_t span = 1u << b;
    const uint32_t longs = 2u * u - span;
    val += u - (longs >> 1);
    if (val > u)
        val -= u;
    const uint32_t shorts = span - u;
    if (val <= shorts) {
        code = val - 1u;
        nbits = b - 1u;
    } else {
        val += shorts;
        code = ((val - 1u) >> 1) | (((val - 1u) & 1u) << (b - 1u));
        nbits = b;
    }
}

__host__ __device__ inline size_t pa_prelude_enc_smem(uint32_t nbins, uint32_t warps = kPaPreEncWarps) { return size_t(warps) * (size_t(nbins) + 72u) * 4u; }

// nfreqs[model][nbins], log2_frame[model]  ->  prelude bytes at pre + model * pre_stride, pre_bytes[model].
// One WARP per model (round 2; the first version was one thread per model, 0.45 ms for 16 384 models next to the
// encoder).  The recursion of the interpolative code is data-independent in shape: which list entry is coded p-th
// (pre-order) and the segment it splits follow from p and n alone, and its bounds are the entries next to that segment.
// So the lanes code 32 consecutive nodes at a time, a warp scan of the code lengths gives the bit positions, and the
// words are assembled in shared memory (the same scheme as remap_alpha_kernel, remap_kernels.cuh).
// `scan_bins` <= nbins: no model has a symbol at or beyond it (the rows are nbins wide, the shared-memory arrays scan_bins).
__global__ void __launch_bounds__(kPaPreEncWarps * 32) pa_prelude_encode_kernel(const uint32_t* __restrict__ nfreqs, uint32_t nbins, uint32_t scan_bins,
    uint32_t n_models, const uint8_t* __restrict__ log2_frame, uint8_t* __restrict__ pre, uint32_t pre_stride, uint32_t* __restrict__ pre_bytes)
{
    extern __shared__ __align__(16) uint32_t pa_pre_smem[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t* v = pa_pre_smem + size_t(warp) * (scan_bins + 72u); // the prefix sums the code is built on
    uint32_t* buf = v + scan_bins;                                 // 72 words of bit stream under construction
    const uint32_t cta_warps = blockDim.x >> 5; // (large alphabets leave room for fewer than kPaPreEncWarps)
    for (uint32_t m = blockIdx.x * cta_warps + warp; m < n_models; m += gridDim.x * cta_warps) {
        const uint32_t* nf = nfreqs + uint64_t(m) * nbins;
        __syncwarp();
        uint32_t carry = 0, max_sym = 0;
        for (uint32_t base = 0; base < scan_bins; base += 32u) {
            const uint32_t s = base + lane;
            const uint32_t f = s < scan_bins ? __ldg(nf + s) : 0u;
            uint32_t x = s < scan_bins ? f + (s ? 1u : 0u) : 0u;
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, x, o);
                if (lane >= uint32_t(o))
                    x += up;
            }
            x += carry;
            if (s < scan_bins)
                v[s] = x;
            carry = __shfl_sync(0xffffffffu, x, 31);
            const uint32_t present = __ballot_sync(0xffffffffu, f != 0u);
            if (present)
                max_sym = base + 31u - uint32_t(__clz(present));
        }
        __syncwarp();
        const uint32_t n = max_sym + 1u;
        const uint32_t lg = log2_frame[m];
        const uint32_t top = (1u << lg) + n + 2u; // interp_put(v, n, low = 1, high = u + 1), u = frame + n + 1
        uint8_t* out = pre + uint64_t(m) * pre_stride;
        uint32_t hdr = 0;
        for (uint32_t x = max_sym;; x >>= 7) { // vbyte (vbyte.hpp:57-75), then log2 M
            const bool more = x >= 0x80u;
            if (lane == 0)
                out[hdr] = uint8_t(more ? (x | 0x80u) : x);
            hdr++;
            if (!more)
                break;
        }
        if (lane == 0)
            out[hdr] = uint8_t(lg);
        hdr++;
        uint8_t* words_out = out + hdr; // 32-bit words of the bit stream, at whatever alignment the header leaves
        uint32_t bitpos = 0;
        buf[lane] = 0u, buf[lane + 32u] = 0u;
        if (lane < 8u)
            buf[64u + lane] = 0u;
        __syncwarp();
        for (uint32_t base = 0; base < n; base += 32u) {
            const uint32_t p = base + lane;
            uint32_t code = 0, nbits = 0;
            if (p < n) {
                uint32_t off = 0, cnt = n, q = p, h, left, right;
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
                const uint32_t mid = v[off + h - 1u] + 1u;
                const uint32_t low = off == 0u ? 1u : v[off - 1u] + 2u;
                const uint32_t high = off + cnt == n ? top : v[off + cnt];
                pa_centered_code(mid - low - left + 1u, high - right - low - left + 1u, code, nbits);
            }
            uint32_t incl = nbits;
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= uint32_t(o))
                    incl += up;
            }
            const uint32_t chunk_bits = __shfl_sync(0xffffffffu, incl, 31);
            const uint32_t word0 = bitpos >> 5; // buf[0] is this word (partly filled)
            if (nbits) {
                const uint32_t at = bitpos + (incl - nbits);
                const uint32_t wi = (at >> 5) - word0, sh = at & 31u;
                const uint64_t shifted = uint64_t(code) << sh;
                atomicOr(buf + wi, uint32_t(shifted));
                if (shifted >> 32)
                    atomicOr(buf + wi + 1u, uint32_t(shifted >> 32));
            }
            __syncwarp();
            bitpos += chunk_bits;
            const uint32_t full = (bitpos >> 5) - word0; // whole words ready (<= 33)
            for (uint32_t i = lane; i < full; i += 32u) {
                const uint32_t w = buf[i];
                uint8_t* o = words_out + 4u * (word0 + i);
                o[0] = uint8_t(w), o[1] = uint8_t(w >> 8), o[2] = uint8_t(w >> 16), o[3] = uint8_t(w >> 24);
            }
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
        uint32_t words = bitpos >> 5;
        if (bitpos & 31u) { // the last, partly filled word goes out whole (bits.hpp)
            if (lane == 0) {
                const uint32_t w = buf[0];
                uint8_t* o = words_out + 4u * words;
                o[0] = uint8_t(w), o[1] = uint8_t(w >> 8), o[2] = uint8_t(w >> 16), o[3] = uint8_t(w >> 24);
            }
            words++;
        }
        const uint32_t bytes = hdr + 4u * words;
        if (lane == 0)
            pre_bytes[m] = bytes;
        for (uint32_t z = bytes + lane; z < ((bytes + 15u) & ~15u); z += 32u) // the compaction copies whole 16-byte chunks: keep them deterministic
            out[z] = 0;
    }
}

// ---- the way back: a serial bit stream per model, one thread each -------------------------------------------------
// Interpolative DEcoding is a chain: where a node's code starts depends on every node before it (pre-order), and how
// wide it is on its ancestors' values.  The first version (one loop trip per node OR per stack pop, the stack in local
// memory, the stream read byte by byte) spent 170 warp instructions per node at 70 % lane efficiency (ncu r02u: 46 M warp
// instructions for 16 384 x 527 nodes, 3.4 warps per SM, 5.9 cycles per instruction: 0.38 ms).  This one keeps the 32
// models of a warp in lockstep -- every loop trip decodes exactly one node in every lane, descent / sibling / pop are
// selects -- reads the stream as aligned 32-bit words two ahead of the accumulator, keeps the stack of pending right
// halves in shared memory (one column per thread: conflict-free), and leaves the frequencies (differences, checks,
// zero fill: coalesced work) to pa_prelude_finish_kernel, one warp per model.
struct PaWordReader {
    const uint32_t* words; // aligned words; word 0 holds the first stream byte
    uint32_t sh;           // bit offset of the stream inside word 0 (0, 8, 16, 24)
    uint32_t idx;          // next aligned word to load
    uint32_t last;         // highest aligned word that may be loaded (reads beyond repeat it)
    uint32_t a0, a1, a2;   // aligned words idx - 3 .. idx - 1: the next stream word is (a1 : a0) >> sh
    uint64_t acc = 0;
    uint32_t fill = 0;
    uint32_t taken = 0;    // stream words moved into acc
    __device__ __forceinline__ uint32_t load(uint32_t i) const { return __ldg(words + min(i, last)); }
    __device__ __forceinline__ void start(const uint8_t* base, uint64_t first_byte, uint64_t last_byte)
    {
        const uintptr_t addr = reinterpret_cast<uintptr_t>(base) + first_byte;
        words = reinterpret_cast<const uint32_t*>(addr & ~uintptr_t(3));
        sh = uint32_t(addr & 3u) * 8u;
        const uintptr_t end = reinterpret_cast<uintptr_t>(base) + last_byte; // last byte that exists
        last = end >= addr ? uint32_t(((end & ~uintptr_t(3)) - (addr & ~uintptr_t(3))) >> 2) : 0u;
        a0 = load(0), a1 = load(1), a2 = load(2);
        idx = 3;
    }
    // at least 33 bits in acc afterwards (one call per node: no code is longer than that)
    __device__ __forceinline__ void top_up()
    {
        if (fill <= 32u) {
            const uint32_t w = __funnelshift_r(a0, a1, sh);
            a0 = a1, a1 = a2, a2 = load(idx);
            idx++;
            acc |= uint64_t(w) << fill;
            fill += 32u;
            taken++;
        }
    }
    __device__ __forceinline__ uint32_t take(uint32_t width) // width <= 32, and at most `fill`
    {
        const uint32_t v = uint32_t(acc) & uint32_t((uint64_t(1) << width) - 1u);
        acc >>= width;
        fill -= width;
        return v;
    }
    __device__ __forceinline__ uint64_t bits_used() const { return uint64_t(taken) * 32u - fill; }
};

// value in [1, u] of the centred minimal binary code (host/prelude.hpp get_centered), u >= 1, codes of at most 33 bits
template <class W> __device__ __forceinline__ W pa_take_centered(PaWordReader& r, W u)
{
    const uint32_t b = sizeof(W) == 8 ? 64u - uint32_t(__clzll((long long)(u - 1u))) : 32u - uint32_t(__clz(int(uint32_t(u) - 1u)));
    const W span = W(1) << b;
    const W longs = 2u * u - span;
    const W shorts = span - u;
    W val = W(r.take(b ? b - 1u : 0u)) + 1u;
    if (b != 0u && val > shorts)
        val = 2u * val + r.take(1u) - shorts - 1u;
    val += longs >> 1;
    if (val > u)
        val -= u;
    return b ? val : W(1);
}

// interp_get (interp.hpp:64-118) of `n` values with bounds [low, high] into v[0 .. n); one node per loop trip.  `stack`
// is this thread's column of a [depth][4 or 6][threads] array of words (stride = threads); depth >= log2(n) + 1.
// Returns false when the stream is not one this coder wrote (bounds that cross).
template <class W>
__device__ __forceinline__ bool pa_interp_decode(PaWordReader& r, uint32_t n, W low, W high, W max_u, uint32_t* __restrict__ v, uint32_t* stack,
    uint32_t stride)
{
    constexpr uint32_t kWords = sizeof(W) == 8 ? 6u : 4u;
    uint32_t off = 0, cnt = n, sp = 0;
    bool ok = true;
    while (cnt != 0u) {
        const uint32_t h = (cnt + 1u) >> 1, left = h - 1u, right = cnt - h;
        const W u = high - low - W(cnt) + 2u; // high - right - low - left + 1
        if (high + 2u <= low + W(cnt) || u > max_u) {
            ok = false;
            break;
        }
        r.top_up();
        const W mid = low + left - 1u + pa_take_centered<W>(r, u);
        v[off + h - 1u] = uint32_t(mid - 1u);
        const bool go_left = left != 0u, go_right = right != 0u;
        if (go_left && go_right) { // the right half waits
            uint32_t* e = stack + size_t(sp) * kWords * stride;
            e[0] = off + h;
            e[stride] = right;
            if (sizeof(W) == 8) {
                const uint64_t lo = uint64_t(mid) + 1u, hi = uint64_t(high);
                e[2 * stride] = uint32_t(lo), e[3 * stride] = uint32_t(lo >> 32);
                e[4 * stride] = uint32_t(hi), e[5 * stride] = uint32_t(hi >> 32);
            } else {
                e[2 * stride] = uint32_t(mid) + 1u;
                e[3 * stride] = uint32_t(high);
            }
            sp++;
        }
        if (go_left) {
            cnt = left;
            high = mid - 1u;
        } else if (go_right) {
            off += h;
            cnt = right;
            low = mid + 1u;
        } else if (sp != 0u) {
            sp--;
            const uint32_t* e = stack + size_t(sp) * kWords * stride;
            off = e[0];
            cnt = e[stride];
            if (sizeof(W) == 8) {
                low = W(uint64_t(e[2 * stride]) | (uint64_t(e[3 * stride]) << 32));
                high = W(uint64_t(e[4 * stride]) | (uint64_t(e[5 * stride]) << 32));
            } else {
                low = W(e[2 * stride]);
                high = W(e[3 * stride]);
            }
        } else {
            cnt = 0u;
        }
    }
    return ok;
}

constexpr int kPaPreDepth = 15; // pending right halves: at most log2(n) + 1, n <= 16384 symbols per model

// prelude bytes of model m at src + offsets[m] (pre_bytes[m] of them)  ->  climb[model][0 .. n) (the prefix sums the code
// carries), log2_frame[model] (0xFF: not a valid prelude), nfreqs[model][0] = n.  pa_prelude_finish_kernel follows.
__global__ void __launch_bounds__(kPaPreThreads) pa_prelude_decode_kernel(const uint8_t* __restrict__ src, const uint64_t* __restrict__ offsets,
    const uint32_t* __restrict__ pre_bytes, uint64_t region_bytes, uint32_t nbins, uint32_t n_models, uint32_t* __restrict__ climb,
    uint32_t* __restrict__ nfreqs, uint8_t* __restrict__ log2_frame)
{
    __shared__ uint32_t s_stack[kPaPreDepth * 4 * kPaPreThreads];
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_models)
        return;
    const bool inside = offsets[m] + pre_bytes[m] <= region_bytes && pre_bytes[m] != 0u; // the size table is untrusted
    const uint8_t* in = src + (inside ? offsets[m] : 0);
    const uint32_t limit = inside ? pre_bytes[m] : 0u;
    uint32_t max_sym = 0, pos = 0;
    for (uint32_t shift = 0; shift < 35u; shift += 7) {
        const uint32_t c = pos < limit ? in[pos] : 0u;
        pos++;
        max_sym += (c & 0x7Fu) << shift;
        if (c < 0x80u)
            break;
    }
    const uint32_t lg = pos < limit ? in[pos] : 0xFFu;
    pos++;
    bool bad = !inside || max_sym >= nbins || max_sym >= (1u << (kPaPreDepth - 1)) || lg > 20u || pos > limit;
    const uint32_t n = bad ? 0u : max_sym + 1u;
    const uint32_t frame = 1u << (bad ? 0u : lg);
    PaWordReader r;
    r.start(src, (inside ? offsets[m] : 0) + (bad ? 0u : pos), region_bytes ? region_bytes - 1u : 0u);
    if (!pa_interp_decode<uint32_t>(r, n, 1u, frame + n + 2u, 1u << 23, climb + uint64_t(m) * nbins, s_stack + threadIdx.x, kPaPreThreads))
        bad = true;
    if (!bad && r.bits_used() > uint64_t(limit - pos) * 8u)
        bad = true;
    nfreqs[uint64_t(m) * nbins] = bad ? 0u : n;
    log2_frame[m] = bad ? uint8_t(0xFF) : uint8_t(lg);
}

// climb[model][0 .. n) -> nfreqs[model][nbins]; anything that is not a valid model of at most nbins symbols sets *error
// (and leaves a harmless one-symbol model).  One warp per model.
constexpr int kPaPreFinishWarps = 8;
__global__ void __launch_bounds__(kPaPreFinishWarps * 32) pa_prelude_finish_kernel(const uint32_t* __restrict__ climb, uint32_t nbins, uint32_t n_models,
    uint32_t* __restrict__ nfreqs, uint8_t* __restrict__ log2_frame, uint32_t* __restrict__ error, uint32_t* __restrict__ max_log2,
    uint32_t* __restrict__ max_symbol)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t m = blockIdx.x * kPaPreFinishWarps + (threadIdx.x >> 5);
    if (m >= n_models)
        return;
    const uint32_t* v = climb + uint64_t(m) * nbins;
    uint32_t* nf = nfreqs + uint64_t(m) * nbins;
    const uint32_t n = min(nf[0], nbins);
    const uint32_t lg = log2_frame[m];
    __syncwarp();
    bool bad = n == 0u || lg > 20u;
    uint64_t sum = 0;
    for (uint32_t base = 0; base < nbins; base += 32u) {
        const uint32_t s = base + lane;
        uint32_t f = 0;
        if (!bad && s < n) {
            const uint32_t here = __ldg(v + s);
            f = s ? here - __ldg(v + s - 1u) - 1u : here;
            if (f > 65535u)
                bad = true;
        }
        if (s < nbins)
            nf[s] = f;
        sum += f;
    }
    for (int o = 16; o > 0; o >>= 1)
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
    bad = __any_sync(0xffffffffu, bad) || sum != (uint64_t(1) << (lg & 31u));
    if (bad) {
        for (uint32_t s = lane; s < nbins; s += 32u)
            nf[s] = s == 0u ? 1u : 0u;
        if (lane == 0) {
            atomicExch(error, 1u);
            log2_frame[m] = 0;
        }
    } else if (lane == 0) {
        atomicMax(max_log2, lg);
        atomicMax(max_symbol, n - 1u); // the decoder sizes its per-model tables by the largest symbol
    }
}


} // namespace b200ans
