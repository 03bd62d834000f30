// b200ans device code (sm_100a): the ANSrfold model on the device ("K4" of SURVEY.md section 8a).
//
// ans_reorder_fold_encode::create (ans_reorder_fold.hpp:75-106) sorts (-count, value) over the whole value range,
// gives the first K = 2^(f+7) values the ids 0..K-1 and moves everything else up by K; the folded histogram is then
// counted over mapping[in[i]] (:110-115).  Here, from the raw device histogram `raw[v]`:
//
//   1. rf_digit_hist / rf_digit_pick : radix select (12 bits per pass, most significant digit first) of the K-th
//      largest count T; n_gt = values with count > T, need_eq = K - n_gt; sigma = distinct values.
//   2. rf_tie_count / rf_tie_scan / rf_collect : the candidates are every value with count > T and the need_eq
//      SMALLEST values with count == T (the reference's tie-break: pairs compare by value after the count) -- an
//      ordered selection: per-tile counts of ties, their exclusive scan, then every tile emits its share.
//   3. rf_rank : id of a candidate = number of candidates that sort before it (K^2 compares, K <= 32768).
//   4. rf_mapping_fill / rf_mapping_hot : mapping[v] = v + K, mapping[most_frequent[r]] = r (identity when
//      sigma < K, ans_reorder_fold.hpp:94-97).
//   5. rf_fold_bins : the folded histogram follows from the raw one -- freqs[fold(mapping[v])] += raw[v] -- a pass
//      over the value range (16 M bins for C3) instead of a second pass over the input (1e9 symbols).
//
// Nothing comes back to the host before step 5; the host then reads the folded bins, most_frequent[] and sigma in
// one synchronisation and builds the model (normalisation and prelude stay on the host, as the north star says).
#pragma once

#include "kernels.cuh"

namespace b200ans {

// device-side state of one selection (all counters zeroed by the host before step 1)
struct RfState {
    unsigned long long prefix;   // bits of T fixed so far (above the current digit)
    unsigned long long k_left;   // how many of the values matching the prefix are still wanted
    unsigned long long sigma;    // distinct values
    unsigned long long n_gt;     // values with count > T  (after the last pick)
    unsigned long long need_eq;  // values with count == T to take, smallest first
    unsigned int n_emitted;      // candidates with count > T written so far (rf_collect)
    unsigned int reorder;        // 1: sigma >= K, the mapping reorders; 0: identity mapping
    unsigned int max_seen;       // largest value with a non-zero count
    unsigned int pad;
};

constexpr int kRfThreads = 256;
constexpr uint32_t kRfTile = 4096; // values per tile of the ordered selection

constexpr uint32_t kRfDigitBits = 12;
constexpr uint32_t kRfDigits = 1u << kRfDigitBits;

// digit histogram of the counts whose bits above `shift + kRfDigitBits` equal the prefix (privatised per CTA: most
// counts of a skewed input share their upper digits); the first pass also counts sigma and the largest value seen
__global__ void __launch_bounds__(kRfThreads) rf_digit_hist_kernel(const unsigned long long* __restrict__ raw, uint32_t nvalues,
    uint32_t shift, bool first, const RfState* __restrict__ st, unsigned int* __restrict__ hist /* kRfDigits, zeroed */,
    RfState* __restrict__ st_out)
{
    __shared__ unsigned int s_hist[kRfDigits];
    for (uint32_t i = threadIdx.x; i < kRfDigits; i += blockDim.x)
        s_hist[i] = 0u;
    __syncthreads();
    const unsigned long long prefix = first ? 0ull : st->prefix;
    const uint32_t hi_shift = shift + kRfDigitBits;
    unsigned int nz = 0, top = 0;
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nvalues; v += gridDim.x * blockDim.x) {
        const unsigned long long c = raw[v];
        nz += c != 0ull;
        if (c != 0ull)
            top = v; // v only grows
        if (hi_shift >= 64u || (c >> hi_shift) == (prefix >> hi_shift))
            atomicAdd(s_hist + uint32_t((c >> shift) & (kRfDigits - 1u)), 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < kRfDigits; i += blockDim.x)
        if (s_hist[i])
            atomicAdd(hist + i, s_hist[i]);
    if (first) {
        for (int o = 16; o > 0; o >>= 1) {
            nz += __shfl_xor_sync(0xffffffffu, nz, o);
            top = max(top, __shfl_xor_sync(0xffffffffu, top, o));
        }
        if ((threadIdx.x & 31) == 0 && nz) {
            atomicAdd(&st_out->sigma, (unsigned long long)nz);
            atomicMax(&st_out->max_seen, top);
        }
    }
}

// one CTA of kRfDigits / 4 threads: find the digit at which the counts, walked from the top, cover k_left values; fix it
__global__ void __launch_bounds__(kRfDigits / 4) rf_digit_pick_kernel(unsigned int* __restrict__ hist, uint32_t shift, bool first, uint32_t K,
    RfState* __restrict__ st)
{
    constexpr uint32_t kT = kRfDigits / 4;
    __shared__ unsigned long long s_warp[kT / 32];
    __shared__ unsigned long long s_above;
    __shared__ unsigned int s_digit, s_found;
    const uint32_t t = threadIdx.x;
    if (t == 0)
        s_found = 0u;
    const unsigned long long k_left = first ? (unsigned long long)K : st->k_left;
    // thread 0 owns the four highest digits, thread kT - 1 the four lowest: a prefix sum over t = "values above"
    const uint32_t base = 4u * (kT - 1u - t);
    const uint4 h = *reinterpret_cast<const uint4*>(hist + base);
    const unsigned long long own = (unsigned long long)h.x + h.y + h.z + h.w;
    unsigned long long incl = own;
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long up = __shfl_up_sync(0xffffffffu, incl, o);
        if ((t & 31u) >= uint32_t(o))
            incl += up;
    }
    if ((t & 31u) == 31u)
        s_warp[t >> 5] = incl;
    __syncthreads();
    unsigned long long above = 0;
    for (uint32_t w = 0; w < (t >> 5); w++)
        above += s_warp[w];
    above += incl - own;
    if (above < k_left && k_left <= above + own) { // exactly one thread, when k_left values exist at all
        const unsigned int hh[4] = { h.w, h.z, h.y, h.x };
        unsigned int d = base + 3u;
        for (int j = 0; j < 4; j++, d--) {
            if (above + hh[j] >= k_left)
                break;
            above += hh[j];
        }
        s_digit = d;
        s_above = above;
        s_found = 1u;
    }
    __syncthreads();
    if (t == 0) {
        // (not found: fewer than k_left values in all -- sigma < K, nothing of this is used)
        const unsigned int d = s_found ? s_digit : 0u;
        const unsigned long long ab = s_found ? s_above : 0ull;
        st->prefix = (first ? 0ull : st->prefix) | ((unsigned long long)d << shift);
        st->k_left = k_left - ab;
        if (shift == 0u) {
            // T = prefix; k_left - ab values with count == T are still wanted
            st->need_eq = k_left - ab;
            st->n_gt = (unsigned long long)K - (k_left - ab);
            st->reorder = (st->sigma >= (unsigned long long)K) ? 1u : 0u;
        }
    }
    __syncthreads();
    *reinterpret_cast<uint4*>(hist + base) = make_uint4(0u, 0u, 0u, 0u); // ready for the next pass
}

// per tile of kRfTile values: how many have count == T
__global__ void __launch_bounds__(kRfThreads) rf_tie_count_kernel(const unsigned long long* __restrict__ raw, uint32_t nvalues,
    const RfState* __restrict__ st, unsigned int* __restrict__ tile_ties)
{
    __shared__ unsigned int s_cnt;
    const unsigned long long T = st->prefix;
    for (uint32_t tile = blockIdx.x; tile * uint64_t(kRfTile) < nvalues; tile += gridDim.x) {
        if (threadIdx.x == 0)
            s_cnt = 0;
        __syncthreads();
        unsigned int c = 0;
        for (uint32_t j = threadIdx.x; j < kRfTile; j += kRfThreads) {
            const uint64_t v = uint64_t(tile) * kRfTile + j;
            if (v < nvalues && raw[v] == T)
                c++;
        }
        for (int o = 16; o > 0; o >>= 1)
            c += __shfl_xor_sync(0xffffffffu, c, o);
        if ((threadIdx.x & 31) == 0 && c)
            atomicAdd(&s_cnt, c);
        __syncthreads();
        if (threadIdx.x == 0)
            tile_ties[tile] = s_cnt;
        __syncthreads();
    }
}

// one CTA: exclusive scan of the per-tile tie counts, in place (64-bit running sum, clamped to 2^32 - 1: only
// ranks below need_eq <= 32768 matter)
__global__ void __launch_bounds__(1024) rf_tie_scan_kernel(unsigned int* __restrict__ tile_ties, uint32_t n_tiles)
{
    __shared__ unsigned long long s_warp[32];
    __shared__ unsigned long long s_carry;
    if (threadIdx.x == 0)
        s_carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_tiles; base += 1024u) {
        const uint32_t i = base + threadIdx.x;
        const unsigned long long mine = i < n_tiles ? tile_ties[i] : 0u;
        unsigned long long incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long up = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= o)
                incl += up;
        }
        if ((threadIdx.x & 31) == 31)
            s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        unsigned long long before = s_carry;
        for (uint32_t w = 0; w < (threadIdx.x >> 5); w++)
            before += s_warp[w];
        const unsigned long long excl = before + incl - mine;
        if (i < n_tiles)
            tile_ties[i] = excl > 0xFFFFFFFFull ? 0xFFFFFFFFu : (unsigned int)excl;
        __syncthreads();
        if (threadIdx.x == 1023)
            s_carry = before + incl;
        __syncthreads();
    }
}

// candidates: {count > T} (any order) and the first need_eq of {count == T} in value order
__global__ void __launch_bounds__(kRfThreads) rf_collect_kernel(const unsigned long long* __restrict__ raw, uint32_t nvalues,
    RfState* __restrict__ st, const unsigned int* __restrict__ tile_ties, uint32_t K, unsigned long long* __restrict__ cand_count,
    uint32_t* __restrict__ cand_value)
{
    __shared__ unsigned int s_warp[kRfThreads / 32];
    if (!st->reorder)
        return;
    const unsigned long long T = st->prefix;
    const unsigned long long need_eq = st->need_eq;
    const uint32_t n_gt = uint32_t(st->n_gt);
    constexpr uint32_t kPer = kRfTile / kRfThreads; // consecutive values per thread: value order inside the tile
    for (uint32_t tile = blockIdx.x; tile * uint64_t(kRfTile) < nvalues; tile += gridDim.x) {
        const uint64_t v0 = uint64_t(tile) * kRfTile + threadIdx.x * kPer;
        unsigned long long cnt[kPer];
        unsigned int ties = 0;
#pragma unroll
        for (uint32_t j = 0; j < kPer; j++) {
            cnt[j] = v0 + j < nvalues ? raw[v0 + j] : 0ull;
            ties += (v0 + j < nvalues && cnt[j] == T);
        }
        // exclusive scan of `ties` over the CTA's threads (thread order = value order)
        unsigned int incl = ties;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int up = __shfl_up_sync(0xffffffffu, incl, o);
            if ((threadIdx.x & 31) >= o)
                incl += up;
        }
        if ((threadIdx.x & 31) == 31)
            s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        unsigned long long rank = tile_ties[tile];
        for (uint32_t w = 0; w < (threadIdx.x >> 5); w++)
            rank += s_warp[w];
        rank += incl - ties;
#pragma unroll
        for (uint32_t j = 0; j < kPer; j++) {
            if (v0 + j >= nvalues)
                continue;
            if (cnt[j] > T) {
                const unsigned int at = atomicAdd(&st->n_emitted, 1u);
                if (at < K) {
                    cand_count[at] = cnt[j];
                    cand_value[at] = uint32_t(v0 + j);
                }
            } else if (cnt[j] == T) {
                if (rank < need_eq) {
                    cand_count[n_gt + uint32_t(rank)] = T;
                    cand_value[n_gt + uint32_t(rank)] = uint32_t(v0 + j);
                }
                rank++;
            }
        }
        __syncthreads();
    }
}

// id of candidate i = number of candidates hotter than it ((-count, value) order, ans_reorder_fold.hpp:86)
__global__ void __launch_bounds__(kRfThreads) rf_rank_kernel(const RfState* __restrict__ st, uint32_t K,
    const unsigned long long* __restrict__ cand_count, const uint32_t* __restrict__ cand_value, uint32_t* __restrict__ most_frequent)
{
    __shared__ unsigned long long s_c[kRfThreads];
    __shared__ uint32_t s_v[kRfThreads];
    if (!st->reorder)
        return;
    const uint32_t i = blockIdx.x * kRfThreads + threadIdx.x;
    const unsigned long long ci = i < K ? cand_count[i] : 0ull;
    const uint32_t vi = i < K ? cand_value[i] : 0u;
    uint32_t rank = 0;
    for (uint32_t base = 0; base < K; base += kRfThreads) {
        const uint32_t j = base + threadIdx.x;
        s_c[threadIdx.x] = j < K ? cand_count[j] : 0ull;
        s_v[threadIdx.x] = j < K ? cand_value[j] : 0xFFFFFFFFu;
        __syncthreads();
        const uint32_t lim = min(uint32_t(kRfThreads), K - base);
        for (uint32_t t = 0; t < lim; t++)
            rank += (s_c[t] > ci) || (s_c[t] == ci && s_v[t] < vi);
        __syncthreads();
    }
    if (i < K)
        most_frequent[rank] = vi;
}

// mapping[v] = v + K (or v when the model keeps the identity mapping)
__global__ void __launch_bounds__(kRfThreads) rf_mapping_fill_kernel(const RfState* __restrict__ st, uint32_t K, uint32_t nvalues,
    uint32_t* __restrict__ mapping)
{
    const uint32_t bias = (!st || st->reorder) ? K : 0u; // (no state: the caller knows that the model reorders)
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nvalues; v += gridDim.x * blockDim.x)
        mapping[v] = v + bias;
}
__global__ void __launch_bounds__(kRfThreads) rf_mapping_hot_kernel(const RfState* __restrict__ st, uint32_t K,
    const uint32_t* __restrict__ most_frequent, uint32_t* __restrict__ mapping)
{
    if (st && !st->reorder)
        return;
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < K)
        mapping[most_frequent[r]] = r;
}

// folded histogram from the raw one: bins[fold(mapping[v])] += raw[v]  (ans_reorder_fold.hpp:110-115)
// A thread walks kRfRun consecutive values: all but the K hot ones map to v + K, whose folded bin changes once every
// 256 (65 536, ...) values, so the thread sums a run in a register and touches a counter when the bin changes.
// SMEM: the folded alphabet (2^(f+9) 64-bit counters) fits the CTA's shared memory.
constexpr uint32_t kRfRun = 16;
template <bool SMEM>
__global__ void __launch_bounds__(kRfThreads) rf_fold_bins_kernel(const unsigned long long* __restrict__ raw, uint32_t nvalues,
    const uint32_t* __restrict__ mapping, uint32_t f, uint32_t nbins, unsigned long long* __restrict__ bins)
{
    extern __shared__ unsigned long long s_fold[];
    if (SMEM) {
        for (uint32_t i = threadIdx.x; i < nbins; i += blockDim.x)
            s_fold[i] = 0ull;
        __syncthreads();
    }
    auto flush = [&](uint32_t bin, unsigned long long sum) {
        if (sum != 0ull && bin < nbins) {
            if (SMEM)
                atomicAdd(s_fold + bin, sum);
            else
                atomicAdd(bins + bin, sum);
        }
    };
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x * kRfRun;
    for (uint64_t v0 = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) * kRfRun; v0 < nvalues; v0 += stride) {
        uint32_t cur = 0xFFFFFFFFu;
        unsigned long long sum = 0ull;
#pragma unroll 4
        for (uint32_t j = 0; j < kRfRun; j++) {
            const uint64_t v = v0 + j;
            if (v >= nvalues)
                break;
            const unsigned long long c = raw[v];
            if (c == 0ull)
                continue;
            uint32_t k;
            const uint32_t bin = fold_map(mapping[v], f, k);
            if (bin != cur) {
                flush(cur, sum);
                cur = bin;
                sum = 0ull;
            }
            sum += c;
        }
        flush(cur, sum);
    }
    if (SMEM) {
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < nbins; i += blockDim.x)
            if (s_fold[i])
                atomicAdd(bins + i, s_fold[i]);
    }
}


// ==========================================================================================
// The fast route: exact selection WITHOUT counting every value
// ==========================================================================================
// The raw histogram costs one global atomic per cold symbol (440 M of them for config 3: 3.3 ms at the L2's atomic
// rate, and 48 ms when the hot values do not sit in the privatised head).  But the model needs exact counts only for
// values that can be among the K most frequent, and of everything else only the folded histogram of v + K.  So:
//   1. rf_sample_kernel      : 2^19 symbols at a stride -> their exact counts in an open-addressing table keyed by value
//   2. (radix select above)  : the kRfCand most frequent values of the sample = the candidates; the host builds their
//                              perfect hash
//   3. rf_count_kernel       : ONE pass over the input, everything in shared memory: a candidate -> its exact counter;
//                              any other value -> a hashed counter of the sketch (an upper bound for every value that
//                              falls into it) and the folded bin of v + K
//   4. rf_cand_rank_kernel   : the candidates ranked by (-count, value): ids, most_frequent[], their folded bins
//      rf_verify_kernel      : T = the K-th largest candidate count.  If every sketch counter is below T, no value
//                              outside the candidates can reach T -- the selection is the reference's, exactly;
//                              otherwise (or with fewer than K candidates) the caller falls back to the raw histogram.
constexpr uint32_t kRfCand = 4096;          // candidates at most
constexpr uint32_t kRfCandSlots = 8192;     // slots of their perfect hash (load <= 1/2: the host finds it at once)
constexpr uint32_t kRfCandBucketsLog2 = 11; // displacement buckets of that hash
constexpr uint32_t kRfSampleBucketsLog2 = 20; // slots of the sample's table (2^19 symbols at most: load <= 1/2)
constexpr uint32_t kRfSketchLog2 = 15;      // sketch counters per CTA (128 KiB: with the candidates and the folded bins it fills the SM)
constexpr uint32_t kRfSampleMul = 0x9E3779B1u, kRfSketchMul = 0x85EBCA6Bu;
constexpr int kRfCountThreads = 1024;

struct RfFast {
    unsigned long long t_count;   // T: count of the candidate with id K - 1
    unsigned int n_valid;         // candidates with a non-zero exact count
    unsigned int max_sketch_hi;   // (unused)
    unsigned long long max_sketch; // largest sketch counter
    unsigned int ok;              // 1: the selection is exact
    unsigned int n_cand;          // candidates collected
};

__global__ void __launch_bounds__(256) rf_sample_kernel(const uint32_t* __restrict__ in, uint64_t n, uint32_t n_sample_vec,
    unsigned long long* __restrict__ buckets /* counts, zeroed */, uint32_t* __restrict__ tags /* keys, 0xFFFFFFFF = empty */)
{
    const uint64_t n4 = n >> 2;
    const uint4* in4 = reinterpret_cast<const uint4*>(in);
    constexpr uint32_t kMask = (1u << kRfSampleBucketsLog2) - 1u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_sample_vec; i += gridDim.x * blockDim.x) {
        const uint64_t at = (uint64_t(i) * n4) / n_sample_vec;
        const uint4 q = __ldg(in4 + at);
        const uint32_t v[4] = { q.x, q.y, q.z, q.w };
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (v[j] == 0xFFFFFFFFu)
                continue; // (outside every coder's domain; the pass over the input reports it)
            uint32_t b = (v[j] * kRfSampleMul) >> (32u - kRfSampleBucketsLog2);
            for (;;) { // linear probing; the table is at most half full
                const uint32_t old = atomicCAS(tags + b, 0xFFFFFFFFu, v[j]);
                if (old == 0xFFFFFFFFu || old == v[j])
                    break;
                b = (b + 1u) & kMask;
            }
            atomicAdd(buckets + b, 1ull);
        }
    }
}

// candidates = the values whose sample count is > T, then those with count == T (T > 0) while there is room
__global__ void __launch_bounds__(256) rf_cand_collect_kernel(const unsigned long long* __restrict__ buckets, const uint32_t* __restrict__ tags,
    const RfState* __restrict__ st, bool ties, RfFast* __restrict__ fs, uint32_t* __restrict__ cand)
{
    const unsigned long long T = st->prefix;
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < (1u << kRfSampleBucketsLog2); b += gridDim.x * blockDim.x) {
        const unsigned long long c = buckets[b];
        if (ties ? (c == T && T != 0ull) : (c > T)) {
            const unsigned int at = atomicAdd(&fs->n_cand, 1u);
            if (at < kRfCand)
                cand[at] = tags[b];
        }
    }
}

// the one pass over the input (see above).  cand_tags / cand_disp: perfect hash of the candidates (host-built).
__global__ void __launch_bounds__(kRfCountThreads, 1) rf_count_kernel(const uint32_t* __restrict__ in, uint64_t n, uint32_t f,
    const uint32_t* __restrict__ cand_tags, const uint16_t* __restrict__ cand_disp, uint32_t mul, uint32_t nfolded,
    unsigned long long* __restrict__ cand_count, unsigned long long* __restrict__ sketch, unsigned long long* __restrict__ bins,
    uint32_t* __restrict__ max_value)
{
    extern __shared__ uint32_t s_rf[];
    uint32_t* s_tags = s_rf;                                   // kRfCandSlots
    uint32_t* s_cnt = s_tags + kRfCandSlots;                   // kRfCandSlots
    uint32_t* s_sketch = s_cnt + kRfCandSlots;                 // 2^kRfSketchLog2
    uint32_t* s_fold = s_sketch + (1u << kRfSketchLog2);       // nfolded
    uint16_t* s_disp = reinterpret_cast<uint16_t*>(s_fold + nfolded); // 2^kRfCandBucketsLog2
    for (uint32_t i = threadIdx.x; i < kRfCandSlots; i += blockDim.x) {
        s_tags[i] = __ldg(cand_tags + i);
        s_cnt[i] = 0u;
    }
    for (uint32_t i = threadIdx.x; i < (1u << kRfSketchLog2); i += blockDim.x)
        s_sketch[i] = 0u;
    for (uint32_t i = threadIdx.x; i < nfolded; i += blockDim.x)
        s_fold[i] = 0u;
    for (uint32_t i = threadIdx.x; i < (1u << kRfCandBucketsLog2); i += blockDim.x)
        s_disp[i] = __ldg(cand_disp + i);
    __syncthreads();
    const uint32_t a_tags = uint32_t(__cvta_generic_to_shared(s_tags)), a_cnt = uint32_t(__cvta_generic_to_shared(s_cnt));
    const uint32_t a_sketch = uint32_t(__cvta_generic_to_shared(s_sketch)), a_fold = uint32_t(__cvta_generic_to_shared(s_fold));
    const uint32_t K = 1u << (f + 7u);
    uint32_t local_max = 0;
    auto count = [&](uint32_t v) {
        const uint32_t h = v * mul;
        const uint32_t d = s_disp[h >> (32u - kRfCandBucketsLog2)];
        const uint32_t s = (h + d) & (kRfCandSlots - 1u);
        local_max = max(local_max, v);
        if (s_tags[s] == v) {
            red_shared_inc(a_cnt + (s << 2));
        } else {
            red_shared_inc(a_sketch + (((v * kRfSketchMul) >> (32u - kRfSketchLog2)) << 2));
            uint32_t k;
            const uint32_t bin = fold_map(v + K, f, k);
            if (bin < nfolded)
                red_shared_inc(a_fold + (bin << 2));
        }
        (void)a_tags;
    };
    const uint64_t n4 = n >> 2;
    const uint4* in4 = reinterpret_cast<const uint4*>(in);
    // a CTA's counters are 32 bits wide: it takes at most 2^32 - 1 symbols (the host sizes the grid accordingly)
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    for (; i + stride < n4; i += 2 * stride) {
        const uint4 q0 = __ldcs(in4 + i);
        const uint4 q1 = __ldcs(in4 + i + stride);
        count(q0.x), count(q0.y), count(q0.z), count(q0.w);
        count(q1.x), count(q1.y), count(q1.z), count(q1.w);
    }
    if (i < n4) {
        const uint4 q0 = __ldcs(in4 + i);
        count(q0.x), count(q0.y), count(q0.z), count(q0.w);
    }
    if (blockIdx.x == 0 && threadIdx.x < 3) {
        const uint64_t t = (n4 << 2) + threadIdx.x;
        if (t < n)
            count(in[t]);
    }
    for (int o = 16; o > 0; o >>= 1)
        local_max = max(local_max, __shfl_xor_sync(0xffffffffu, local_max, o));
    if ((threadIdx.x & 31) == 0 && local_max != 0)
        atomicMax(max_value, local_max);
    __syncthreads();
    for (uint32_t j = threadIdx.x; j < kRfCandSlots; j += blockDim.x)
        if (s_cnt[j])
            atomicAdd(cand_count + j, (unsigned long long)s_cnt[j]);
    for (uint32_t j = threadIdx.x; j < (1u << kRfSketchLog2); j += blockDim.x)
        if (s_sketch[j])
            atomicAdd(sketch + j, (unsigned long long)s_sketch[j]);
    for (uint32_t j = threadIdx.x; j < nfolded; j += blockDim.x)
        if (s_fold[j])
            atomicAdd(bins + j, (unsigned long long)s_fold[j]);
}

// candidates (candidate i sits in slot cand_slot[i]: value tags[slot], exact count cand_count[slot]) ranked by
// (-count, value), one warp per candidate: the first K get the ids
__global__ void __launch_bounds__(kRfThreads) rf_cand_rank_kernel(const uint32_t* __restrict__ tags, const unsigned long long* __restrict__ cand_count,
    const uint16_t* __restrict__ cand_slot, uint32_t n_cand, uint32_t K, uint32_t f, uint32_t nfolded, uint32_t* __restrict__ most_frequent,
    unsigned long long* __restrict__ bins, RfFast* __restrict__ fs)
{
    const uint32_t i = (blockIdx.x * kRfThreads + threadIdx.x) >> 5;
    if (i >= n_cand)
        return;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t si = cand_slot[i];
    const uint32_t vi = tags[si];
    const unsigned long long ci = cand_count[si];
    uint32_t rank = 0;
    for (uint32_t j = lane; j < n_cand; j += 32u) {
        const uint32_t sj = cand_slot[j];
        const unsigned long long cj = cand_count[sj];
        rank += (cj > ci) || (cj == ci && tags[sj] < vi);
    }
    for (int o = 16; o > 0; o >>= 1)
        rank += __shfl_xor_sync(0xffffffffu, rank, o);
    if (lane != 0 || ci == 0ull)
        return; // (candidates that never occur rank behind every one that does: they take no id)
    atomicAdd(&fs->n_valid, 1u);
    if (rank < K) {
        most_frequent[rank] = vi;
        atomicAdd(bins + rank, ci); // fold(id) = id below the threshold
        if (rank == K - 1u)
            fs->t_count = ci;
    } else {
        uint32_t k;
        const uint32_t bin = fold_map(vi + K, f, k);
        if (bin < nfolded)
            atomicAdd(bins + bin, ci);
    }
}

__global__ void __launch_bounds__(1024) rf_verify_kernel(const unsigned long long* __restrict__ sketch, uint32_t K, RfFast* __restrict__ fs)
{
    __shared__ unsigned long long s_max[32];
    unsigned long long m = 0;
    for (uint32_t j = threadIdx.x; j < (1u << kRfSketchLog2); j += blockDim.x)
        m = max(m, sketch[j]);
    for (int o = 16; o > 0; o >>= 1)
        m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0)
        s_max[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 32; w++)
            m = max(m, s_max[w]);
        fs->max_sketch = m;
        // strictly below T: a value outside the candidates with count == T could win a tie on the value
        fs->ok = (fs->n_valid >= K && fs->t_count != 0ull && m < fs->t_count) ? 1u : 0u;
    }
}

} // namespace b200ans
