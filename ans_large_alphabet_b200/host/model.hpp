// b200ans host-side model utilities: histogram -> normalised frequencies.
//
// Drop-in for the reference's ans_util.hpp model API (same free-function names, argument
// meaning and results, bit for bit):
//   adjust_freqs        <- /root/reference/include/ans_util.hpp:100-157
//   scale_freqs         <- ans_util.hpp:77-95
//   next_power_of_two   <- ans_util.hpp:65-72
//   is_power_of_two     <- ans_util.hpp:74
//   entropy             <- util.hpp:271-282
//   cross_entropy       <- util.hpp:284-298
// This step stays on the host: the histogram it consumes is
// tiny for the folded coders and the accept/reject decision hangs on glibc's log2.
//
// "Bit for bit" includes the reference's quirks:
//  * symbols are scaled rarest-first against a *running* ratio;
//  * cross_entropy sums both distributions into a 32-bit signed accumulator;
//  * the three multiply-adds below are fused, as GCC's default -ffp-contract=fast does to
//    the reference on every FMA-capable x86-64 (its CMake build uses -march=native).
#pragma once

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <utility>
#include <array>
#include <vector>

namespace b200ans {

inline bool is_power_of_two(uint64_t x) { return x != 0 && (x & (x - 1)) == 0; }

// smallest power of two strictly greater than x's top bit, i.e. 2^(floor(log2 x) + 1); 1 for x = 0.
// (callers only use it for non-powers of two, where this is "round up")
inline uint64_t next_power_of_two(uint64_t x)
{
    if (x == 0)
        return 1;
    return uint64_t(2) << (63 - __builtin_clzll(x));
}

template <class X> inline double entropy(const std::vector<X>& freqs, size_t freqs_sum)
{
    const double total = double(freqs_sum);
    double acc = 0.0;
    for (const X& f : freqs) {
        if (f == 0)
            continue;
        const double p = double(f) / total;
        acc = std::fma(p, std::log2(p), acc);
    }
    return -acc;
}

namespace detail {
    // std::accumulate(first, last, 0): the accumulator is an `int`
    template <class V> inline double int_accumulated_sum(const V& v)
    {
        int32_t acc = 0;
        for (const auto& e : v)
            acc = int32_t(uint32_t(uint64_t(int64_t(acc)) + uint64_t(e)));
        return double(acc);
    }
}

template <class X, class Y> inline double cross_entropy(const std::vector<X>& P, const std::vector<Y>& Q)
{
    const double n = detail::int_accumulated_sum(P);
    const double m = detail::int_accumulated_sum(Q);
    double acc = 0.0;
    for (size_t i = 0; i < P.size(); i++) {
        if (P[i] == 0 || Q[i] == 0) // P is zero beyond Q's end, so Q[i] is never read out of range
            continue;
        const double p = double(P[i]) / n;
        const double q = double(Q[i]) / m;
        acc = std::fma(p, std::log2(q), acc);
    }
    return -acc;
}

namespace detail {
    // cross_entropy(P, Q) for normalised frequencies Q: log2(double(Q[i]) / m) is the same call with the same
    // argument for every symbol that got the same share, so it is evaluated once per distinct small Q[i]
    // (same bits, fewer libm calls: the tail of a large alphabet shares a handful of values).
    // `prob` (optional): P[i] / n for every i, computed once by the caller -- the same doubles this function would
    // compute (adjust_freqs evaluates up to a dozen candidate frames against the same counts)
    inline double cross_entropy_shared_log(const std::vector<uint64_t>& P, const std::vector<uint32_t>& Q, const std::vector<double>* prob = nullptr)
    {
        constexpr uint32_t kMemo = 4096;
        static thread_local double memo[kMemo];
        static thread_local uint32_t stamp[kMemo];
        static thread_local uint32_t epoch = 0;
        if (++epoch == 0) {
            std::fill(stamp, stamp + kMemo, 0u);
            epoch = 1;
        }
        const double n = prob ? 0.0 : int_accumulated_sum(P);
        const double m = int_accumulated_sum(Q);
        double acc = 0.0;
        for (size_t i = 0; i < P.size(); i++) {
            if (P[i] == 0 || Q[i] == 0)
                continue;
            const double p = prob ? (*prob)[i] : double(P[i]) / n;
            const uint32_t share = Q[i];
            double lg;
            if (share < kMemo) {
                if (stamp[share] != epoch) {
                    memo[share] = std::log2(double(share) / m);
                    stamp[share] = epoch;
                }
                lg = memo[share];
            } else {
                lg = std::log2(double(share) / m);
            }
            acc = std::fma(p, lg, acc);
        }
        return -acc;
    }
}

// Fit F (symbols visited in `mapping` order, rarest first) into a frame of exactly M slots.
// Returns true when the fit FAILED (slots left over or overdrawn), like the reference.
inline bool scale_freqs(std::vector<uint32_t>& S, const std::vector<uint64_t>& F, std::vector<uint32_t>& mapping,
    int64_t M, size_t sigma, size_t freq_sum)
{
    for (size_t rank = 0; rank < sigma; rank++) {
        const uint32_t sym = mapping[rank];
        const double ratio = double(M) / double(freq_sum);
        uint32_t share = uint32_t(std::fma(ratio, double(F[sym]), 0.5));
        if (share == 0)
            share = 1;
        S[sym] = share;
        M -= share;
        freq_sum -= F[sym];
        if (M < 0)
            break;
    }
    return M != 0;
}

namespace detail {
    // scale_freqs for the frames M0, 2·M0, ... 2^(K-1)·M0 in one sweep (K independent dependency chains).
    // failed[j] and S[j] are what scale_freqs(S, F, mapping, M0 << j, ...) gives on a zeroed S; entries a failed fit
    // leaves behind are never looked at.
    template <int K>
    inline void scale_freqs_batch(std::vector<uint32_t> (&S)[K], bool (&failed)[K], const std::vector<uint64_t>& F,
        const std::vector<uint32_t>& mapping, int64_t M0, size_t sigma, size_t freq_sum)
    {
        int64_t M[K];
        bool live[K];
        for (int j = 0; j < K; j++) {
            M[j] = M0 << j;
            live[j] = true;
        }
        for (size_t rank = 0; rank < sigma; rank++) {
            const uint32_t sym = mapping[rank];
            const double f = double(F[sym]);
            const double rest = double(freq_sum);
            for (int j = 0; j < K; j++) {
                if (!live[j])
                    continue;
                const double ratio = double(M[j]) / rest;
                uint32_t share = uint32_t(std::fma(ratio, f, 0.5));
                if (share == 0)
                    share = 1;
                S[j][sym] = share;
                M[j] -= share;
                live[j] = M[j] >= 0;
            }
            freq_sum -= F[sym];
        }
        for (int j = 0; j < K; j++)
            failed[j] = M[j] != 0;
    }
}

// Smallest power-of-two frame whose cross entropy is within H_approx/1000 of the entropy
// (or, with require_u16, the last frame before some normalised frequency reaches 65535).
inline std::vector<uint32_t> adjust_freqs(
    const std::vector<uint64_t>& freqs, uint32_t largest_sym, bool require_u16, uint32_t H_approx = 1)
{
    size_t sigma = 0, freq_sum = 0;
    for (uint64_t f : freqs) {
        freq_sum += f;
        sigma += (f != 0);
    }
    size_t frame = is_power_of_two(sigma) ? sigma : size_t(next_power_of_two(sigma));

    // symbols by (frequency, symbol) ascending -- std::sort of pairs in the reference; the order is total, so any
    // sort gives the same sequence.  Packed 64-bit keys when they fit (faster compares), pairs otherwise.
    std::vector<uint32_t> order(sigma);
    uint64_t top_freq = 0;
    for (uint64_t f : freqs)
        top_freq = std::max(top_freq, f);
    if (top_freq < (uint64_t(1) << 40) && freqs.size() <= (size_t(1) << 24)) {
        std::vector<uint64_t> keys;
        keys.reserve(sigma);
        for (size_t s = 0; s < freqs.size(); s++)
            if (freqs[s] != 0)
                keys.push_back((freqs[s] << 24) | uint64_t(s));
        std::sort(keys.begin(), keys.end());
        for (size_t i = 0; i < sigma; i++)
            order[i] = uint32_t(keys[i] & 0xFFFFFFu);
    } else {
        std::vector<std::pair<uint64_t, uint32_t>> by_freq;
        by_freq.reserve(sigma);
        for (size_t s = 0; s < freqs.size(); s++)
            if (freqs[s] != 0)
                by_freq.emplace_back(freqs[s], uint32_t(s));
        std::sort(by_freq.begin(), by_freq.end());
        for (size_t i = 0; i < sigma; i++)
            order[i] = by_freq[i].second;
    }

    const double H = entropy(freqs, freq_sum);
    const double limit = H * (1.0 + double(H_approx) / double(1000));
    // The reference tries frame, 2·frame, 4·frame, ... one after the other.  Each fit is a serial chain
    // (divide -> round -> subtract per symbol) that does not depend on the other frames, so kBatch consecutive
    // frames are fitted in one interleaved sweep and then examined in the reference's order.
    constexpr int kBatch = 8;
    std::vector<uint32_t> cand[kBatch];
    for (auto& v : cand)
        v.assign(size_t(largest_sym) + 1, 0);
    std::vector<uint32_t> last_good(size_t(largest_sym) + 1, 0);
    std::vector<double> prob;
    for (;;) {
        bool failed[kBatch];
        detail::scale_freqs_batch<kBatch>(cand, failed, freqs, order, int64_t(frame), sigma, freq_sum);
        for (int j = 0; j < kBatch; j++) {
            if (failed[j])
                continue;
            std::vector<uint32_t>& cur = cand[j];
            const uint32_t top = *std::max_element(cur.begin(), cur.end());
            if (require_u16 && top >= std::numeric_limits<uint16_t>::max()) // (the reference computes XH first; it has no effect here)
                return last_good;
            if (prob.empty()) { // P[i] / n, once for all candidate frames
                const double n = detail::int_accumulated_sum(freqs);
                prob.resize(freqs.size());
                for (size_t i = 0; i < freqs.size(); i++)
                    prob[i] = double(freqs[i]) / n;
            }
            const double XH = detail::cross_entropy_shared_log(freqs, cur, &prob);
            if (XH < limit)
                return std::move(cur);
            last_good = cur;
        }
        frame <<= kBatch;
    }
}


// ans_byte.hpp:41-97: normalise byte counts to a frame of min(4096, 64 x distinct symbols) rounded up to a power of two.
// Scale down with a fudge factor that shrinks by 0.01 per try until the truncated shares fit, then hand the slack out
// from the top symbol down in proportion, whatever is left to the largest share.  Same double operations in the same
// order as the reference (no expression here can be contracted into an FMA).
inline std::array<uint16_t, 256> byte_adjust_freqs(const uint64_t* freqs /* 256 */)
{
    std::array<uint16_t, 256> adj {};
    size_t uniq = 0, initial_sum = 0;
    for (size_t i = 0; i < 256; i++) {
        initial_sum += freqs[i];
        uniq += freqs[i] != 0;
    }
    size_t target = uniq * 64;
    if (target > 4096)
        target = 4096;
    if (!is_power_of_two(target))
        target = size_t(next_power_of_two(target));
    const double c = double(target) / double(initial_sum);
    size_t cur = ~size_t(0);
    double fudge = 1.0;
    while (cur > target) {
        fudge -= 0.01;
        cur = 0;
        for (size_t sym = 0; sym < 256; sym++) {
            adj[sym] = uint16_t(fudge * double(freqs[sym]) * c);
            if (adj[sym] == 0 && freqs[sym] != 0)
                adj[sym] = 1;
            cur += adj[sym];
        }
    }
    int64_t excess = int64_t(target) - int64_t(cur);
    for (size_t sym = 0; sym < 256; sym++) {
        const uint16_t ncnt = adj[255 - sym];
        if (ncnt == 0)
            continue;
        const double ratio = double(excess) / double(cur);
        size_t adder = size_t(ratio * double(ncnt));
        if (adder > size_t(excess))
            adder = size_t(excess);
        excess -= int64_t(adder);
        cur -= ncnt;
        adj[255 - sym] = uint16_t(adj[255 - sym] + adder);
    }
    if (excess != 0) {
        size_t max_freq = 0, max_sym = 0;
        for (size_t sym = 0; sym < 256; sym++)
            if (adj[sym] > max_freq) {
                max_freq = adj[sym];
                max_sym = sym;
            }
        adj[max_sym] = uint16_t(adj[max_sym] + excess);
    }
    return adj;
}

} // namespace b200ans
