// Test program (built with -fsanitize=address,undefined by tests/test_host_robustness.py): the bounded prelude reader
// of host/prelude.hpp on well-formed, truncated and hostile bytes.  Every buffer is heap-allocated at its exact size, so
// one byte read past the end is an ASan failure.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <random>
#include <vector>

#include "prelude.hpp"

using namespace b200ans;

static bool load(const std::vector<uint8_t>& bytes, size_t max_symbols, std::vector<uint32_t>& out, size_t* used)
{
    // exact-size heap copy: the sanitizer catches any over-read
    std::unique_ptr<uint8_t[]> buf(new uint8_t[bytes.size() ? bytes.size() : 1]);
    if (!bytes.empty())
        std::memcpy(buf.get(), bytes.data(), bytes.size());
    return ans_load_interp_checked(buf.get(), buf.get() + bytes.size(), max_symbols, 28u, out, used);
}

int main()
{
    std::mt19937 gen(7);
    int accepted = 0, rejected = 0;
    for (int round = 0; round < 40; round++) {
        // a well-formed model: sigma symbols, a power-of-two frame
        const size_t sigma = 2 + gen() % 1500;
        const unsigned lg = 11 + gen() % 8;
        std::vector<uint32_t> nf(sigma, 0);
        uint64_t left = uint64_t(1) << lg;
        for (size_t s = 0; s + 1 < sigma && left > 1; s++) {
            if (gen() % 4 == 0)
                continue; // absent symbol
            const uint32_t share = uint32_t(1 + gen() % std::max<uint64_t>(1, std::min<uint64_t>(left - 1, 4 * (uint64_t(1) << lg) / sigma)));
            nf[s] = share;
            left -= share;
        }
        nf[sigma - 1] = uint32_t(left);
        std::vector<uint8_t> bytes(sigma * 8 + 64);
        uint8_t* p = bytes.data();
        ans_serialize_interp(nf, size_t(1) << lg, p);
        bytes.resize(size_t(p - bytes.data()));

        std::vector<uint32_t> back;
        size_t used = 0;
        if (!load(bytes, 1 << 20, back, &used) || back != nf || used != bytes.size()) {
            std::printf("round %d: a well-formed prelude was refused or misread\n", round);
            return 1;
        }
        accepted++;
        // alphabet limit
        if (load(bytes, sigma - 1, back, nullptr)) {
            std::printf("round %d: alphabet limit ignored\n", round);
            return 1;
        }
        // every truncation must be refused (or, when only padding bits were cut, decode to the same model) without over-reading
        for (size_t cut = 0; cut < bytes.size(); cut += 1 + cut / 16) {
            std::vector<uint8_t> t(bytes.begin(), bytes.begin() + cut);
            std::vector<uint32_t> got;
            if (load(t, 1 << 20, got, nullptr)) {
                if (got != nf) {
                    uint64_t sum = 0;
                    for (uint32_t v : got)
                        sum += v;
                    (void)sum; // a different model is acceptable here: the caller checks the frame sum; no over-read happened
                }
            } else {
                rejected++;
            }
        }
        // hostile headers: huge alphabets, over-long vbyte, big log2
        std::vector<std::vector<uint8_t>> hostile = {
            { 0xFF, 0xFF, 0xFF, 0x7F, 20 },             // max_sym ~ 2^28 with nothing behind it
            { 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 12 },  // vbyte longer than a u32
            { 0x05, 63, 1, 2, 3, 4 },                    // log2 M = 63
            { 0x80 },                                    // vbyte cut off
            {},
        };
        for (auto& h : hostile) {
            std::vector<uint32_t> got;
            if (load(h, size_t(1) << 28, got, nullptr)) {
                std::printf("hostile header accepted\n");
                return 1;
            }
            rejected++;
        }
        // random garbage of random length
        std::vector<uint8_t> junk(gen() % 200);
        for (auto& b : junk)
            b = uint8_t(gen());
        std::vector<uint32_t> got;
        load(junk, 4096, got, nullptr); // either way: no crash, no over-read
    }
    std::printf("ok accepted=%d rejected=%d\n", accepted, rejected);
    return 0;
}
