// A driver in the shape of the reference's src/table_efficiency.cpp (run<t_compressor>, :64-121): best-of-N timing of
// T::encode / T::decode over synthetic inputs, REQUIRE_EQUAL-style verification -- instantiated with the b200ans
// adaptor structs instead of the reference's.  Shows that the adaptors are drop-ins for the `run<>` templates and is
// run by tests/test_cpp_adaptors.py on the GPU box.  Inputs follow src/generate_inputs.cpp:70-117 (mt19937 seed 0).
//
// build: g++ -std=c++17 -O2 -I include tests/cpp/table_efficiency_b200.cpp -L ans_large_alphabet_b200 -lb200ans
#include <algorithm>
#include <chrono>
#include <cstring>
#include <iostream>
#include <limits>
#include <random>

#include "b200ans_methods.hpp"

const int NUM_RUNS = 3;

static double compute_ips(size_t n, size_t ns) { return double(n) / (double(ns) / 1e9) / 1e6; } // million ints / s

template <class t_compressor> bool run(const std::vector<std::vector<uint32_t>>& inputs)
{
    bool ok = true;
    for (const auto& input : inputs) {
        std::vector<uint8_t> encoded_data(input.size() * 8 + 4096);
        std::vector<uint8_t> tmp_buf(16);
        size_t encoded_bytes = 0;
        size_t enc_ns = std::numeric_limits<size_t>::max();
        for (int i = 0; i < NUM_RUNS; i++) {
            auto t0 = std::chrono::high_resolution_clock::now();
            encoded_bytes = t_compressor::encode(input.data(), input.size(), encoded_data.data(), encoded_data.size(), tmp_buf.data());
            auto t1 = std::chrono::high_resolution_clock::now();
            enc_ns = std::min(size_t((t1 - t0).count()), enc_ns);
        }
        encoded_data.resize(encoded_bytes);
        std::vector<uint32_t> recover(input.size());
        size_t dec_ns = std::numeric_limits<size_t>::max();
        for (int i = 0; i < NUM_RUNS; i++) {
            auto t0 = std::chrono::high_resolution_clock::now();
            t_compressor::decode(encoded_data.data(), encoded_data.size(), recover.data(), recover.size(), tmp_buf.data());
            auto t1 = std::chrono::high_resolution_clock::now();
            dec_ns = std::min(size_t((t1 - t0).count()), dec_ns);
        }
        const bool same = std::memcmp(input.data(), recover.data(), input.size() * 4) == 0; // REQUIRE_EQUAL (cutil.hpp:30-50)
        ok = ok && same;
        std::printf("%-12s n=%9zu  bytes=%10zu  bpi=%7.4f  enc=%9.2f Mint/s  dec=%9.2f Mint/s  %s\n", t_compressor::name().c_str(),
            input.size(), encoded_bytes, 8.0 * encoded_bytes / input.size(), compute_ips(input.size(), enc_ns),
            compute_ips(input.size(), dec_ns), same ? "OK" : "MISMATCH");
    }
    return ok;
}

int main(int argc, char** argv)
{
    const size_t n = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 2000003;
    std::vector<std::vector<uint32_t>> inputs;
    {
        std::mt19937 gen(0);
        std::geometric_distribution<uint32_t> d(0.01); // generate_inputs.cpp:97-98
        std::vector<uint32_t> v(n);
        for (auto& x : v)
            x = d(gen);
        inputs.push_back(std::move(v));
    }
    {
        std::mt19937 gen(0);
        std::uniform_int_distribution<uint32_t> d(0, (1u << 20) - 1); // generate_inputs.cpp:88-95 (uniform 20 bit)
        std::vector<uint32_t> v(n);
        for (auto& x : v)
            x = d(gen) >> (gen() % 12); // skewed towards small values so that folding has something to do
        inputs.push_back(std::move(v));
    }
    bool ok = true;
    ok = run<ANSint_b200>(inputs) && ok;
    ok = run<ANSmsb_b200>(inputs) && ok;
    ok = run<ANSfold_b200<1>>(inputs) && ok;
    ok = run<ANSfold_b200<2>>(inputs) && ok;
    ok = run<ANSsint_b200<20>>({ inputs[0] }) && ok;  // benchmark.cpp:165-172
    ok = run<ANSsmsb_b200<80>>(inputs) && ok;         // benchmark.cpp:156-163
    ok = run<ANSrfold_b200<1>>({ inputs[1] }) && ok; // rfold needs sigma >= 2^(f+7)

    // the ans_util.hpp model API: histogram -> adjust_freqs -> serialize -> load
    std::vector<uint64_t> freqs(300, 0);
    for (uint32_t x : inputs[0])
        freqs[std::min<uint32_t>(x, 299)]++;
    uint32_t largest = 0;
    for (uint32_t s = 0; s < freqs.size(); s++)
        if (freqs[s])
            largest = s;
    std::vector<uint32_t> nfreqs = b200::adjust_freqs(freqs, largest, true);
    uint64_t frame = 0;
    for (uint32_t f : nfreqs)
        frame += f;
    std::vector<uint8_t> prelude(nfreqs.size() * 8 + 128);
    uint8_t* p = prelude.data();
    const size_t interp_bytes = b200::ans_serialize_interp(nfreqs, frame, p);
    const std::vector<uint32_t> back = b200::ans_load_interp(prelude.data());
    const bool model_ok = b200::is_power_of_two(frame) && back == nfreqs && interp_bytes > 0 && size_t(p - prelude.data()) > interp_bytes;
    std::printf("model API: sigma=%zu frame=%llu prelude=%zu bytes (interp %zu)  %s\n", nfreqs.size(), (unsigned long long)frame,
        size_t(p - prelude.data()), interp_bytes, model_ok ? "OK" : "MISMATCH");
    ok = ok && model_ok;

    // scale_freqs (ans_util.hpp:77-95): one fit of the same counts to the frame adjust_freqs chose reproduces its answer
    {
        std::vector<std::pair<uint64_t, uint32_t>> order;
        uint64_t total = 0;
        for (uint32_t s = 0; s <= largest; s++)
            if (freqs[s]) {
                order.emplace_back(freqs[s], s);
                total += freqs[s];
            }
        std::sort(order.begin(), order.end());
        std::vector<uint32_t> mapping;
        for (auto& pr : order)
            mapping.push_back(pr.second);
        std::vector<uint32_t> S(largest + 1, 0);
        std::vector<uint64_t> F(freqs.begin(), freqs.begin() + largest + 1);
        const bool failed = b200::scale_freqs(S, F, mapping, int64_t(frame), mapping.size(), total);
        const bool sf_ok = !failed && S == nfreqs;
        std::printf("scale_freqs: %s\n", sf_ok ? "OK" : "MISMATCH");
        ok = ok && sf_ok;
    }

    // the model structs (ans_fold.hpp:67-125,177-235 and siblings): create -> serialize -> load, fields as in the reference
    {
        auto enc = b200::ans_fold_encode<1>::create(inputs[1].data(), inputs[1].size());
        std::vector<uint8_t> pre(enc.nfreqs.size() * 8 + 128);
        uint8_t* q = pre.data();
        const size_t ret = enc.serialize(q);
        size_t used = 0;
        auto dec = b200::ans_fold_decode<1>::load(pre.data(), size_t(q - pre.data()), &used);
        uint64_t sum = 0;
        for (uint32_t v : enc.nfreqs)
            sum += v;
        bool m_ok = dec.nfreqs == enc.nfreqs && sum == enc.frame_size && enc.lower_bound == 16 * enc.frame_size
            && enc.initial_state() == enc.lower_bound && dec.frame_mask == enc.frame_size - 1 && (uint64_t(1) << dec.frame_log2) == dec.frame_size
            && used == size_t(q - pre.data()) && ret > 0 && ret < used && enc.table.size() == enc.nfreqs.size()
            && enc.table.back().base + enc.table.back().freq == enc.frame_size;
        // the struct's model is the one the one-shot coder uses: same prelude inside the container
        std::vector<uint8_t> blob(inputs[1].size() * 8 + 4096);
        const size_t nb = enc.encode(blob.data(), blob.size(), inputs[1].data(), inputs[1].size());
        b200ans_container info;
        m_ok = m_ok && b200ans_container_info(blob.data(), nb, &info) == B200ANS_OK && info.prelude_bytes == used
            && std::memcmp(blob.data() + B200ANS_HEADER_BYTES, pre.data(), used) == 0;
        std::vector<uint32_t> out(inputs[1].size());
        b200::ans_fold_decompress<1>(out.data(), out.size(), blob.data(), nb);
        m_ok = m_ok && out == inputs[1];
        std::printf("model structs ANSfold-1: sigma'=%zu frame=%llu prelude=%zu  %s\n", enc.nfreqs.size(),
            (unsigned long long)enc.frame_size, used, m_ok ? "OK" : "MISMATCH");
        ok = ok && m_ok;

        auto renc = b200::ans_reorder_fold_encode<1>::create(inputs[1].data(), inputs[1].size());
        std::vector<uint8_t> rpre(renc.nfreqs.size() * 8 + 128 + 4 * renc.most_frequent.size());
        uint8_t* rq = rpre.data();
        renc.serialize(rq);
        auto rdec = b200::ans_reorder_fold_decode<1>::load(rpre.data(), size_t(rq - rpre.data()));
        const bool r_ok = rdec.nfreqs == renc.nfreqs && rdec.most_frequent == renc.most_frequent && renc.most_frequent.size() == 256
            && renc.sigma >= 256;
        auto ienc = b200::ans_int_encode::create(inputs[0].data(), inputs[0].size());
        auto menc = b200::ans_msb_encode::create(inputs[1].data(), inputs[1].size());
        const bool i_ok = b200::is_power_of_two(ienc.frame_size) && b200::is_power_of_two(menc.frame_size) && menc.nfreqs.size() <= 1280;
        std::printf("model structs ANSrfold-1 / ANS / ANSmsb: %s\n", (r_ok && i_ok) ? "OK" : "MISMATCH");
        ok = ok && r_ok && i_ok;
    }
    // ans_byte (ans_byte.hpp:236,279): the byte coder of the vbyteANS / streamvbyteANS baselines
    {
        std::vector<uint8_t> bytes(inputs[0].size());
        for (size_t i = 0; i < bytes.size(); i++)
            bytes[i] = uint8_t(std::min<uint32_t>(inputs[0][i], 255));
        std::vector<uint8_t> blob(bytes.size() * 4 + 4096), back(bytes.size());
        const size_t nb = b200::ans_byte_compress(blob.data(), blob.size(), bytes.data(), bytes.size());
        b200::ans_byte_decompress(back.data(), back.size(), blob.data(), nb);
        const bool b_ok = back == bytes && nb < bytes.size();
        std::printf("ans_byte: %zu -> %zu bytes  %s\n", bytes.size(), nb, b_ok ? "OK" : "MISMATCH");
        ok = ok && b_ok;
    }
    std::printf(ok ? "ALL OK\n" : "FAILED\n");
    return ok ? 0 : 1;
}
