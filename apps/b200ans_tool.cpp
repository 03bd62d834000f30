// b200ans_tool -- file-level driver over the b200ans adaptors, Boost-free.
//
// Mirrors the reference's drivers for the ANS methods of this path (mpetri/ans-large-alphabet):
//   efficiency     src/table_efficiency.cpp:64-121,123-182    best-of-NUM_RUNS million ints/s, same row layout
//   effectiveness  src/table_effectiveness.cpp:64-90,92-152   bits per integer, same row layout
//   compress / decompress                                      the on-disk b200ans container
// Input handling follows the reference: `-i` is a directory (all *.u32, or *.txt with -t, sorted by name) or a single
// file (table_efficiency.cpp:129-160); .u32 = raw little-endian uint32 (util.hpp:172-192), .txt = one decimal per line
// (util.hpp:160-170).
//
// build: g++ -std=c++17 -O2 -I include apps/b200ans_tool.cpp -L ans_large_alphabet_b200 -lb200ans -o apps/b200ans_tool
#include <dirent.h>
#include <sys/stat.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <iostream>
#include <limits>
#include <string>
#include <vector>

#include "b200ans_methods.hpp"

namespace {

const int NUM_RUNS = 5; // table_efficiency.cpp:32

[[noreturn]] void die(const std::string& msg)
{
    std::fprintf(stderr, "b200ans_tool: %s\n", msg.c_str());
    std::exit(EXIT_FAILURE);
}

bool is_regular_file(const std::string& p)
{
    struct stat st;
    return stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode);
}

bool ends_with(const std::string& s, const std::string& suffix)
{
    return s.size() >= suffix.size() && s.compare(s.size() - suffix.size(), suffix.size(), suffix) == 0;
}

std::vector<uint32_t> read_file_u32(const std::string& name)
{
    FILE* f = std::fopen(name.c_str(), "rb");
    if (!f)
        die("opening file '" + name + "' failed");
    std::fseek(f, 0, SEEK_END);
    const long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (bytes < 0 || bytes % 4 != 0)
        die("reading file content failed: file size % 32bit != 0");
    std::vector<uint32_t> v(size_t(bytes) / 4);
    if (std::fread(v.data(), 4, v.size(), f) != v.size())
        die("reading file content failed: " + name);
    std::fclose(f);
    return v;
}

std::vector<uint32_t> read_file_text(const std::string& name)
{
    FILE* f = std::fopen(name.c_str(), "r");
    if (!f)
        die("opening file '" + name + "' failed");
    std::vector<uint32_t> v;
    uint32_t num;
    while (std::fscanf(f, "%u\n", &num) == 1)
        v.push_back(num);
    std::fclose(f);
    return v;
}

std::vector<uint8_t> read_file_u8(const std::string& name)
{
    FILE* f = std::fopen(name.c_str(), "rb");
    if (!f)
        die("opening file '" + name + "' failed");
    std::fseek(f, 0, SEEK_END);
    const long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> v(bytes < 0 ? 0 : size_t(bytes));
    if (std::fread(v.data(), 1, v.size(), f) != v.size())
        die("reading file content failed: " + name);
    std::fclose(f);
    return v;
}

void write_file(const std::string& name, const void* data, size_t bytes)
{
    FILE* f = std::fopen(name.c_str(), "wb");
    if (!f)
        die("opening file '" + name + "' for writing failed");
    if (std::fwrite(data, 1, bytes, f) != bytes)
        die("writing '" + name + "' failed");
    std::fclose(f);
}

// the directory / single-file logic of table_efficiency.cpp:129-160
std::vector<std::string> list_inputs(const std::string& input, bool text)
{
    std::vector<std::string> files;
    if (is_regular_file(input)) {
        files.push_back(input);
        return files;
    }
    DIR* d = opendir(input.c_str());
    if (!d)
        die("cannot open input directory '" + input + "'");
    const std::string suffix = text ? ".txt" : ".u32";
    while (dirent* e = readdir(d)) {
        const std::string path = input + "/" + e->d_name;
        if (ends_with(e->d_name, suffix) && is_regular_file(path))
            files.push_back(path);
    }
    closedir(d);
    std::sort(files.begin(), files.end());
    return files;
}

double compute_ips(size_t n, size_t ns) { return double(n) / (double(ns) / 1e9) / 1e6; } // table_efficiency.cpp:34-39

template <class t_compressor> bool run_efficiency(const std::vector<std::vector<uint32_t>>& inputs)
{
    std::cout << "\\method{" << t_compressor::name() << "}  &" << std::endl;
    std::vector<double> enc_speed, dec_speed;
    bool ok = true;
    for (const auto& input : inputs) {
        std::vector<uint8_t> encoded_data(input.size() * 8 + 4096);
        size_t encoded_bytes = 0;
        size_t enc_ns = std::numeric_limits<size_t>::max();
        for (int i = 0; i < NUM_RUNS; i++) {
            auto t0 = std::chrono::high_resolution_clock::now();
            encoded_bytes = t_compressor::encode(input.data(), input.size(), encoded_data.data(), encoded_data.size());
            auto t1 = std::chrono::high_resolution_clock::now();
            enc_ns = std::min(size_t((t1 - t0).count()), enc_ns);
        }
        encoded_data.resize(encoded_bytes);
        std::vector<uint32_t> recover(input.size());
        size_t dec_ns = std::numeric_limits<size_t>::max();
        for (int i = 0; i < NUM_RUNS; i++) {
            auto t0 = std::chrono::high_resolution_clock::now();
            t_compressor::decode(encoded_data.data(), encoded_data.size(), recover.data(), recover.size());
            auto t1 = std::chrono::high_resolution_clock::now();
            dec_ns = std::min(size_t((t1 - t0).count()), dec_ns);
        }
        if (std::memcmp(input.data(), recover.data(), input.size() * 4) != 0) { // REQUIRE_EQUAL, cutil.hpp:30-50
            std::fprintf(stderr, "ERROR: %s: decoded data differs from the input\n", t_compressor::name().c_str());
            ok = false;
        }
        enc_speed.push_back(compute_ips(input.size(), enc_ns));
        dec_speed.push_back(compute_ips(input.size(), dec_ns));
    }
    for (size_t i = 0; i < enc_speed.size(); i++) {
        for (size_t j = 0; j < i * 4; j++)
            std::printf(" ");
        std::printf("%15.4f  &  %15.4f  ", enc_speed[i], dec_speed[i]);
        std::printf(i + 1 == enc_speed.size() ? "\\\\ \n\n" : "&\n");
    }
    return ok;
}

template <class t_compressor> bool run_effectiveness(const std::vector<std::vector<uint32_t>>& inputs)
{
    std::cout << t_compressor::name() << "  &" << std::endl;
    std::vector<double> BPIs;
    for (const auto& input : inputs) {
        std::vector<uint8_t> encoded_data(input.size() * 8 + 4096);
        const size_t encoded_bytes = t_compressor::encode(input.data(), input.size(), encoded_data.data(), encoded_data.size());
        BPIs.push_back(double(encoded_bytes * 8) / double(input.size()));
    }
    for (size_t i = 0; i < BPIs.size(); i++) {
        for (size_t j = 0; j < i * 4; j++)
            std::printf(" ");
        std::printf("%2.4f  ", BPIs[i]);
        std::printf(i + 1 == BPIs.size() ? "\\\\ \n" : "&\n");
    }
    return true;
}

struct Method {
    int id;
    uint32_t fidelity;
};

bool parse_method(const std::string& s, Method& m)
{
    char name[64];
    for (int id = 0; id <= B200ANS_METHOD_RFOLD; id++)
        for (uint32_t f = 0; f <= 8; f++) {
            const bool folded = id == B200ANS_METHOD_FOLD || id == B200ANS_METHOD_RFOLD;
            if (folded != (f > 0))
                continue;
            b200ans_method_name(id, f, name, sizeof name);
            if (s == name) {
                m = { id, f };
                return true;
            }
        }
    return false;
}

void usage()
{
    std::fprintf(stderr,
        "usage: b200ans_tool efficiency    -i <dir|file> [-t]\n"
        "       b200ans_tool effectiveness -i <dir|file> [-t]\n"
        "       b200ans_tool compress   -m <ANS|ANSmsb|ANSfold-f|ANSrfold-f> -i <file.u32|file.txt> [-t] -o <file.b2an>\n"
        "       b200ans_tool decompress -i <file.b2an> -o <file.u32>\n"
        "  -t: inputs are text files, one unsigned decimal per line (default: raw uint32)\n"
        "  environment: B200ANS_DEVICE (GPU index), B200ANS_BLOCK_SYMS (symbols per block; 0 = reference wire format)\n");
    std::exit(EXIT_FAILURE);
}

} // namespace

int main(int argc, char** argv)
{
    if (argc < 2)
        usage();
    const std::string cmd = argv[1];
    std::string input, output, method_name;
    bool text = false;
    for (int i = 2; i < argc; i++) {
        const std::string a = argv[i];
        if ((a == "-i" || a == "--input") && i + 1 < argc)
            input = argv[++i];
        else if ((a == "-o" || a == "--output") && i + 1 < argc)
            output = argv[++i];
        else if ((a == "-m" || a == "--method") && i + 1 < argc)
            method_name = argv[++i];
        else if (a == "-t" || a == "--text")
            text = true;
        else
            usage();
    }
    if (input.empty())
        usage();

    if (cmd == "efficiency" || cmd == "effectiveness") {
        std::vector<std::vector<uint32_t>> inputs;
        for (const auto& file : list_inputs(input, text)) {
            std::cerr << "loading " << file << std::endl;
            inputs.push_back(text || ends_with(file, ".txt") ? read_file_text(file) : read_file_u32(file));
        }
        if (inputs.empty())
            die("no input files");
        bool ok = true;
        if (cmd == "efficiency") { // the ANS rows of table_efficiency.cpp:175-179 (+ msb, fold-2, rfold-2)
            ok = run_efficiency<ANSint_b200>(inputs) && ok;
            ok = run_efficiency<ANSmsb_b200>(inputs) && ok;
            ok = run_efficiency<ANSfold_b200<1>>(inputs) && ok;
            ok = run_efficiency<ANSfold_b200<2>>(inputs) && ok;
            ok = run_efficiency<ANSfold_b200<5>>(inputs) && ok;
        } else { // the ANS rows of table_effectiveness.cpp:145-146 and fold_effectiveness.cpp
            ok = run_effectiveness<ANSint_b200>(inputs) && ok;
            ok = run_effectiveness<ANSmsb_b200>(inputs) && ok;
            ok = run_effectiveness<ANSfold_b200<1>>(inputs) && ok;
            ok = run_effectiveness<ANSfold_b200<2>>(inputs) && ok;
            ok = run_effectiveness<ANSfold_b200<5>>(inputs) && ok;
        }
        return ok ? EXIT_SUCCESS : EXIT_FAILURE;
    }
    if (cmd == "compress") {
        Method m;
        if (output.empty() || !parse_method(method_name, m))
            usage();
        const std::vector<uint32_t> in = text ? read_file_text(input) : read_file_u32(input);
        std::vector<uint8_t> out(b200ans_compress_bound(b200::context(), m.id, in.size()));
        const size_t bytes = b200::compress(m.id, m.fidelity, out.data(), out.size(), in.data(), in.size());
        write_file(output, out.data(), bytes);
        std::printf("%s;%s;%zu;%zu;%.4f\n", input.c_str(), method_name.c_str(), in.size(), bytes,
            in.empty() ? 0.0 : 8.0 * double(bytes) / double(in.size()));
        return EXIT_SUCCESS;
    }
    if (cmd == "decompress") {
        if (output.empty())
            usage();
        const std::vector<uint8_t> blob = read_file_u8(input);
        b200ans_container info;
        if (b200ans_container_info(blob.data(), blob.size(), &info) != B200ANS_OK)
            die("'" + input + "' is not a b200ans container (reference wire-format streams carry no length: use the API)");
        std::vector<uint32_t> out(info.n_symbols);
        b200::decompress(info.method, info.fidelity, out.data(), out.size(), blob.data(), blob.size());
        write_file(output, out.data(), out.size() * 4);
        char name[64];
        std::printf("%s;%s;%zu;%zu\n", input.c_str(), b200ans_method_name(info.method, info.fidelity, name, sizeof name),
            out.size(), blob.size());
        return EXIT_SUCCESS;
    }
    usage();
}
