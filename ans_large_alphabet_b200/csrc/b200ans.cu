// b200ans C-ABI implementation (include/b200ans.h): host orchestration around the sm_100a
// kernels of kernels.cuh.  No CPU coding path exists here: the histogram, the encoder and the
// decoder run on the GPU or the call fails.
#include "../../include/b200ans.h"

#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <new>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../host/fold.hpp"
#include "../host/model.hpp"
#include "../host/prelude.hpp"
#include "kernels.cuh"
#include "lpb_kernels.cuh"
#include "lpb3_kernels.cuh"
#include "remap_kernels.cuh"
#include "rfold_kernels.cuh"
#include "pa_kernels.cuh"

using namespace b200ans;

namespace {

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap)
            return cudaSuccess;
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            cudaGetLastError();
            want = bytes;
            e = cudaMalloc(&p, want);
        }
        if (e == cudaSuccess)
            cap = want;
        return e;
    }
    void release()
    {
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T* as() const { return static_cast<T*>(p); }
};

struct Header {
    uint32_t magic;
    uint16_t version;
    uint8_t method;
    uint8_t fidelity;
    uint64_t n_symbols;
    uint32_t block_syms;
    uint32_t n_blocks;
    uint32_t prelude_bytes;
    uint32_t flags;
    uint64_t payload_bytes; // padded sum
    uint8_t reserved[24];
};
static_assert(sizeof(Header) == B200ANS_HEADER_BYTES, "container header is 64 bytes");
constexpr uint32_t kFlagPseudoAdaptive = 1u; // Header::flags bit 0: one model per block (pseudo-adaptive container)
constexpr uint32_t kFlagRemap = 2u;          // Header::flags bit 1: per-block dense-rank alphabets in front of a pseudo-adaptive container of the ranks

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- NCCL, opened at run time (no link-time dependency: the library loads on machines without NCCL, and a process
// that already holds libnccl.so.2 -- torch's copy, say -- shares it).  Only the handful of entry points the histogram
// exchange needs; enum values are NCCL 2.x's (nccl.h: ncclSum = 0, ncclMax = 2, ncclUint32 = 3, ncclUint64 = 5).
struct NcclUniqueId {
    char internal[B200ANS_COMM_ID_BYTES];
};
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(void**, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};
constexpr int kNcclSum = 0, kNcclMax = 2, kNcclUint32 = 3, kNcclUint64 = 5;

const NcclApi& nccl_api()
{
    static const NcclApi api = []() {
        NcclApi a;
        const char* override_path = std::getenv("B200ANS_NCCL_LIB");
        for (const char* name : { override_path, "libnccl.so.2", "libnccl.so" }) {
            if (!name || !*name)
                continue;
            a.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (a.lib)
                break;
        }
        if (!a.lib)
            return a;
        auto sym = [&](const char* n) { return dlsym(a.lib, n); };
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
        a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
        a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
        a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
        a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllReduce && a.GroupStart && a.GroupEnd && a.GetErrorString;
        return a;
    }();
    return api;
}

} // namespace

// No C++ exception crosses the C ABI (include/b200ans.h): every extern "C" function is a function-try-block.
#define B200ANS_CATCH_INT                                                                                              \
    catch (const std::bad_alloc&) { return B200ANS_E_NOMEM; }                                                          \
    catch (const std::length_error&) { return B200ANS_E_NOMEM; }                                                       \
    catch (...) { return B200ANS_E_INTERNAL; }

struct b200ans_ctx {
    int device = 0;
    int num_sms = 0;
    size_t max_smem = 0;
    size_t max_smem_sm = 0; // shared memory per SM
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::vector<uint8_t> front; // container front of the encode in flight (must outlive its H2D copy)
    uint32_t block_syms = B200ANS_DEFAULT_BLOCK_SYMS;
    int kernel_path = B200ANS_PATH_AUTO;
    bool dec_have_pure = false;     // dec_bits_pure holds the rank records with whole-group answers (DEC_INT_SMEM_PURE)
    bool dec_pure_possible = false; // ... or can be built from the host copies below (frame <= 2^19)
    std::vector<RankRec> h_rank_rec;
    std::vector<RankEntry> h_rank_ent;
    size_t h_rank_words = 0;
    bool dec_fold_rank = false;     // folded decoder model with a large frame: rank/select index, no slot table
    bool dec_words_aligned = false; // the decode in flight reads container blocks: payloads start 16-byte aligned
    uint32_t h_approx = 1; // adjust_freqs' H_approx (ANSsint<H> / ANSsmsb<H>)
    uint32_t max_nfreq = 0; // largest normalised frequency of the encoder model
    std::string err;
    b200ans_stats stats {};
    cudaEvent_t ev[8] {};
    // host-buffer decode pipeline: a copy-in and a copy-out stream next to `stream`, one event pair per chunk
    static constexpr int kPipeChunks = 8;
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    cudaEvent_t ev_in[kPipeChunks + 1] {}, ev_dec[kPipeChunks] {};
    bool pa3_off = false; // pa_compress: the retry without the lane-per-block encoder is running
    // pageable host buffers (what std::vector hands the C++ adaptors): large copies go through page-locked staging
    // buffers filled / drained by a few host threads (staged_h2d / staged_d2h)
    static constexpr int kStageThreads = 8;   // at most (B200ANS_STAGE_THREADS picks fewer)
    static constexpr size_t kStageChunk = size_t(8) << 20;
    uint8_t* stage_buf[kStageThreads][2] {};
    cudaStream_t stage_st[kStageThreads] {};
    cudaEvent_t stage_ev[kStageThreads][2] {}, stage_done[kStageThreads] {}, stage_fork = nullptr;
    bool stage_ready = false;
    // histogram exchange over NVLink peer memory (b200ans_comm_ipc_export / _attach): my exchange buffer, everybody's
    // (ipc_peer[ipc_rank] is mine), how many exchanges have run
    static constexpr int kIpcMaxWorld = 16;
    void* ipc_buf = nullptr;
    void* ipc_peer[kIpcMaxWorld] {};
    int ipc_rank = 0, ipc_world = 0;
    uint32_t ipc_epoch = 0;
    cudaEvent_t ev_side[2] {}; // fork / join of work that runs beside the main stream inside one call (remap: alphabets beside ranks)

    // grow-only device buffers
    DevBuf bins, small, scratch, block_bytes, offsets, enc_table, mapping, dec_slots, dec_syms, dec_bits, dec_bits_pure, dec_cnt, dec_ent, io_in, io_out, io_bytes, pa_nfreqs, pa_pre, pa_keys, pa_log2tab, rm_ranks, rm_vals, rm_sigma, rm_alpha, rm_bytes, rm_offsets, rm_err;
    bool pa_log2_ready = false;
    // ANSrfold on the device (rfold_kernels.cuh): raw histogram, selection state, candidates, hash of the hot values
    DevBuf rf_sbuckets, rf_stags, rf_fast, rf_cand, rf_ctags, rf_cdisp, rf_cslot, rf_ccount, rf_sketch; // the fast route (rfold_kernels.cuh)
    DevBuf rf_raw, rf_raw32, rf_state, rf_hist, rf_tiles, rf_cand_count, rf_cand_value, rf_most, rf_tags, rf_disp, enc_table_rf;
    size_t rf_guess = 0;          // value range the next optimistic raw histogram is sized for
    bool rf_hash_ok = false;      // the L3 encoder's perfect hash of most_frequent[] is on the device
    bool rf_fast_used = false;    // the current ANSrfold mapping came from the fast route (rfold_fast_model)
    uint32_t rf_mul = 0, rf_bshift = 0;
    uint32_t pa_undecided = 0;    // models of the last pa_compress the device normaliser handed to the host
    uint8_t* pa_pinned = nullptr; // pseudo-adaptive mode: counts / frequencies of all models (pinned, grows on demand)
    size_t pa_pinned_cap = 0;
    // `small` layout: [0] u32 max, [1] u32 error flag, [2..] pad; totals (2 x u64) at byte 16
    uint32_t* d_max() { return small.as<uint32_t>(); }
    uint32_t* d_err() { return small.as<uint32_t>() + 1; }
    uint64_t* d_totals() { return reinterpret_cast<uint64_t*>(small.as<uint8_t>() + 16); }
    // pinned host mirror: kPinnedBins bytes of histogram bins, then the 32 bytes of `small`; lets the small
    // device->host reads of one step share a single synchronisation
    static constexpr size_t kPinnedBins = size_t(64) << 10;
    uint8_t* pinned = nullptr;
    struct SmallHost {
        uint32_t max_value, error_flag, max_log2, max_symbol;
        uint64_t totals[2];
        uint64_t totals_b[2]; // a second scan in flight (pseudo-adaptive: preludes and payloads)
    };
    uint32_t* d_max_log2() { return small.as<uint32_t>() + 2; }
    uint32_t* d_max_symbol() { return small.as<uint32_t>() + 3; }
    uint64_t* d_totals_b() { return reinterpret_cast<uint64_t*>(small.as<uint8_t>() + 32); }
    SmallHost* h_small() { return reinterpret_cast<SmallHost*>(pinned + kPinnedBins); }

    // encoder model
    bool have_enc = false;
    int enc_method = -1;
    uint32_t enc_f = 0;
    std::vector<uint32_t> nfreqs;
    uint64_t frame = 0;
    uint32_t log2_frame = 0;
    std::vector<uint8_t> prelude;
    // rfold value reordering
    bool have_mapping = false;
    uint32_t map_f = 0;
    uint32_t n_values = 0;
    uint64_t rfold_sigma = 0;
    std::vector<uint32_t> most_frequent;

    // decoder model
    bool have_dec = false;
    int dec_method = -1;
    uint32_t dec_f = 0;
    uint32_t dec_log2_frame = 0;
    uint32_t dec_n_syms = 0;
    uint32_t dec_n_present = 0;
    uint32_t dec_dense = 0;
    std::vector<uint8_t> dec_prelude; // the prelude the decoder tables on the device were built from (model cache key)

    // multi-GPU: the communicator the histogram exchange runs on (b200ans_comm_init / _attach)
    void* comm = nullptr;
    cudaEvent_t ev_ar[2] {};
    bool ar_timed = false;
    bool own_comm = false;
    int comm_rank = 0, comm_world = 1;

    int fail(int code, const std::string& what)
    {
        err = what;
        return code;
    }
    int nccl_fail(int e, const char* where)
    {
        const NcclApi& a = nccl_api();
        err = std::string(where) + ": " + (a.ok ? a.GetErrorString(e) : "NCCL not loaded");
        return B200ANS_E_NCCL;
    }
    int cuda_fail(cudaError_t e, const char* where)
    {
        err = std::string(where) + ": " + cudaGetErrorString(e);
        cudaGetLastError();
        return B200ANS_E_CUDA;
    }
};

#define CU(ctx, call)                                                                                                  \
    do {                                                                                                               \
        cudaError_t e__ = (call);                                                                                      \
        if (e__ != cudaSuccess)                                                                                        \
            return (ctx)->cuda_fail(e__, #call);                                                                       \
    } while (0)

#define RET(call)                                                                                                      \
    do {                                                                                                               \
        int r__ = (call);                                                                                              \
        if (r__ != B200ANS_OK)                                                                                         \
            return r__;                                                                                                \
    } while (0)

namespace {

bool valid_method(int method, uint32_t f)
{
    if (method == METHOD_INT || method == METHOD_MSB || method == METHOD_BYTE)
        return true;
    if (method == METHOD_FOLD || method == METHOD_RFOLD)
        return f >= 1 && f <= 8;
    return false;
}

size_t alphabet_bins(int method, uint32_t f, uint32_t max_value)
{
    switch (method) {
    case METHOD_INT: return size_t(max_value) + 1;
    case METHOD_BYTE: return kByteSigma;
    case METHOD_MSB: return kMsbMaxSigma;
    default: return size_t(1) << (f + 9);
    }
}

// worst-case payload bytes of one block (+32 so whole-sector stores never leave the slot); slots are 32-byte aligned
size_t block_stride(int method, size_t block_syms)
{
    const size_t per_sym = (coder_of(method) == METHOD_INT) ? 4 : 7;
    return align_up(per_sym * block_syms + 32, 32) + 32;
}

float elapsed(cudaEvent_t a, cudaEvent_t b)
{
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

// smallest double >= 1/f (kernels.cuh ans_push / lpb_kernels.cuh ans_push52 rely on the direction)
double reciprocal_up(uint32_t f)
{
    double r = 1.0 / double(f);
    if (std::fma(r, double(f), -1.0) < 0.0) // r * f < 1 exactly: one ulp up
        r = std::nextafter(r, 2.0);
    return r;
}

void set_name(char (&dst)[32], const char* name)
{
    std::snprintf(dst, sizeof dst, "%s", name);
}

double now_ms()
{
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

} // namespace

// ==========================================================================================
// context
// ==========================================================================================
extern "C" int b200ans_create(b200ans_ctx** out, int device, void* stream)
try {
    if (!out)
        return B200ANS_E_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0 || device < 0 || device >= count) {
        cudaGetLastError();
        return B200ANS_E_CUDA; // no CPU fallback: without a GPU there is no coder
    }
    b200ans_ctx* c = new b200ans_ctx();
    c->device = device;
    if (cudaSetDevice(device) != cudaSuccess) {
        delete c;
        return B200ANS_E_CUDA;
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    c->num_sms = prop.multiProcessorCount;
    c->max_smem = prop.sharedMemPerBlockOptin;
    c->max_smem_sm = prop.sharedMemPerMultiprocessor;
    if (stream) {
        c->stream = static_cast<cudaStream_t>(stream);
    } else {
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
            delete c;
            return B200ANS_E_CUDA;
        }
        c->own_stream = true;
    }
    for (auto& e : c->ev)
        cudaEventCreate(&e);
    for (auto& e : c->ev_ar)
        cudaEventCreate(&e);
    if (cudaHostAlloc(reinterpret_cast<void**>(&c->pinned), b200ans_ctx::kPinnedBins + 64, cudaHostAllocDefault) != cudaSuccess)
        c->pinned = nullptr;
    if (!c->pinned || c->small.reserve(256) != cudaSuccess) {
        b200ans_destroy(c);
        return B200ANS_E_CUDA;
    }
    *out = c;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" void b200ans_destroy(b200ans_ctx* c)
{
    if (!c)
        return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    for (DevBuf* b : { &c->bins, &c->small, &c->scratch, &c->block_bytes, &c->offsets, &c->enc_table, &c->mapping,
             &c->dec_slots, &c->dec_syms, &c->dec_bits, &c->dec_bits_pure, &c->dec_cnt, &c->dec_ent, &c->io_in, &c->io_out, &c->io_bytes, &c->pa_nfreqs, &c->pa_pre, &c->pa_keys, &c->pa_log2tab, &c->rm_ranks, &c->rm_vals, &c->rm_sigma, &c->rm_alpha, &c->rm_bytes, &c->rm_offsets, &c->rm_err,
             &c->rf_sbuckets, &c->rf_stags, &c->rf_fast, &c->rf_cand, &c->rf_ctags, &c->rf_cdisp, &c->rf_cslot, &c->rf_ccount, &c->rf_sketch,
             &c->rf_raw, &c->rf_raw32, &c->rf_state, &c->rf_hist, &c->rf_tiles, &c->rf_cand_count, &c->rf_cand_value, &c->rf_most, &c->rf_tags, &c->rf_disp,
             &c->enc_table_rf })
        b->release();
    for (auto& e : c->ev)
        if (e)
            cudaEventDestroy(e);
    for (auto& e : c->ev_ar)
        if (e)
            cudaEventDestroy(e);
    for (auto& e : c->ev_in)
        if (e)
            cudaEventDestroy(e);
    for (auto& e : c->ev_dec)
        if (e)
            cudaEventDestroy(e);
    for (auto& e : c->ev_side)
        if (e)
            cudaEventDestroy(e);
    if (c->copy_in)
        cudaStreamDestroy(c->copy_in);
    if (c->copy_out)
        cudaStreamDestroy(c->copy_out);
    if (c->pinned)
        cudaFreeHost(c->pinned);
    if (c->pa_pinned)
        cudaFreeHost(c->pa_pinned);
    if (c->comm && c->own_comm && nccl_api().ok)
        nccl_api().CommDestroy(c->comm);
    for (int w = 0; w < b200ans_ctx::kStageThreads; w++) {
        for (int b = 0; b < 2; b++) {
            if (c->stage_buf[w][b])
                cudaFreeHost(c->stage_buf[w][b]);
            if (c->stage_ev[w][b])
                cudaEventDestroy(c->stage_ev[w][b]);
        }
        if (c->stage_done[w])
            cudaEventDestroy(c->stage_done[w]);
        if (c->stage_st[w])
            cudaStreamDestroy(c->stage_st[w]);
    }
    if (c->stage_fork)
        cudaEventDestroy(c->stage_fork);
    if (c->ipc_buf) {
        for (int p = 0; p < c->ipc_world; p++)
            if (p != c->ipc_rank && c->ipc_peer[p])
                cudaIpcCloseMemHandle(c->ipc_peer[p]);
        cudaFree(c->ipc_buf);
        c->ipc_buf = nullptr;
    }
    if (c->own_stream)
        cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" int b200ans_set_block_syms(b200ans_ctx* c, uint32_t block_syms)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (block_syms != 0 && (block_syms % 128 != 0 || block_syms > (1u << 24)))
        return c->fail(B200ANS_E_ARG, "block_syms must be 0 or a multiple of 128 up to 2^24");
    c->block_syms = block_syms;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" uint32_t b200ans_suggest_block_syms(const b200ans_ctx* c, size_t n, uint32_t lo, uint32_t hi)
{
    // Lane-per-block coders: a warp owns 32 blocks for the whole launch and every SM holds kL3Warps warps, so a
    // launch lasts rounds x (time of one block) with rounds = ceil(tasks / (SMs * warps)).  Cost ~ rounds * block size.
    const uint64_t kL3Warps = 16;
    lo = std::max<uint32_t>(128u, (lo + 127u) & ~127u);
    hi = std::min<uint32_t>(hi & ~127u, 1u << 24);
    if (lo > hi)
        return 0;
    const uint64_t slots = uint64_t(c ? c->num_sms : 148) * kL3Warps;
    uint32_t best = hi;
    uint64_t best_cost = ~uint64_t(0);
    for (uint32_t bs = hi;; bs -= 128) {
        const uint64_t blocks = (uint64_t(n) + bs - 1) / bs;
        const uint64_t tasks = (blocks + 31) / 32;
        const uint64_t rounds = std::max<uint64_t>(1, (tasks + slots - 1) / slots);
        const uint64_t cost = rounds * bs;
        if (cost < best_cost) { // strict: the larger block size keeps a tie (less container overhead)
            best_cost = cost;
            best = bs;
        }
        if (bs == lo)
            break;
    }
    return best;
}

extern "C" int b200ans_set_kernel_path(b200ans_ctx* c, int path)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (path != B200ANS_PATH_AUTO && path != B200ANS_PATH_GROUP && path != B200ANS_PATH_LPB)
        return c->fail(B200ANS_E_ARG, "unknown kernel path");
    c->kernel_path = path;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_container_info(const uint8_t* src, size_t src_bytes, b200ans_container* out)
try {
    if (!src || !out)
        return B200ANS_E_ARG;
    if (src_bytes < B200ANS_HEADER_BYTES)
        return B200ANS_E_FORMAT;
    Header h;
    std::memcpy(&h, src, sizeof h);
    if (h.magic != B200ANS_MAGIC || h.version != 1 || h.method > METHOD_BYTE)
        return B200ANS_E_FORMAT;
    out->method = h.method;
    out->fidelity = h.fidelity;
    out->n_symbols = h.n_symbols;
    out->block_syms = h.block_syms;
    out->n_blocks = h.n_blocks;
    out->prelude_bytes = h.prelude_bytes;
    out->payload_bytes = h.payload_bytes;
    out->total_bytes = h.n_symbols == 0
        ? B200ANS_HEADER_BYTES
        : align_up(B200ANS_HEADER_BYTES + align_up(h.prelude_bytes, 4) + size_t(h.n_blocks) * 4, 16) + h.payload_bytes;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_set_h_approx(b200ans_ctx* c, uint32_t h_approx)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (h_approx == 0 || h_approx > 1000)
        return c->fail(B200ANS_E_ARG, "H_approx must be in 1..1000");
    c->h_approx = h_approx;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_get_stats(const b200ans_ctx* c, b200ans_stats* out)
try {
    if (!c || !out)
        return B200ANS_E_ARG;
    *out = c->stats;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" const char* b200ans_last_error(const b200ans_ctx* c) { return c ? c->err.c_str() : "null context"; }

extern "C" const char* b200ans_strerror(int code)
{
    switch (code) {
    case B200ANS_OK: return "ok";
    case B200ANS_E_ARG: return "bad argument";
    case B200ANS_E_CUDA: return "CUDA failure (or no CUDA device: this library has no CPU fallback)";
    case B200ANS_E_CAPACITY: return "destination too small";
    case B200ANS_E_DOMAIN: return "input outside the coder's valid domain";
    case B200ANS_E_FORMAT: return "not a b200ans container";
    case B200ANS_E_STATE: return "call sequence error";
    case B200ANS_E_NCCL: return "NCCL unavailable or a collective failed";
    case B200ANS_E_NOMEM: return "host allocation failed";
    case B200ANS_E_INTERNAL: return "internal error (exception stopped at the ABI boundary)";
    }
    return "unknown error";
}

// methods.hpp:485,500,530-533,550-553
extern "C" const char* b200ans_method_name(int method, uint32_t fidelity, char* buf, size_t buflen)
{
    if (!buf || buflen == 0)
        return "";
    switch (method) {
    case METHOD_INT: snprintf(buf, buflen, "ANS"); break;
    case METHOD_BYTE: snprintf(buf, buflen, "ANSbyte"); break;
    case METHOD_MSB: snprintf(buf, buflen, "ANSmsb"); break;
    case METHOD_FOLD: snprintf(buf, buflen, "ANSfold-%u", fidelity); break;
    case METHOD_RFOLD: snprintf(buf, buflen, "ANSrfold-%u", fidelity); break;
    default: snprintf(buf, buflen, "?"); break;
    }
    return buf;
}

extern "C" size_t b200ans_compress_bound(const b200ans_ctx* c, int method, size_t n)
{
    const uint32_t bs = c ? c->block_syms : B200ANS_DEFAULT_BLOCK_SYMS;
    const size_t per_sym = (coder_of(method) == METHOD_INT) ? 4 : 7;
    const size_t n_blocks = bs ? (n + bs - 1) / bs : 1;
    // header + prelude (worst case ~ 5 bytes per distinct symbol, bounded by n and 2^30) + table + payloads
    const size_t prelude = 4096 + 8 * std::min<size_t>(n, size_t(1) << 26) + (size_t(4) << 12);
    return B200ANS_HEADER_BYTES + prelude + 4 * n_blocks + 16 + per_sym * n + n_blocks * 48 + 64;
}

// ==========================================================================================
// host-only model API
// ==========================================================================================
extern "C" int b200ans_adjust_freqs(const uint64_t* freqs, size_t nbins, uint32_t largest_sym, int require_u16,
    uint32_t H_approx, uint32_t* out)
try {
    if (!freqs || !out || largest_sym >= nbins)
        return B200ANS_E_ARG;
    std::vector<uint64_t> F(freqs, freqs + nbins);
    const std::vector<uint32_t> s = adjust_freqs(F, largest_sym, require_u16 != 0, H_approx);
    std::copy(s.begin(), s.end(), out);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// scale_freqs (ans_util.hpp:77-95)
extern "C" int b200ans_scale_freqs(uint32_t* S, const uint64_t* F, size_t nbins, const uint32_t* mapping, size_t sigma,
    int64_t M, uint64_t freq_sum, int* fits)
try {
    if (!S || !F || !mapping || !fits || sigma > nbins)
        return B200ANS_E_ARG;
    for (size_t r = 0; r < sigma; r++)
        if (mapping[r] >= nbins)
            return B200ANS_E_ARG;
    std::vector<uint32_t> s(S, S + nbins), map(mapping, mapping + sigma);
    const std::vector<uint64_t> f(F, F + nbins);
    const bool failed = scale_freqs(s, f, map, M, sigma, size_t(freq_sum));
    std::copy(s.begin(), s.end(), S);
    *fits = failed ? 0 : 1;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_serialize_interp(const uint32_t* nfreqs, size_t count, uint64_t frame_size, uint8_t* dst,
    size_t capacity, size_t* out_bytes)
try {
    if (!nfreqs || !dst || !out_bytes || count == 0)
        return B200ANS_E_ARG;
    if (capacity < 8 * count + 16)
        return B200ANS_E_CAPACITY;
    std::vector<uint32_t> v(nfreqs, nfreqs + count);
    uint8_t* p = dst;
    ans_serialize_interp(v, size_t(frame_size), p);
    *out_bytes = size_t(p - dst);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_load_interp(const uint8_t* src, uint32_t* dst, size_t capacity, size_t* out_count, size_t* consumed)
try {
    if (!src || !out_count)
        return B200ANS_E_ARG;
    const uint8_t* p = src;
    const uint32_t max_sym = vbyte_decode_u32(p);
    *out_count = size_t(max_sym) + 1;
    if (!dst)
        return B200ANS_OK;
    if (capacity < *out_count)
        return B200ANS_E_CAPACITY;
    const std::vector<uint32_t> v = ans_load_interp(src, consumed);
    std::copy(v.begin(), v.end(), dst);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_interp_encode(uint32_t* dst, const uint32_t* in, size_t n, size_t universe, size_t* out_bytes)
try {
    if (!dst || (!in && n) || !out_bytes)
        return B200ANS_E_ARG;
    *out_bytes = interp_encode(dst, in, n, universe);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_interp_decode(const uint32_t* src, uint32_t* out, size_t n, size_t universe, size_t* words_read)
try {
    if (!src || (!out && n))
        return B200ANS_E_ARG;
    const size_t w = interp_decode(src, out, n, universe);
    if (words_read)
        *words_read = w;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" uint32_t b200ans_fold_map(uint32_t f, uint32_t x, uint32_t* nbytes)
{
    uint32_t k;
    const uint32_t s = fold_map(x, f, k);
    if (nbytes)
        *nbytes = k;
    return s;
}
extern "C" uint32_t b200ans_fold_unmap(uint32_t f, uint32_t sym, uint32_t* nbytes)
{
    uint32_t k;
    const uint32_t v = fold_unmap(sym, f, k);
    if (nbytes)
        *nbytes = k;
    return v;
}
extern "C" uint32_t b200ans_msb_map(uint32_t x, uint32_t* nbytes)
{
    uint32_t k;
    const uint32_t s = msb_map(x, k);
    if (nbytes)
        *nbytes = k;
    return s;
}
extern "C" uint32_t b200ans_msb_unmap(uint32_t sym, uint32_t* nbytes)
{
    uint32_t k;
    const uint32_t v = msb_unmap(sym, k);
    if (nbytes)
        *nbytes = k;
    return v;
}

// ==========================================================================================
// K1 launches: maximum + histogram
// ==========================================================================================
namespace {

template <int METHOD>
int launch_histogram(b200ans_ctx* c, uint32_t f, const uint32_t* d_src, size_t n, uint64_t* d_bins, size_t nbins,
    uint32_t* d_max, unsigned int* d_bins32 = nullptr)
{
    HistArgs a {};
    a.bins32 = d_bins32;
    a.in = d_src;
    a.n = n;
    a.fidelity = f;
    a.mapping = c->mapping.as<uint32_t>();
    a.n_values = c->n_values;
    a.bins = reinterpret_cast<unsigned long long*>(d_bins);
    a.nbins = uint32_t(nbins);
    a.max_value = d_max;
    // whole alphabet in shared memory when it fits, else only the hot head of the raw alphabet
    a.smem_bins = nbins <= kHistSmemWords ? uint32_t(nbins) : kHistHeadBins;
    uint32_t replicas = 8;
    while (replicas > 1 && replicas * (a.smem_bins + 1) > kHistSmemWords)
        replicas >>= 1;
    a.replicas = replicas;
    a.stride = (a.smem_bins & 31u) == 0 ? a.smem_bins + 1 : a.smem_bins | 1u; // one bank of skew per replica
    if ((a.stride & 31u) == 0)
        a.stride += 1;
    const size_t smem_bytes = size_t(a.replicas) * a.stride * sizeof(uint32_t);
    const bool all_smem = a.smem_bins == a.nbins;
    auto k = all_smem ? histogram_kernel<METHOD, true> : (d_bins32 && METHOD == METHOD_INT) ? histogram_kernel<METHOD, false, METHOD == METHOD_INT> : histogram_kernel<METHOD, false>;
    if (all_smem && d_bins32)
        return c->fail(B200ANS_E_INTERNAL, "histogram: 32-bit bins are for alphabets beyond the shared-memory head");
    CU(c, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kHistSmemWords * 4 + 1024)));
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, kHistThreads, smem_bytes) != cudaSuccess || per_sm < 1) {
        cudaGetLastError();
        per_sm = 1;
    }
    const size_t vec = (n >> 2) + 1;
    int grid = int(std::min<size_t>((vec + kHistThreads - 1) / kHistThreads, size_t(c->num_sms) * per_sm));
    grid = std::max(grid, 1);
    k<<<grid, kHistThreads, smem_bytes, c->stream>>>(a);
    c->stats.launches++;
    CU(c, cudaGetLastError());
    return B200ANS_OK;
}

int histogram_dispatch(b200ans_ctx* c, int method, uint32_t f, const uint32_t* d_src, size_t n, uint64_t* d_bins,
    size_t nbins, uint32_t* d_max)
{
    if (nbins == 0 || nbins > (size_t(1) << 31))
        return c->fail(B200ANS_E_ARG, "histogram: bad bin count");
    switch (method) {
    case METHOD_INT:
    case METHOD_BYTE: return launch_histogram<METHOD_INT>(c, f, d_src, n, d_bins, nbins, d_max);
    case METHOD_FOLD: return launch_histogram<METHOD_FOLD>(c, f, d_src, n, d_bins, nbins, d_max);
    case METHOD_MSB: return launch_histogram<METHOD_MSB>(c, f, d_src, n, d_bins, nbins, d_max);
    case METHOD_RFOLD:
        if (!c->have_mapping || c->map_f != f)
            return c->fail(B200ANS_E_STATE, "histogram(RFOLD) needs b200ans_rfold_prepare first");
        return launch_histogram<METHOD_RFOLD>(c, f, d_src, n, d_bins, nbins, d_max);
    }
    return c->fail(B200ANS_E_ARG, "unknown method");
}

// maximum of the raw values: the histogram kernel in INT mode with a single dummy bin range of 0
__global__ void __launch_bounds__(512) max_kernel(const uint32_t* __restrict__ in, uint64_t n, uint32_t* __restrict__ out)
{
    uint32_t m = 0;
    const uint64_t n4 = n >> 2;
    const uint4* in4 = reinterpret_cast<const uint4*>(in);
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const uint4 q = __ldg(in4 + i);
        m = max(max(m, q.x), max(q.y, max(q.z, q.w)));
    }
    if (blockIdx.x == 0 && threadIdx.x < 3) {
        const uint64_t t = (n4 << 2) + threadIdx.x;
        if (t < n)
            m = max(m, in[t]);
    }
    for (int o = 16; o > 0; o >>= 1)
        m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m)
        atomicMax(out, m);
}

} // namespace

extern "C" int b200ans_max_device(b200ans_ctx* c, const uint32_t* d_src, size_t n, uint32_t* d_max)
try {
    if (!c || !d_src || !d_max || (reinterpret_cast<uintptr_t>(d_src) & 15))
        return c ? c->fail(B200ANS_E_ARG, "max: null or unaligned pointer") : B200ANS_E_ARG;
    CU(c, cudaSetDevice(c->device));
    const size_t vec = (n >> 2) + 1;
    const int grid = int(std::max<size_t>(1, std::min<size_t>((vec + 511) / 512, size_t(c->num_sms) * 4)));
    max_kernel<<<grid, 512, 0, c->stream>>>(d_src, n, d_max);
    c->stats.launches++;
    CU(c, cudaGetLastError());
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_histogram_device(b200ans_ctx* c, int method, uint32_t f, const uint32_t* d_src, size_t n,
    uint64_t* d_bins, size_t nbins)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_src || !d_bins || !valid_method(method, f) || (reinterpret_cast<uintptr_t>(d_src) & 15))
        return c->fail(B200ANS_E_ARG, "histogram: bad argument (d_src must be 16-byte aligned)");
    CU(c, cudaSetDevice(c->device));
    return histogram_dispatch(c, method, f, d_src, n, d_bins, nbins, c->d_max());
}
B200ANS_CATCH_INT

// As b200ans_histogram_device, and the largest raw value of the input goes to *d_max (device, zeroed here).  Values
// whose bin is beyond nbins are not counted, so a caller that guesses the alphabet (plain ANS: ans_int.hpp:42-49
// takes the maximum first) counts optimistically in ONE pass and looks at the maximum afterwards.
extern "C" int b200ans_histogram_max_device(b200ans_ctx* c, int method, uint32_t f, const uint32_t* d_src, size_t n,
    uint64_t* d_bins, size_t nbins, uint32_t* d_max)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_src || !d_bins || !d_max || !valid_method(method, f) || (reinterpret_cast<uintptr_t>(d_src) & 15))
        return c->fail(B200ANS_E_ARG, "histogram: bad argument (d_src must be 16-byte aligned)");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaMemsetAsync(d_max, 0, 4, c->stream));
    return histogram_dispatch(c, method, f, d_src, n, d_bins, nbins, d_max);
}
B200ANS_CATCH_INT

// ==========================================================================================
// models
// ==========================================================================================


namespace {

// Perfect hash of the K = 2^(f+7) most frequent values for the L3 encoder (lpb3_kernels.cuh L3Rfold): "hash and
// displace" -- bucket b = (v * mul) >> bshift carries a displacement d_b, slot(v) = (v * mul + d_b) & (K - 1), every
// hot value gets a slot of its own (K slots for K values: minimal).  Buckets are placed largest first.
bool rfold_build_hash(const std::vector<uint32_t>& hot, uint32_t slots, uint32_t mul, uint32_t bshift, std::vector<uint32_t>& tags,
    std::vector<uint16_t>& disp, std::vector<uint32_t>& slot_of_rank)
{
    const uint32_t K = slots; // a power of two, at least hot.size()
    const uint32_t B = 1u << (32u - bshift);
    std::vector<std::vector<uint32_t>> buckets(B);
    for (uint32_t r = 0; r < hot.size(); r++)
        buckets[(hot[r] * mul) >> bshift].push_back(r);
    std::vector<uint32_t> order(B);
    for (uint32_t b = 0; b < B; b++)
        order[b] = b;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) { return buckets[x].size() > buckets[y].size(); });
    tags.assign(K, 0xFFFFFFFFu);
    disp.assign(B, 0);
    slot_of_rank.assign(hot.size(), 0);
    std::vector<uint8_t> used(K, 0);
    for (uint32_t b : order) {
        const auto& keys = buckets[b];
        if (keys.empty())
            break;
        bool placed = false;
        for (uint32_t d = 0; d < K && !placed; d++) {
            bool ok = true;
            size_t i = 0;
            for (; i < keys.size(); i++) {
                const uint32_t s = (hot[keys[i]] * mul + d) & (K - 1);
                if (used[s]) {
                    ok = false;
                    break;
                }
                used[s] = 2; // tentatively
            }
            if (ok) {
                for (uint32_t r : keys) {
                    const uint32_t s = (hot[r] * mul + d) & (K - 1);
                    used[s] = 1;
                    tags[s] = hot[r];
                    slot_of_rank[r] = s;
                }
                disp[b] = uint16_t(d);
                placed = true;
            } else {
                for (size_t j = 0; j < i; j++) {
                    const uint32_t s = (hot[keys[j]] * mul + d) & (K - 1);
                    if (used[s] == 2)
                        used[s] = 0;
                }
            }
        }
        if (!placed)
            return false;
    }
    return true;
}

} // namespace

// ans_reorder_fold.hpp:75-106: order values by (-count, value), the first 2^(f+7) become
// 0..2^(f+7)-1 and everything else moves up by 2^(f+7); identity if fewer distinct values.
extern "C" int b200ans_rfold_prepare(b200ans_ctx* c, uint32_t f, const uint64_t* raw_bins, size_t nvalues)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!raw_bins || nvalues == 0 || f < 1 || f > 8 || nvalues > (size_t(1) << 31))
        return c->fail(B200ANS_E_ARG, "rfold_prepare: bad argument");
    CU(c, cudaSetDevice(c->device));
    const uint32_t thres = 1u << (f + 7);
    std::vector<std::pair<uint64_t, uint32_t>> seen; // (count, value)
    for (size_t v = 0; v < nvalues; v++)
        if (raw_bins[v])
            seen.emplace_back(raw_bins[v], uint32_t(v));
    c->rfold_sigma = seen.size();
    // Fewer than 2^(f+7) distinct values: the reference's encoder keeps the identity mapping (ans_reorder_fold.hpp:94-97)
    // but its decoder subtracts 2^(f+7) from every value regardless (:301), so such an input only round-trips when all
    // values are below the threshold (then nothing is folded and the subtraction cancels the table's own offset).
    if (seen.size() < thres && !seen.empty() && seen.back().second >= thres)
        return c->fail(B200ANS_E_DOMAIN,
            "ANSrfold: fewer than 2^(f+7) distinct values and a value >= 2^(f+7): the reference's decoder returns wrong values for this input");
    std::vector<uint32_t> mapping(nvalues);
    c->most_frequent.clear();
    if (seen.size() < thres) {
        for (size_t v = 0; v < nvalues; v++)
            mapping[v] = uint32_t(v);
    } else {
        auto hotter = [](const std::pair<uint64_t, uint32_t>& a, const std::pair<uint64_t, uint32_t>& b) {
            return a.first != b.first ? a.first > b.first : a.second < b.second;
        };
        std::partial_sort(seen.begin(), seen.begin() + thres, seen.end(), hotter);
        for (size_t v = 0; v < nvalues; v++)
            mapping[v] = uint32_t(v) + thres;
        c->most_frequent.resize(thres);
        for (uint32_t i = 0; i < thres; i++) {
            mapping[seen[i].second] = i;
            c->most_frequent[i] = seen[i].second;
        }
    }
    CU(c, c->mapping.reserve(nvalues * sizeof(uint32_t)));
    CU(c, cudaMemcpyAsync(c->mapping.p, mapping.data(), nvalues * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->n_values = uint32_t(nvalues);
    c->map_f = f;
    c->have_mapping = true;
    c->rf_fast_used = false;
    return B200ANS_OK;
}
B200ANS_CATCH_INT


namespace {

// rfold_kernels.cuh steps 1-5 on the context's stream: from the raw device histogram to most_frequent[] (rf_most),
// mapping[] and the folded histogram; nothing is read back here.  `count_bits`: no count needs more bits.
int rfold_select_launch(b200ans_ctx* c, uint32_t f, const uint64_t* d_raw, uint32_t nvalues, uint32_t count_bits, uint64_t* d_folded)
{
    const uint32_t K = 1u << (f + 7);
    const uint32_t n_tiles = (nvalues + kRfTile - 1) / kRfTile;
    const size_t nfolded = alphabet_bins(METHOD_RFOLD, f, 0);
    CU(c, c->rf_state.reserve(sizeof(RfState)));
    CU(c, c->rf_hist.reserve(kRfDigits * 4));
    CU(c, c->rf_tiles.reserve(size_t(n_tiles) * 4));
    CU(c, c->rf_cand_count.reserve(size_t(K) * 8));
    CU(c, c->rf_cand_value.reserve(size_t(K) * 4));
    CU(c, c->rf_most.reserve(size_t(K) * 4));
    CU(c, c->mapping.reserve(size_t(nvalues) * 4));
    CU(c, cudaMemsetAsync(c->rf_state.p, 0, sizeof(RfState), c->stream));
    CU(c, cudaMemsetAsync(c->rf_hist.p, 0, kRfDigits * 4, c->stream));
    CU(c, cudaMemsetAsync(d_folded, 0, nfolded * 8, c->stream));
    RfState* st = c->rf_state.as<RfState>();
    const auto* raw = reinterpret_cast<const unsigned long long*>(d_raw);
    const int grid = int(std::max<uint32_t>(1, std::min<uint32_t>((nvalues + kRfThreads - 1) / kRfThreads, uint32_t(c->num_sms) * 8)));
    const int tile_grid = int(std::max<uint32_t>(1, std::min<uint32_t>(n_tiles, uint32_t(c->num_sms) * 8)));
    const uint32_t passes = std::max<uint32_t>(1, (std::min<uint32_t>(count_bits, 64) + kRfDigitBits - 1) / kRfDigitBits);
    for (uint32_t p = 0; p < passes; p++) {
        const uint32_t shift = (passes - 1 - p) * kRfDigitBits;
        rf_digit_hist_kernel<<<grid, kRfThreads, 0, c->stream>>>(raw, nvalues, shift, p == 0, st, c->rf_hist.as<unsigned int>(), st);
        rf_digit_pick_kernel<<<1, kRfDigits / 4, 0, c->stream>>>(c->rf_hist.as<unsigned int>(), shift, p == 0, K, st);
        c->stats.launches += 2;
    }
    rf_tie_count_kernel<<<tile_grid, kRfThreads, 0, c->stream>>>(raw, nvalues, st, c->rf_tiles.as<unsigned int>());
    rf_tie_scan_kernel<<<1, 1024, 0, c->stream>>>(c->rf_tiles.as<unsigned int>(), n_tiles);
    rf_collect_kernel<<<tile_grid, kRfThreads, 0, c->stream>>>(raw, nvalues, st, c->rf_tiles.as<unsigned int>(), K,
        c->rf_cand_count.as<unsigned long long>(), c->rf_cand_value.as<uint32_t>());
    rf_rank_kernel<<<(K + kRfThreads - 1) / kRfThreads, kRfThreads, 0, c->stream>>>(st, K, c->rf_cand_count.as<unsigned long long>(),
        c->rf_cand_value.as<uint32_t>(), c->rf_most.as<uint32_t>());
    rf_mapping_fill_kernel<<<grid, kRfThreads, 0, c->stream>>>(st, K, nvalues, c->mapping.as<uint32_t>());
    rf_mapping_hot_kernel<<<(K + kRfThreads - 1) / kRfThreads, kRfThreads, 0, c->stream>>>(st, K, c->rf_most.as<uint32_t>(), c->mapping.as<uint32_t>());
    if (nfolded * 8 <= 48 * 1024)
        rf_fold_bins_kernel<true><<<grid, kRfThreads, nfolded * 8, c->stream>>>(raw, nvalues, c->mapping.as<uint32_t>(), f, uint32_t(nfolded),
            reinterpret_cast<unsigned long long*>(d_folded));
    else
        rf_fold_bins_kernel<false><<<grid, kRfThreads, 0, c->stream>>>(raw, nvalues, c->mapping.as<uint32_t>(), f, uint32_t(nfolded),
            reinterpret_cast<unsigned long long*>(d_folded));
    c->stats.launches += 7;
    CU(c, cudaGetLastError());
    return B200ANS_OK;
}

} // namespace

// Device counterpart of b200ans_rfold_prepare (K4 of SURVEY.md section 8a; ans_reorder_fold.hpp:75-115): from the raw
// histogram of the value range ON THE DEVICE to the value reordering (mapping[] stays on the device, most_frequent[]
// comes back for the prelude) and the folded histogram over the reordered values, which is written to d_folded_bins
// (2^(f+9) x u64, device) -- no second pass over the input.  One stream synchronisation, at the end.
extern "C" int b200ans_rfold_prepare_device(b200ans_ctx* c, uint32_t f, const uint64_t* d_raw_bins, size_t nvalues, uint64_t* d_folded_bins)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_raw_bins || !d_folded_bins || nvalues == 0 || f < 1 || f > 8 || nvalues > (size_t(1) << 30))
        return c->fail(B200ANS_E_ARG, "rfold_prepare_device: bad argument");
    CU(c, cudaSetDevice(c->device));
    c->have_mapping = false;
    const uint32_t K = 1u << (f + 7);
    RET(rfold_select_launch(c, f, d_raw_bins, uint32_t(nvalues), 64, d_folded_bins));
    RfState st;
    std::vector<uint32_t> most(K);
    CU(c, cudaMemcpyAsync(&st, c->rf_state.p, sizeof st, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaMemcpyAsync(most.data(), c->rf_most.p, size_t(K) * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->rfold_sigma = st.sigma;
    c->most_frequent.clear();
    if (st.reorder) {
        c->most_frequent = std::move(most);
    } else {
        // identity mapping (ans_reorder_fold.hpp:94-97); the reference's decoder still subtracts 2^(f+7) (:301), so the
        // input only round-trips when every value is below the threshold (b200ans_rfold_prepare has the details)
        if (st.sigma != 0 && st.max_seen >= K)
            return c->fail(B200ANS_E_DOMAIN,
                "ANSrfold: fewer than 2^(f+7) distinct values and a value >= 2^(f+7): the reference's decoder returns wrong values for this input");
    }
    c->n_values = uint32_t(nvalues);
    c->map_f = f;
    c->have_mapping = true;
    c->rf_fast_used = false;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// ans_X_encode::create after the counting loop (ans_fold.hpp:79-92) + serialize (ans_fold.hpp:96-99,
// ans_reorder_fold.hpp:132-154) + the device encoder table.
extern "C" int b200ans_model_build(b200ans_ctx* c, int method, uint32_t f, const uint64_t* bins, size_t nbins)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!bins || !valid_method(method, f) || nbins == 0)
        return c->fail(B200ANS_E_ARG, "model_build: bad argument");
    CU(c, cudaSetDevice(c->device));
    const double t0 = now_ms();
    c->have_enc = false;
    if (method == METHOD_RFOLD && (!c->have_mapping || c->map_f != f))
        return c->fail(B200ANS_E_STATE, "model_build(RFOLD) needs b200ans_rfold_prepare first");
    std::vector<uint64_t> freqs(bins, bins + nbins);
    size_t sigma = 0;
    uint32_t max_sym = 0;
    for (size_t s = 0; s < nbins; s++)
        if (freqs[s]) {
            sigma++;
            max_sym = uint32_t(s);
        }
    if (sigma == 0)
        return c->fail(B200ANS_E_DOMAIN, "model_build: empty histogram");
    if (method == METHOD_INT && sigma < 2)
        return c->fail(B200ANS_E_DOMAIN,
            "plain ANS needs at least two distinct symbols (the reference's adjust_freqs never terminates otherwise)");
    if (method == METHOD_BYTE) { // ans_byte.hpp:99-120: all 256 entries, its own normaliser
        if (nbins > kByteSigma)
            return c->fail(B200ANS_E_ARG, "model_build(BYTE): at most 256 bins");
        freqs.resize(kByteSigma, 0);
        const std::array<uint16_t, 256> nf = byte_adjust_freqs(freqs.data());
        c->nfreqs.assign(nf.begin(), nf.end());
    } else {
        c->nfreqs = adjust_freqs(freqs, max_sym, method != METHOD_INT, c->h_approx);
    }
    uint64_t frame = 0;
    for (uint32_t v : c->nfreqs)
        frame += v;
    if (!is_power_of_two(frame) || frame > (uint64_t(1) << 28))
        return c->fail(B200ANS_E_DOMAIN, "model_build: normalisation produced an unusable frame size");
    c->frame = frame;
    c->log2_frame = uint32_t(63 - __builtin_clzll(frame));

    // prelude
    c->prelude.assign(c->nfreqs.size() * 8 + 64 + (method == METHOD_RFOLD ? 4 * c->most_frequent.size() + 8 : 0), 0);
    uint8_t* p = c->prelude.data();
    if (method == METHOD_RFOLD) {
        const uint32_t flag = c->most_frequent.empty() ? 0u : 1u;
        std::memcpy(p, &flag, 4);
        p += 4;
        if (flag) {
            std::memcpy(p, c->most_frequent.data(), 4 * c->most_frequent.size());
            p += 4 * c->most_frequent.size();
        }
    }
    if (method == METHOD_BYTE) { // ans_byte.hpp:122-137: interp code of the running sums of (nfreq + 1), no header
        uint32_t inc[kByteSigma];
        inc[0] = c->nfreqs[0];
        for (uint32_t sym = 1; sym < kByteSigma; sym++)
            inc[sym] = inc[sym - 1] + c->nfreqs[sym] + 1;
        p += interp_encode(p, inc, kByteSigma, kByteMaxFrame + kByteSigma);
    } else {
        ans_serialize_interp(c->nfreqs, size_t(frame), p);
    }
    c->prelude.resize(size_t(p - c->prelude.data()));

    // encoder table: {freq, base = exclusive prefix sum, smallest double >= 1/freq}
    std::vector<EncEntry> table(c->nfreqs.size());
    c->max_nfreq = 0;
    uint64_t base = 0;
    for (size_t s = 0; s < table.size(); s++) {
        const uint32_t fr = c->nfreqs[s];
        table[s].freq = fr;
        table[s].base = uint32_t(base);
        table[s].rcp = fr ? reciprocal_up(fr) : 0.0;
        base += fr;
        c->max_nfreq = std::max(c->max_nfreq, fr);
    }
    CU(c, c->enc_table.reserve(table.size() * sizeof(EncEntry)));
    // pageable source: cudaMemcpyAsync returns once the buffer is staged, so `table` may go out of scope
    CU(c, cudaMemcpyAsync(c->enc_table.p, table.data(), table.size() * sizeof(EncEntry), cudaMemcpyHostToDevice, c->stream));
    // ANSrfold, L3 encoder: hash of the hot values + a copy of the table whose first K entries sit at the hash slots
    c->rf_hash_ok = false;
    if (method == METHOD_RFOLD && !c->most_frequent.empty() && f <= 3 && table.size() >= c->most_frequent.size()) {
        const uint32_t K = uint32_t(c->most_frequent.size());
        const uint32_t bshift = 32u - (f + 6u); // K / 2 buckets
        std::vector<uint32_t> tags, slot_of_rank;
        std::vector<uint16_t> disp;
        uint32_t mul = 0x9E3779B1u;
        bool ok = false;
        for (int attempt = 0; attempt < 16 && !ok; attempt++, mul = mul * 0x01000193u + 0x7F4A7C15u) {
            mul |= 1u;
            ok = rfold_build_hash(c->most_frequent, K, mul, bshift, tags, disp, slot_of_rank);
            if (ok)
                break;
        }
        if (ok) {
            std::vector<EncEntry> table_rf(table);
            for (uint32_t r = 0; r < K; r++)
                table_rf[slot_of_rank[r]] = table[r];
            CU(c, c->enc_table_rf.reserve(table_rf.size() * sizeof(EncEntry)));
            CU(c, c->rf_tags.reserve(tags.size() * 4));
            CU(c, c->rf_disp.reserve(disp.size() * 2));
            CU(c, cudaMemcpyAsync(c->enc_table_rf.p, table_rf.data(), table_rf.size() * sizeof(EncEntry), cudaMemcpyHostToDevice, c->stream));
            CU(c, cudaMemcpyAsync(c->rf_tags.p, tags.data(), tags.size() * 4, cudaMemcpyHostToDevice, c->stream));
            CU(c, cudaMemcpyAsync(c->rf_disp.p, disp.data(), disp.size() * 2, cudaMemcpyHostToDevice, c->stream));
            CU(c, cudaStreamSynchronize(c->stream)); // the sources are locals
            c->rf_mul = mul;
            c->rf_bshift = bshift;
            c->rf_hash_ok = true;
        }
    }
    c->enc_method = method;
    c->enc_f = f;
    c->have_enc = true;
    c->stats.frame_size = frame;
    c->stats.n_symbols_model = uint32_t(c->nfreqs.size());
    c->stats.prelude_bytes = c->prelude.size();
    {
        uint64_t total = 0;
        for (uint64_t v : freqs)
            total += v;
        c->stats.model_flags = (total >= (uint64_t(1) << 31) ? B200ANS_MODEL_INT_ACCUMULATOR_QUIRK : 0u)
            | (method == METHOD_RFOLD && c->rf_fast_used ? B200ANS_MODEL_RFOLD_FAST : 0u);
    }
    c->stats.ms_model = float(now_ms() - t0);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_model_prelude(b200ans_ctx* c, uint8_t* dst, size_t capacity, size_t* out_bytes)
try {
    if (!c || !out_bytes)
        return B200ANS_E_ARG;
    if (!c->have_enc)
        return c->fail(B200ANS_E_STATE, "no encoder model");
    *out_bytes = c->prelude.size();
    if (!dst)
        return B200ANS_OK;
    if (capacity < c->prelude.size())
        return c->fail(B200ANS_E_CAPACITY, "prelude buffer too small");
    std::memcpy(dst, c->prelude.data(), c->prelude.size());
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_model_nfreqs(b200ans_ctx* c, uint32_t* dst, size_t capacity, size_t* out_count, uint64_t* frame_size)
try {
    if (!c || !out_count)
        return B200ANS_E_ARG;
    if (!c->have_enc)
        return c->fail(B200ANS_E_STATE, "no encoder model");
    *out_count = c->nfreqs.size();
    if (frame_size)
        *frame_size = c->frame;
    if (!dst)
        return B200ANS_OK;
    if (capacity < c->nfreqs.size())
        return c->fail(B200ANS_E_CAPACITY, "nfreqs buffer too small");
    std::copy(c->nfreqs.begin(), c->nfreqs.end(), dst);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// ans_X_decode::load (ans_fold.hpp:179-207, ans_int.hpp:116-164, ans_msb.hpp:179-204,
// ans_reorder_fold.hpp:235-279): prelude -> nfreqs -> slot tables, uploaded to the device.
extern "C" int b200ans_model_load(b200ans_ctx* c, int method, uint32_t f, const uint8_t* prelude, size_t prelude_bytes,
    size_t* consumed)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!prelude || !valid_method(method, f) || prelude_bytes < 3)
        return c->fail(B200ANS_E_ARG, "model_load: bad argument");
    CU(c, cudaSetDevice(c->device));
    const double t0 = now_ms();
    // The decoder tables on the device are a pure function of (method, fidelity, prelude bytes): a prelude that starts
    // with the bytes of the model already loaded -- the shards of one job, a container decoded again -- costs a memcmp.
    if (c->have_dec && c->dec_method == method && c->dec_f == f && !c->dec_prelude.empty() && prelude_bytes >= c->dec_prelude.size()
        && std::memcmp(prelude, c->dec_prelude.data(), c->dec_prelude.size()) == 0 && !std::getenv("B200ANS_NO_MODEL_CACHE")) {
        if (consumed)
            *consumed = c->dec_prelude.size();
        c->stats.frame_size = uint64_t(1) << c->dec_log2_frame;
        c->stats.n_symbols_model = c->dec_n_syms;
        c->stats.model_cache_hits++;
        c->stats.ms_model = float(now_ms() - t0);
        return B200ANS_OK;
    }
    c->have_dec = false;
    c->dec_prelude.clear();
    // untrusted bytes: every read is bounded by the caller's buffer (host/prelude.hpp ans_load_interp_checked)
    const uint8_t* p = prelude;
    const uint8_t* const end = prelude + prelude_bytes;
    const uint32_t thres = 1u << (f + 7);
    std::vector<uint32_t> most_frequent;
    if (method == METHOD_RFOLD) {
        if (prelude_bytes < 4 + 3)
            return c->fail(B200ANS_E_FORMAT, "model_load: truncated rfold prelude");
        uint32_t flag;
        std::memcpy(&flag, p, 4);
        p += 4;
        most_frequent.resize(thres);
        if (flag == 1) {
            if (prelude_bytes < 4 + size_t(4) * thres + 3)
                return c->fail(B200ANS_E_FORMAT, "model_load: truncated rfold prelude");
            std::memcpy(most_frequent.data(), p, size_t(4) * thres);
            p += size_t(4) * thres;
        } else if (flag == 0) {
            for (uint32_t i = 0; i < thres; i++)
                most_frequent[i] = i;
        } else {
            return c->fail(B200ANS_E_FORMAT, "model_load: bad rfold prelude flag");
        }
    }
    // one limit on both sides: b200ans_compress refuses plain-ANS alphabets beyond 2^28 values, frames beyond 2^28 slots
    const size_t limit = method == METHOD_INT ? (size_t(1) << 28)
                                              : (method == METHOD_MSB ? size_t(kMsbMaxSigma) : (size_t(1) << (f + 9)));
    size_t used = 0;
    std::vector<uint32_t> nfreqs;
    if (method == METHOD_BYTE) { // ans_byte_decode::load (ans_byte.hpp:178-190), bounded
        nfreqs.assign(kByteSigma, 0u);
        detail::word_unpacker r(p, end);
        detail::interp_get(r, nfreqs.data(), kByteSigma, 1, uint64_t(kByteMaxFrame + kByteSigma) + 1);
        if (r.overrun())
            return c->fail(B200ANS_E_FORMAT, "model_load: prelude cut off");
        uint32_t below = nfreqs[0];
        for (uint32_t sym = 1; sym < kByteSigma; sym++) {
            const uint32_t here = nfreqs[sym];
            if (here <= below)
                return c->fail(B200ANS_E_FORMAT, "model_load: byte model's running sums do not climb");
            nfreqs[sym] = here - below - 1;
            below = here;
        }
        used = r.words() * sizeof(uint32_t);
    } else if (!ans_load_interp_checked(p, end, limit, 28u, nfreqs, &used))
        return c->fail(B200ANS_E_FORMAT, "model_load: prelude cut off, or its header is implausible for this method");
    const size_t total_used = size_t(p - prelude) + used;
    if (consumed)
        *consumed = total_used;
    uint64_t frame = 0;
    for (uint32_t v : nfreqs)
        frame += v;
    if (!is_power_of_two(frame) || frame > (uint64_t(1) << 28) || (method == METHOD_BYTE && frame > kByteMaxFrame))
        return c->fail(B200ANS_E_FORMAT, "model_load: frequencies do not sum to a power of two");
    const uint32_t lg = uint32_t(63 - __builtin_clzll(frame));

    // rank/select index over the frame (kernels.cuh RankRec / RankEntry); `word(s)` is what the kernels get back for
    // present symbol s: its id (plain ANS) or the slot table's packed word (folded coders with a large frame)
    auto upload_rank = [&](auto&& word) -> int {
        const size_t words = (size_t(frame) + 31) / 32;
        std::vector<RankRec> rec(align_up(words, 2), RankRec { 0u, 0u });
        std::vector<RankEntry> ent;
        std::vector<uint32_t> ids;
        uint64_t base = 0;
        for (size_t s = 0; s < nfreqs.size(); s++) {
            if (!nfreqs[s])
                continue;
            rec[base >> 5].bits |= 1u << (base & 31);
            ent.push_back(RankEntry { uint32_t(base), nfreqs[s] });
            ids.push_back(word(uint32_t(s)));
            base += nfreqs[s];
        }
        uint32_t run = 0;
        for (size_t w = 0; w < words; w++) {
            rec[w].cnt = run;
            run += uint32_t(__builtin_popcount(rec[w].bits));
        }
        ent.resize(align_up(ent.size(), 2), RankEntry { 0u, 0u });
        CU(c, c->dec_bits.reserve(rec.size() * sizeof(RankRec)));
        CU(c, c->dec_cnt.reserve(std::max<size_t>(ids.size() * 4, 16)));
        CU(c, c->dec_ent.reserve(ent.size() * sizeof(RankEntry)));
        CU(c, cudaMemcpyAsync(c->dec_bits.p, rec.data(), rec.size() * sizeof(RankRec), cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpyAsync(c->dec_cnt.p, ids.data(), ids.size() * 4, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpyAsync(c->dec_ent.p, ent.data(), ent.size() * sizeof(RankEntry), cudaMemcpyHostToDevice, c->stream));
        // (pageable sources are staged before cudaMemcpyAsync returns: no synchronisation needed for the locals)
        c->dec_n_present = uint32_t(ids.size());
        // the same records with whole-group answers (kernels.cuh DEC_INT_SMEM_PURE) are built when a launch wants them
        c->dec_have_pure = false;
        c->dec_pure_possible = frame <= (uint64_t(1) << 19) && ids.size() < (size_t(1) << 23);
        if (c->dec_pure_possible) {
            c->h_rank_rec.swap(rec);
            c->h_rank_ent.swap(ent);
            c->h_rank_words = words;
        }
        return B200ANS_OK;
    };
    c->dec_fold_rank = false;
    if (coder_of(method) == METHOD_INT) {
        RET(upload_rank([](uint32_t s) { return s; }));
        c->dec_dense = c->dec_n_present == nfreqs.size() ? 1u : 0u;
    } else {
        auto packed = [&](uint32_t s) { // the slot table's word for symbol s: unfolded base value | raw bytes << 30
            uint32_t k = 0, value;
            if (method == METHOD_MSB)
                value = msb_unmap(s, k);
            else if (method == METHOD_RFOLD && s < thres)
                value = most_frequent[s] + thres; // ans_reorder_fold.hpp:214-215
            else
                value = fold_unmap(s, f, k);
            return value + (k << 30);
        };
        for (uint32_t fr : nfreqs)
            if (fr > 65535)
                return c->fail(B200ANS_E_FORMAT, "model_load: folded model with a frequency above 16 bits");
        if (frame * sizeof(DecEntry) > (size_t(128) << 10)) {
            // A frame beyond 2^14 slots: the 8-byte slot table no longer fits next to the decoder's rings in shared
            // memory (M = 2^18 .. 2^19 once the job reaches 2^32 symbols: 2 .. 4 MB to build here, upload and gather
            // from L2 per symbol).  Such models get the rank/select index instead: M / 4 + 12 sigma bytes.
            RET(upload_rank(packed));
            c->dec_dense = 0;
            c->dec_fold_rank = true;
        } else {
            // the M-entry slot table is filled on the device (kernels.cuh dec_slots_fill_kernel) from one record per
            // present symbol
            std::vector<SlotSym> syms;
            uint64_t base = 0;
            for (size_t s = 0; s < nfreqs.size(); s++) {
                if (nfreqs[s])
                    syms.push_back(SlotSym { uint32_t(base), nfreqs[s], packed(uint32_t(s)) });
                base += nfreqs[s];
            }
            CU(c, c->dec_slots.reserve(size_t(frame) * sizeof(DecEntry)));
            CU(c, c->dec_syms.reserve(syms.size() * sizeof(SlotSym)));
            CU(c, cudaMemcpyAsync(c->dec_syms.p, syms.data(), syms.size() * sizeof(SlotSym), cudaMemcpyHostToDevice, c->stream));
            dec_slots_fill_kernel<<<int(std::max<uint64_t>(1, (frame + 255) / 256)), 256, 0, c->stream>>>(c->dec_syms.as<SlotSym>(), uint32_t(syms.size()),
                uint32_t(frame), c->dec_slots.as<DecEntry>());
            c->stats.launches++;
            CU(c, cudaGetLastError());
        }
    }
    c->dec_method = method;
    c->dec_f = f;
    c->dec_log2_frame = lg;
    c->dec_n_syms = uint32_t(nfreqs.size());
    c->dec_prelude.assign(prelude, prelude + total_used);
    c->have_dec = true;
    c->stats.frame_size = frame;
    c->stats.n_symbols_model = uint32_t(nfreqs.size());
    c->stats.ms_model = float(now_ms() - t0);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// ==========================================================================================
// K2 / K3 launches
// ==========================================================================================
namespace {

int env_int(const char* name, int dflt)
{
    const char* v = std::getenv(name);
    return v && *v ? std::atoi(v) : dflt;
}

template <class K> int persistent_grid(b200ans_ctx* c, K kernel, int threads, size_t smem, uint32_t n_tasks, int warps)
{
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) {
        cudaGetLastError();
        per_sm = 1;
    }
    const uint32_t need = (n_tasks + uint32_t(warps) - 1) / uint32_t(warps);
    return int(std::max<uint32_t>(1, std::min<uint32_t>(need, uint32_t(per_sm) * uint32_t(c->num_sms))));
}

template <int METHOD> int launch_encode_t(b200ans_ctx* c, const EncArgs& a)
{
    const size_t rows = size_t(kEncWarps) * 8 * kEncRowBytes;
    const uint32_t n_tasks = (a.n_blocks + 7) / 8;
    const bool smem_table = a.n_entries <= uint32_t(kEncMaxSmemEntries);
    if (smem_table) {
        auto k = encode_kernel<METHOD, true>;
        const size_t smem = rows + size_t(a.n_entries) * sizeof(EncEntry);
        CU(c, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        const int grid = persistent_grid(c, k, kEncWarps * 32, smem, n_tasks, kEncWarps);
        k<<<grid, kEncWarps * 32, smem, c->stream>>>(a);
    } else {
        auto k = encode_kernel<METHOD, false>;
        const int grid = persistent_grid(c, k, kEncWarps * 32, rows, n_tasks, kEncWarps);
        k<<<grid, kEncWarps * 32, rows, c->stream>>>(a);
    }
    set_name(c->stats.encode_kernel, "encode_kernel");
    c->stats.launches++;
    CU(c, cudaGetLastError());
    return B200ANS_OK;
}

// Lane-per-block kernels need 32 blocks per warp: worth it once that still fills the GPU.
bool use_lpb(const b200ans_ctx* c, uint32_t n_blocks, bool encode)
{
    if (c->kernel_path == B200ANS_PATH_GROUP)
        return false;
    if (c->kernel_path == B200ANS_PATH_LPB)
        return true;
    // The L3 encoder beats the 4-lanes-per-block encoder from one full warp of blocks upwards (measured: 1e7 and
    // 1e8 symbols, 0.44 ms against 0.67 / 0.69 ms); the L3 decoder from ~3.5 warps of blocks per SM (1e8 symbols:
    // 24 414 blocks 0.39 ms against 0.54 ms, 12 207 blocks 0.78 ms against 0.74 ms).
    return encode ? n_blocks >= 32u : n_blocks >= uint32_t(c->num_sms) * 112u;
}

// Lane-per-block launch shape: ONE CTA per SM.  A task (32 blocks) runs for the whole kernel, so
// the tasks are spread in balanced waves: waves = ceil(tasks / (SMs * max warps)), and every CTA
// gets ceil(tasks / (SMs * waves)) warps -- 2^30 symbols are 4096 tasks = one wave of 28 warps per SM.
template <class K>
int launch_lpb(b200ans_ctx* c, K k, size_t table_bytes, size_t warp_bytes, uint32_t n_blocks, const void* args, int warp_cap = kLpbMaxWarps)
{
    const size_t room = c->max_smem > table_bytes + 1024 ? c->max_smem - table_bytes - 1024 : 0;
    const uint32_t max_warps = uint32_t(std::max<size_t>(1, std::min<size_t>(size_t(warp_cap), room / warp_bytes)));
    const uint32_t n_tasks = (n_blocks + 31) / 32;
    const uint32_t sms = uint32_t(c->num_sms);
    const uint32_t waves = (n_tasks + sms * max_warps - 1) / (sms * max_warps);
    const uint32_t warps = std::max<uint32_t>(1, std::min<uint32_t>(max_warps, (n_tasks + sms * waves - 1) / (sms * waves)));
    const uint32_t grid = std::min<uint32_t>(sms, (n_tasks + warps - 1) / warps);
    const size_t smem = table_bytes + 128 + size_t(warps) * warp_bytes;
    CU(c, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    void* params[] = { const_cast<void*>(args) };
    CU(c, cudaLaunchKernel(reinterpret_cast<const void*>(k), dim3(grid), dim3(warps * 32), params, smem, c->stream));
    c->stats.launches++;
    return B200ANS_OK;
}

template <int METHOD> int launch_lpb_encode_t(b200ans_ctx* c, const EncArgs& a)
{
    set_name(c->stats.encode_kernel, "lpb_encode_kernel");
    const bool smem_table = a.n_entries <= uint32_t(kLpbEncMaxSmemEntries);
    const size_t tb = smem_table ? size_t(a.n_entries) * sizeof(EncEntry) : 0;
    if (smem_table)
        return a.fast_div ? launch_lpb(c, lpb_encode_kernel<METHOD, true, true>, tb, kLpbEncWarpBytes, a.n_blocks, &a)
                          : launch_lpb(c, lpb_encode_kernel<METHOD, true, false>, tb, kLpbEncWarpBytes, a.n_blocks, &a);
    return a.fast_div ? launch_lpb(c, lpb_encode_kernel<METHOD, false, true>, tb, kLpbEncWarpBytes, a.n_blocks, &a)
                      : launch_lpb(c, lpb_encode_kernel<METHOD, false, false>, tb, kLpbEncWarpBytes, a.n_blocks, &a);
}

template <int METHOD, int TABLE> int launch_lpb_decode_tt(b200ans_ctx* c, const DecArgs& a, size_t table_bytes)
{
    // L3 decoder (lpb3_kernels.cuh) needs 16-byte aligned output; QI = 2 stores whole 32-byte sectors
    const uintptr_t out_addr = reinterpret_cast<uintptr_t>(a.out);
    set_name(c->stats.decode_kernel, "lpb_decode_kernel");
    if ((out_addr & 15) == 0 && a.block_syms % 8u == 0u && !env_int("B200ANS_NO_L3", 0)) {
        set_name(c->stats.decode_kernel, "lpb3_decode_kernel");
        const size_t tb = align_up(table_bytes, 16);
        const int shape = env_int("B200ANS_L3_DEC", 0);
        if ((out_addr & 31) == 0 && shape == 0)
            return launch_lpb(c, lpb3_decode_kernel<METHOD, TABLE, 2, 16>, tb + 256, L3DecGeom<2>::kWarpBytes, a.n_blocks, &a, 16);
        if ((out_addr & 31) == 0 && shape == 2)
            return launch_lpb(c, lpb3_decode_kernel<METHOD, TABLE, 2, 28>, tb + 256, L3DecGeom<2>::kWarpBytes, a.n_blocks, &a, 28);
        return launch_lpb(c, lpb3_decode_kernel<METHOD, TABLE, 1, 28>, tb + 128, L3DecGeom<1>::kWarpBytes, a.n_blocks, &a, 28);
    }
    return launch_lpb(c, lpb_decode_kernel<METHOD, TABLE>, align_up(table_bytes, 16), kLpbDecWarpBytes, a.n_blocks, &a);
}

// rank/select records with whole-group answers: worth it where the decoder is bound by shared-memory wavefronts, i.e.
// from about two rounds of tasks per SM upwards (below that the kernel waits on latency and the extra selects cost)
template <int METHOD> int launch_lpb3_decode_pure(b200ans_ctx* c, DecArgs a, size_t table_bytes)
{
    if (!c->dec_have_pure) { // first large decode with this model: rewrite the records of groups inside one symbol
        std::vector<RankRec>& pure = c->h_rank_rec;
        const std::vector<RankEntry>& ent = c->h_rank_ent;
        for (size_t w = 0; w < c->h_rank_words; w++) {
            if (pure[w].bits & ~1u)
                continue; // a symbol starts inside the group
            const uint32_t r = pure[w].cnt + (pure[w].bits & 1u) - 1u; // the symbol that holds slot 32 w
            pure[w].bits = ent[r].base | ((r & 0xFFFu) << 20);
            pure[w].cnt = kRankPureFlag | ((r >> 12) << 20) | ent[r].freq;
        }
        CU(c, c->dec_bits_pure.reserve(pure.size() * sizeof(RankRec)));
        CU(c, cudaMemcpyAsync(c->dec_bits_pure.p, pure.data(), pure.size() * sizeof(RankRec), cudaMemcpyHostToDevice, c->stream));
        c->dec_have_pure = true;
    }
    a.rank_rec = c->dec_bits_pure.as<RankRec>();
    set_name(c->stats.decode_kernel, "lpb3_decode_kernel");
    const size_t tb = align_up(table_bytes, 16);
    return launch_lpb(c, lpb3_decode_kernel<METHOD, DEC_INT_SMEM_PURE, 2, 16>, tb + 256, L3DecGeom<2>::kWarpBytes, a.n_blocks, &a, 16);
}

bool use_pure_records(const b200ans_ctx* c, const DecArgs& a)
{
    const int mode = env_int("B200ANS_PURE", -1); // 0 / 1 force, default by size
    if (!(c->dec_have_pure || c->dec_pure_possible) || mode == 0 || env_int("B200ANS_NO_L3", 0) || env_int("B200ANS_L3_DEC", 0))
        return false;
    if ((reinterpret_cast<uintptr_t>(a.out) & 31) != 0 || a.block_syms % 8u != 0u)
        return false;
    return mode == 1 || a.n_blocks >= uint32_t(c->num_sms) * 448u;
}

// ---- L3 encoder (lpb3_kernels.cuh): TMA input tiles + replicated shared-memory table ---------
typedef CUresult (*tensor_map_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

tensor_map_encode_fn tensor_map_encoder()
{
    static tensor_map_encode_fn fn = []() -> tensor_map_encode_fn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<tensor_map_encode_fn>(p);
    }();
    return fn;
}

// largest input value whose (folded) symbol is below n_entries; the mappings are monotone
uint32_t max_value_below(int method, uint32_t f, uint32_t n_entries)
{
    uint64_t lo = 0, hi = 0xFFFFFFFFull; // invariant: map(lo) < n_entries
    uint32_t k;
    auto sym = [&](uint32_t v) { return method == METHOD_MSB ? msb_map(v, k) : fold_map(v, f, k); };
    if (sym(0xFFFFFFFFu) < n_entries)
        return 0xFFFFFFFFu;
    while (lo + 1 < hi) {
        const uint64_t mid = (lo + hi) / 2;
        if (sym(uint32_t(mid)) < n_entries)
            lo = mid;
        else
            hi = mid;
    }
    return uint32_t(lo);
}

// returns B200ANS_OK after a launch, 1 when this input / model is not for the L3 kernel
template <int METHOD, class CFG> int launch_lpb3_encode_cfg(b200ans_ctx* c, const EncArgs& a)
{
    constexpr int TQ = CFG::TQ;
    if (!a.fast_div || a.block_syms % 128u != 0u || a.n / a.block_syms == 0 || env_int("B200ANS_NO_L3", 0))
        return 1;
    tensor_map_encode_fn encode = tensor_map_encoder();
    if (!encode)
        return 1;
    if (METHOD == METHOD_RFOLD && !c->rf_hash_ok)
        return 1;
    const uint32_t alphabet = METHOD == METHOD_INT ? a.n_entries + 1u : uint32_t(alphabet_bins(METHOD, a.fidelity, 0));
    // RFOLD: tags + displacements of the hot values' hash sit between the table and the tiles
    const size_t aux_bytes = METHOD == METHOD_RFOLD ? align_up((size_t(4) << (a.fidelity + 7)) + (size_t(2) << (a.fidelity + 6)), 1024) : 0;
    const uint32_t alloc = METHOD == METHOD_INT ? a.n_entries + 1u : a.n_entries;
    const uint32_t n_tasks = (a.n_blocks + 31) / 32;
    const uint32_t sms = uint32_t(c->num_sms);
    // replicas: as many as leave room for one balanced wave of warps (or 16 warps, whichever is less)
    const uint32_t want_warps = std::min<uint32_t>(CFG::MAXW, std::max<uint32_t>(1, (n_tasks + sms - 1) / sms));
    int rep = -1;
    uint32_t max_warps = 0;
    const int rep_hi = std::min(3, env_int("B200ANS_L3_REP", 3));
    for (int r = rep_hi; r >= 0; r--) {
        const size_t table_bytes = align_up((size_t(alloc) << r) * 16, 1024);
        // every address the mapping can produce must stay inside the CTA's shared-memory window
        if ((size_t(alphabet) << r) * 16 + 64 > c->max_smem || table_bytes + aux_bytes + 2048 > c->max_smem)
            continue;
        const uint32_t fit = uint32_t(std::min<size_t>(CFG::MAXW, (c->max_smem - table_bytes - aux_bytes - 1024) / L3Geom<CFG>::kWarpBytes));
        if (fit >= std::min<uint32_t>(want_warps, uint32_t(CFG::MAXW)) || r == 0) {
            rep = r;
            max_warps = fit;
            break;
        }
    }
    if (rep < 0 || max_warps == 0)
        return 1;
    max_warps = std::min<uint32_t>(max_warps, uint32_t(std::max(1, env_int("B200ANS_L3_WARPS", CFG::MAXW))));
    const uint32_t waves = (n_tasks + sms * max_warps - 1) / (sms * max_warps);
    const uint32_t warps = std::max<uint32_t>(1, std::min<uint32_t>(max_warps, (n_tasks + sms * waves - 1) / (sms * waves)));
    const uint32_t grid = std::min<uint32_t>(sms, (n_tasks + warps - 1) / warps);

    Enc3Args g {};
    g.rep_log2 = uint32_t(rep);
    g.table_alloc = alloc;
    g.n_full_blocks = uint32_t(a.n / a.block_syms);
    // folded coders: the largest value the model covers, and never 2^30 or more (ans_fold.hpp:198-200)
    g.max_value = METHOD == METHOD_INT ? 0xFFFFFFFFu : std::min(max_value_below(METHOD, a.fidelity, a.n_entries), (1u << 30) - 1u);
    EncArgs a2 = a;
    if (METHOD == METHOD_RFOLD) {
        g.rf_tags = c->rf_tags.as<uint32_t>();
        g.rf_disp = c->rf_disp.as<uint16_t>();
        g.rf_mul = c->rf_mul;
        g.rf_bshift = c->rf_bshift;
        a2.table = c->enc_table_rf.as<EncEntry>();
    }

    CUtensorMap tmap;
    const cuuint64_t dims[2] = { a.block_syms, g.n_full_blocks };
    const cuuint64_t strides[1] = { cuuint64_t(a.block_syms) * 4 };
    const cuuint32_t box[2] = { 4 * TQ, 32 };
    const cuuint32_t estr[2] = { 1, 1 };
    const int promo = env_int("B200ANS_L3_PROMO", 3);
    const CUresult rc = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint32_t*>(a.in), dims, strides, box, estr,
        CU_TENSOR_MAP_INTERLEAVE_NONE,
        TQ == 2 ? CU_TENSOR_MAP_SWIZZLE_32B : TQ == 4 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
        promo == 3 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : promo == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS)
        return 1;

    // the full window: symbols beyond the model read (and flag) whatever lies behind the table
    const size_t smem = c->max_smem;
    auto k = lpb3_encode_kernel<METHOD, 0, CFG>;
    auto k1 = lpb3_encode_kernel<METHOD, 1, CFG>;
    const bool f1 = METHOD == METHOD_FOLD && a.fidelity == 1;
    CU(c, cudaFuncSetAttribute(f1 ? k1 : k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    if (f1)
        k1<<<grid, warps * 32, smem, c->stream>>>(a2, g, tmap);
    else
        k<<<grid, warps * 32, smem, c->stream>>>(a2, g, tmap);
    set_name(c->stats.encode_kernel, "lpb3_encode_kernel");
    c->stats.launches++;
    CU(c, cudaGetLastError());
    return B200ANS_OK;
}

// Pseudo-adaptive mode through the L3 encoder (lpb3_kernels.cuh, PA = true): every lane keeps its block's model in shared
// memory.  Returns B200ANS_OK after a launch, 1 when the models do not fit one wave of warps (the caller then takes the
// four-lanes-per-block coder of pa_kernels.cuh).  B200ANS_NO_PA3=1 (testing) declines always.
struct Pa3Cfg {
    static constexpr int TQ = 4, STAGES = 2, DQ = 2, MAXW = 8;
};
template <int METHOD> int launch_pa3_encode(b200ans_ctx* c, const EncArgs& a, uint32_t max_lg)
{
    using CFG = Pa3Cfg;
    constexpr int TQ = CFG::TQ;
    if (a.block_syms % 128u != 0u || a.n / a.block_syms == 0 || max_lg > 15u || c->pa3_off || env_int("B200ANS_NO_PA3", 0)
        || env_int("B200ANS_NO_L3", 0))
        return 1;
    tensor_map_encode_fn encode = tensor_map_encoder();
    if (!encode)
        return 1;
    Enc3Args g {};
    // 32 KiB of reciprocals cover the frequencies up to 4096 (B200ANS_PA3_RCPN, testing: a shorter table)
    g.pa_rcp_n = std::min<uint32_t>(uint32_t(std::max(2, env_int("B200ANS_PA3_RCPN", 4097))), (1u << max_lg) + 1u);
    g.pa_tab_stride = uint32_t(align_up((size_t(a.pa_syms) + 2) * 2, 8));
    if (((g.pa_tab_stride >> 3) & 1u) == 0u)
        g.pa_tab_stride += 8; // an odd number of 8-byte bank pairs between the lanes' tables
    const uint32_t n_tasks = (a.n_blocks + 31) / 32;
    const uint32_t sms = uint32_t(c->num_sms);
    const size_t fixed = size_t(g.pa_rcp_n) * 8 + 2048; // reciprocals + alignment slack of the tiles
    const size_t per_warp = size_t(32) * g.pa_tab_stride + L3Geom<CFG>::kWarpBytes;
    if (c->max_smem < fixed + per_warp)
        return 1;
    const uint32_t max_warps = uint32_t(std::min<size_t>(CFG::MAXW, (c->max_smem - fixed) / per_warp));
    const uint32_t want_warps = (n_tasks + sms - 1) / sms;
    if (max_warps < std::min<uint32_t>(want_warps, 4u)) // more than one wave of tasks, or fewer than four warps per SM
        return 1;
    const uint32_t waves = (n_tasks + sms * max_warps - 1) / (sms * max_warps);
    const uint32_t warps = std::max<uint32_t>(1, std::min<uint32_t>(max_warps, (n_tasks + sms * waves - 1) / (sms * waves)));
    const uint32_t grid = std::min<uint32_t>(sms, (n_tasks + warps - 1) / warps);
    g.n_full_blocks = uint32_t(a.n / a.block_syms);
    g.max_value = METHOD == METHOD_INT ? 0xFFFFFFFFu : std::min(max_value_below(METHOD, a.fidelity, a.pa_syms), (1u << 30) - 1u);

    CUtensorMap tmap;
    const cuuint64_t dims[2] = { a.block_syms, g.n_full_blocks };
    const cuuint64_t strides[1] = { cuuint64_t(a.block_syms) * 4 };
    const cuuint32_t box[2] = { 4 * TQ, 32 };
    const cuuint32_t estr[2] = { 1, 1 };
    const CUresult rc = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint32_t*>(a.in), dims, strides, box, estr,
        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS)
        return 1;
    const size_t smem = align_up(size_t(g.pa_rcp_n) * 8 + size_t(warps) * 32 * g.pa_tab_stride, 1024) + 1024 + size_t(warps) * L3Geom<CFG>::kWarpBytes;
    auto k = lpb3_encode_kernel<METHOD, 0, CFG, true>;
    auto k1 = lpb3_encode_kernel<METHOD, 1, CFG, true>;
    const bool f1 = METHOD == METHOD_FOLD && a.fidelity == 1;
    CU(c, cudaFuncSetAttribute(f1 ? k1 : k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    if (f1)
        k1<<<grid, warps * 32, smem, c->stream>>>(a, g, tmap);
    else
        k<<<grid, warps * 32, smem, c->stream>>>(a, g, tmap);
    set_name(c->stats.encode_kernel, "lpb3_encode_kernel<PA>");
    c->stats.launches++;
    CU(c, cudaGetLastError());
    return B200ANS_OK;
}

template <int METHOD> int launch_lpb3_encode_t(b200ans_ctx* c, const EncArgs& a)
{
    return launch_lpb3_encode_cfg<METHOD, L3Cfg>(c, a);
}

// encode d_src into scratch slots + block_bytes using the current encoder model
int launch_encode(b200ans_ctx* c, const uint32_t* d_src, size_t n, uint32_t block_syms, uint32_t n_blocks,
    uint8_t* d_scratch, size_t stride, uint32_t* d_block_bytes)
{
    EncArgs a {};
    a.in = d_src;
    a.n = n;
    a.block_syms = block_syms;
    a.n_blocks = n_blocks;
    a.table = c->enc_table.as<EncEntry>();
    a.n_entries = uint32_t(c->nfreqs.size());
    a.log2_frame = c->log2_frame;
    a.fidelity = c->enc_f;
    a.mapping = c->mapping.as<uint32_t>();
    a.n_values = c->n_values;
    a.scratch = d_scratch;
    a.stride = stride;
    a.block_bytes = d_block_bytes;
    a.error_flag = c->d_err();
    a.fast_div = c->max_nfreq < 65536u ? 1u : 0u;
    CU(c, cudaMemsetAsync(c->d_err(), 0, 4, c->stream));
    if (use_lpb(c, n_blocks, true)) {
        int rc3 = 1;
        switch (coder_of(c->enc_method)) {
        case METHOD_INT: rc3 = launch_lpb3_encode_t<METHOD_INT>(c, a); break;
        case METHOD_FOLD: rc3 = launch_lpb3_encode_t<METHOD_FOLD>(c, a); break;
        case METHOD_MSB: rc3 = launch_lpb3_encode_t<METHOD_MSB>(c, a); break;
        case METHOD_RFOLD: rc3 = launch_lpb3_encode_t<METHOD_RFOLD>(c, a); break;
        default: break;
        }
        if (rc3 <= 0)
            return rc3;
        switch (coder_of(c->enc_method)) {
        case METHOD_INT: return launch_lpb_encode_t<METHOD_INT>(c, a);
        case METHOD_FOLD: return launch_lpb_encode_t<METHOD_FOLD>(c, a);
        case METHOD_MSB: return launch_lpb_encode_t<METHOD_MSB>(c, a);
        case METHOD_RFOLD: return launch_lpb_encode_t<METHOD_RFOLD>(c, a);
        }
    }
    switch (coder_of(c->enc_method)) {
    case METHOD_INT: return launch_encode_t<METHOD_INT>(c, a);
    case METHOD_FOLD: return launch_encode_t<METHOD_FOLD>(c, a);
    case METHOD_MSB: return launch_encode_t<METHOD_MSB>(c, a);
    case METHOD_RFOLD: return launch_encode_t<METHOD_RFOLD>(c, a);
    }
    return c->fail(B200ANS_E_STATE, "encode: no model");
}

template <int METHOD, int TABLE> int launch_decode_tt(b200ans_ctx* c, const DecArgs& a, size_t table_bytes)
{
    auto k = decode_kernel<METHOD, TABLE>;
    const size_t smem = table_bytes + size_t(kDecWarps) * 8 * kDecRowBytes;
    CU(c, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    const uint32_t n_tasks = (a.n_blocks + 7) / 8;
    const int grid = persistent_grid(c, k, kDecWarps * 32, smem, n_tasks, kDecWarps);
    k<<<grid, kDecWarps * 32, smem, c->stream>>>(a);
    set_name(c->stats.decode_kernel, "decode_kernel");
    c->stats.launches++;
    CU(c, cudaGetLastError());
    return B200ANS_OK;
}

template <int TABLE> int launch_int_decode_t(b200ans_ctx* c, const DecArgs& a, size_t table_bytes)
{
    auto k = int_decode_kernel<TABLE>;
    const size_t smem = table_bytes + size_t(kDecWarps) * 8 * kDecRowBytes;
    CU(c, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    const uint32_t n_tasks = (a.n_blocks + 7) / 8;
    const int grid = persistent_grid(c, k, kDecWarps * 32, smem, n_tasks, kDecWarps);
    k<<<grid, kDecWarps * 32, smem, c->stream>>>(a);
    set_name(c->stats.decode_kernel, "int_decode_kernel");
    c->stats.launches++;
    CU(c, cudaGetLastError());
    return B200ANS_OK;
}

int launch_decode(b200ans_ctx* c, const uint8_t* d_payload, const uint64_t* d_offsets, const uint32_t* d_block_bytes,
    size_t n, uint32_t block_syms, uint32_t n_blocks, uint32_t* d_dst)
{
    DecArgs a {};
    a.payload = d_payload;
    a.offsets = d_offsets;
    a.block_bytes = d_block_bytes;
    a.out = d_dst;
    a.n = n;
    a.block_syms = block_syms;
    a.n_blocks = n_blocks;
    a.log2_frame = c->dec_log2_frame;
    a.fidelity = c->dec_f;
    a.slots = c->dec_slots.as<DecEntry>();
    a.rank_rec = c->dec_bits.as<RankRec>();
    a.rank_sym = c->dec_cnt.as<uint32_t>();
    a.rank_ent = c->dec_ent.as<RankEntry>();
    a.n_present = c->dec_n_present;
    a.dense = c->dec_dense;
    a.error_flag = c->d_err(); // the caller zeroes it before the first launch of a call and reads it after the last
    const size_t frame = size_t(1) << c->dec_log2_frame;
    if (use_lpb(c, n_blocks, false)) {
        const size_t lbudget = c->max_smem > size_t(16) * kLpbDecWarpBytes + 4096 ? c->max_smem - size_t(16) * kLpbDecWarpBytes - 4096 : 0;
        if (coder_of(c->dec_method) == METHOD_INT) {
            const size_t tb = DecTables<METHOD_INT, DEC_INT_SMEM>::smem_bytes(uint32_t(frame), c->dec_n_present, c->dec_dense);
            if (tb <= std::min<size_t>(lbudget, size_t(100) << 10))
                return use_pure_records(c, a) ? launch_lpb3_decode_pure<METHOD_INT>(c, a, tb) : launch_lpb_decode_tt<METHOD_INT, DEC_INT_SMEM>(c, a, tb);
            return launch_lpb_decode_tt<METHOD_INT, DEC_INT_GLOBAL>(c, a, 0);
        }
        const size_t tb = align_up(frame * sizeof(DecEntry), 16);
        const bool smem = tb <= std::min<size_t>(lbudget, size_t(128) << 10);
        if (c->dec_fold_rank) { // large frame: rank/select index in shared memory (fewer warps if need be), else in L2
            const size_t rb = DecTables<METHOD_FOLD, DEC_INT_SMEM>::smem_bytes(uint32_t(frame), c->dec_n_present, 0);
            const size_t room = c->max_smem > size_t(4) * L3DecGeom<2>::kWarpBytes + 4096 ? c->max_smem - size_t(4) * L3DecGeom<2>::kWarpBytes - 4096 : 0;
            if (rb <= room && !env_int("B200ANS_RANK_GLOBAL", 0)) {
                if (c->dec_method == METHOD_FOLD && use_pure_records(c, a))
                    return launch_lpb3_decode_pure<METHOD_FOLD>(c, a, rb);
                switch (coder_of(c->dec_method)) {
                case METHOD_FOLD: return launch_lpb_decode_tt<METHOD_FOLD, DEC_INT_SMEM>(c, a, rb);
                case METHOD_MSB: return launch_lpb_decode_tt<METHOD_MSB, DEC_INT_SMEM>(c, a, rb);
                case METHOD_RFOLD: return launch_lpb_decode_tt<METHOD_RFOLD, DEC_INT_SMEM>(c, a, rb);
                }
            }
            switch (coder_of(c->dec_method)) {
            case METHOD_FOLD: return launch_lpb_decode_tt<METHOD_FOLD, DEC_INT_GLOBAL>(c, a, 0);
            case METHOD_MSB: return launch_lpb_decode_tt<METHOD_MSB, DEC_INT_GLOBAL>(c, a, 0);
            case METHOD_RFOLD: return launch_lpb_decode_tt<METHOD_RFOLD, DEC_INT_GLOBAL>(c, a, 0);
            }
        }
        // two copies of the slot table (kernels.cuh DEC_FOLD_SMEM2) when they leave room for 16 warps of rings and the
        // L3 decoder takes the input; B200ANS_DEC_TWICE=0 keeps one
        const bool l3 = (reinterpret_cast<uintptr_t>(a.out) & 31) == 0 && a.block_syms % 8u == 0u && !env_int("B200ANS_NO_L3", 0) && !env_int("B200ANS_L3_DEC", 0);
        if (smem && l3 && env_int("B200ANS_DEC_TWICE", 1) && 2 * tb + 256 + 1024 + size_t(16) * L3DecGeom<2>::kWarpBytes <= c->max_smem) {
            switch (coder_of(c->dec_method)) {
            case METHOD_FOLD: return launch_lpb_decode_tt<METHOD_FOLD, DEC_FOLD_SMEM2>(c, a, 2 * tb);
            case METHOD_MSB: return launch_lpb_decode_tt<METHOD_MSB, DEC_FOLD_SMEM2>(c, a, 2 * tb);
            case METHOD_RFOLD: return launch_lpb_decode_tt<METHOD_RFOLD, DEC_FOLD_SMEM2>(c, a, 2 * tb);
            }
        }
        switch (coder_of(c->dec_method)) {
        case METHOD_FOLD:
            return smem ? launch_lpb_decode_tt<METHOD_FOLD, DEC_FOLD_SMEM>(c, a, tb)
                        : launch_lpb_decode_tt<METHOD_FOLD, DEC_FOLD_GLOBAL>(c, a, 0);
        case METHOD_MSB:
            return smem ? launch_lpb_decode_tt<METHOD_MSB, DEC_FOLD_SMEM>(c, a, tb)
                        : launch_lpb_decode_tt<METHOD_MSB, DEC_FOLD_GLOBAL>(c, a, 0);
        case METHOD_RFOLD:
            return smem ? launch_lpb_decode_tt<METHOD_RFOLD, DEC_FOLD_SMEM>(c, a, tb)
                        : launch_lpb_decode_tt<METHOD_RFOLD, DEC_FOLD_GLOBAL>(c, a, 0);
        }
        return c->fail(B200ANS_E_STATE, "decode: no model");
    }
    const size_t rows = size_t(kDecWarps) * 8 * kDecRowBytes;
    const size_t budget = c->max_smem > rows + 1024 ? c->max_smem - rows - 1024 : 0;
    if (coder_of(c->dec_method) == METHOD_INT) {
        const size_t tb = DecTables<METHOD_INT, DEC_INT_SMEM>::smem_bytes(uint32_t(frame), c->dec_n_present, c->dec_dense);
        const bool words = c->dec_words_aligned && (reinterpret_cast<uintptr_t>(d_payload) & 3) == 0 && !env_int("B200ANS_NO_INTDEC", 0);
        if (tb <= std::min<size_t>(budget, size_t(100) << 10))
            return words ? launch_int_decode_t<DEC_INT_SMEM>(c, a, tb) : launch_decode_tt<METHOD_INT, DEC_INT_SMEM>(c, a, tb);
        return words ? launch_int_decode_t<DEC_INT_GLOBAL>(c, a, 0) : launch_decode_tt<METHOD_INT, DEC_INT_GLOBAL>(c, a, 0);
    }
    const size_t tb = align_up(frame * sizeof(DecEntry), 16);
    const bool smem = tb <= std::min<size_t>(budget, size_t(128) << 10);
    if (c->dec_fold_rank) {
        const size_t rb = DecTables<METHOD_FOLD, DEC_INT_SMEM>::smem_bytes(uint32_t(frame), c->dec_n_present, 0);
        if (rb <= budget && !env_int("B200ANS_RANK_GLOBAL", 0)) {
            switch (coder_of(c->dec_method)) {
            case METHOD_FOLD: return launch_decode_tt<METHOD_FOLD, DEC_INT_SMEM>(c, a, rb);
            case METHOD_MSB: return launch_decode_tt<METHOD_MSB, DEC_INT_SMEM>(c, a, rb);
            case METHOD_RFOLD: return launch_decode_tt<METHOD_RFOLD, DEC_INT_SMEM>(c, a, rb);
            }
        }
        switch (coder_of(c->dec_method)) {
        case METHOD_FOLD: return launch_decode_tt<METHOD_FOLD, DEC_INT_GLOBAL>(c, a, 0);
        case METHOD_MSB: return launch_decode_tt<METHOD_MSB, DEC_INT_GLOBAL>(c, a, 0);
        case METHOD_RFOLD: return launch_decode_tt<METHOD_RFOLD, DEC_INT_GLOBAL>(c, a, 0);
        }
    }
    switch (coder_of(c->dec_method)) {
    case METHOD_FOLD:
        return smem ? launch_decode_tt<METHOD_FOLD, DEC_FOLD_SMEM>(c, a, tb)
                    : launch_decode_tt<METHOD_FOLD, DEC_FOLD_GLOBAL>(c, a, 0);
    case METHOD_MSB:
        return smem ? launch_decode_tt<METHOD_MSB, DEC_FOLD_SMEM>(c, a, tb)
                    : launch_decode_tt<METHOD_MSB, DEC_FOLD_GLOBAL>(c, a, 0);
    case METHOD_RFOLD:
        return smem ? launch_decode_tt<METHOD_RFOLD, DEC_FOLD_SMEM>(c, a, tb)
                    : launch_decode_tt<METHOD_RFOLD, DEC_FOLD_GLOBAL>(c, a, 0);
    }
    return c->fail(B200ANS_E_STATE, "decode: no model");
}

// the two side streams (copy-in / copy-out of the host-buffer decode pipeline; the first also runs the pseudo-adaptive
// prelude coder next to the encoder) and their events, created on first use
int ensure_side_streams(b200ans_ctx* c)
{
    if (c->copy_in)
        return B200ANS_OK;
    CU(c, cudaStreamCreateWithFlags(&c->copy_in, cudaStreamNonBlocking));
    CU(c, cudaStreamCreateWithFlags(&c->copy_out, cudaStreamNonBlocking));
    for (auto& e : c->ev_in)
        CU(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : c->ev_dec)
        CU(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : c->ev_side)
        CU(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    return B200ANS_OK;
}

// one asynchronous read of the device control block (maximum, error flag, scan totals) + one synchronisation
int read_small(b200ans_ctx* c)
{
    CU(c, cudaMemcpyAsync(c->h_small(), c->small.p, sizeof(b200ans_ctx::SmallHost), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return B200ANS_OK;
}

int write_empty_container(b200ans_ctx* c, int method, uint32_t f, uint8_t* d_dst, size_t cap, size_t* out_bytes)
{
    if (c->block_syms == 0) { // the reference cannot represent an empty input either
        *out_bytes = 0;
        return B200ANS_OK;
    }
    if (cap < B200ANS_HEADER_BYTES)
        return c->fail(B200ANS_E_CAPACITY, "destination too small");
    Header h {};
    h.magic = B200ANS_MAGIC;
    h.version = 1;
    h.method = uint8_t(method);
    h.fidelity = uint8_t(f);
    h.block_syms = c->block_syms;
    CU(c, cudaMemcpyAsync(d_dst, &h, sizeof h, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    *out_bytes = B200ANS_HEADER_BYTES;
    return B200ANS_OK;
}

int check_error_flag(b200ans_ctx* c)
{
    uint32_t flag = 0;
    CU(c, cudaMemcpyAsync(&flag, c->d_err(), 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    if (flag)
        return c->fail(B200ANS_E_DOMAIN, "encode: input holds a symbol the model has no frequency for");
    return B200ANS_OK;
}

} // namespace

// ==========================================================================================
// staged encode / decode
// ==========================================================================================
extern "C" int b200ans_block_stride(const b200ans_ctx* c, int method, size_t* stride)
try {
    if (!c || !stride || c->block_syms == 0)
        return B200ANS_E_ARG;
    *stride = block_stride(method, c->block_syms);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_encode_blocks_device(b200ans_ctx* c, const uint32_t* d_src, size_t n, uint8_t* d_scratch,
    size_t stride, uint32_t* d_block_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_src || !d_scratch || !d_block_bytes || n == 0 || (reinterpret_cast<uintptr_t>(d_src) & 15)
        || (reinterpret_cast<uintptr_t>(d_scratch) & 31) || (stride & 31))
        return c->fail(B200ANS_E_ARG, "encode_blocks: null / unaligned argument (scratch and stride: 32 bytes)");
    if (!c->have_enc)
        return c->fail(B200ANS_E_STATE, "encode_blocks: build a model first");
    if (c->block_syms == 0 && n >= (uint64_t(1) << 32))
        return c->fail(B200ANS_E_ARG, "single-block mode is limited to 2^32-1 symbols");
    CU(c, cudaSetDevice(c->device));
    const uint32_t bs = c->block_syms ? c->block_syms : uint32_t(n);
    const uint64_t nb = (n + bs - 1) / bs;
    if (stride < block_stride(c->enc_method, bs))
        return c->fail(B200ANS_E_CAPACITY, "encode_blocks: stride below the worst-case block size");
    CU(c, cudaEventRecord(c->ev[0], c->stream));
    RET(launch_encode(c, d_src, n, bs, uint32_t(nb), d_scratch, stride, d_block_bytes));
    CU(c, cudaEventRecord(c->ev[1], c->stream));
    RET(check_error_flag(c));
    c->stats.ms_encode = elapsed(c->ev[0], c->ev[1]);
    c->stats.n_blocks = uint32_t(nb);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

namespace {
// `expected_total` (or ~0): what the size scan in front of this call (its sum is in the context's control block) must
// have added up to.  The size checks run on the device and the decoders return at once when one fails: ONE
// synchronisation, behind the decode (round 2; there were three -- 40-50 us of a 1.1 ms config-2 step).
int decode_blocks_impl(b200ans_ctx* c, const uint8_t* d_payload, const uint64_t* d_offsets, const uint32_t* d_block_bytes, size_t n,
    uint32_t* d_dst, uint64_t expected_total)
{
    if (!d_payload || !d_offsets || !d_block_bytes || !d_dst || n == 0)
        return c->fail(B200ANS_E_ARG, "decode_blocks: null argument");
    if (!c->have_dec)
        return c->fail(B200ANS_E_STATE, "decode_blocks: load a model first");
    if (c->block_syms == 0 && n >= (uint64_t(1) << 32))
        return c->fail(B200ANS_E_ARG, "single-block mode is limited to 2^32-1 symbols");
    CU(c, cudaSetDevice(c->device));
    const uint32_t bs = c->block_syms ? c->block_syms : uint32_t(n);
    const uint64_t nb = (n + bs - 1) / bs;
    CU(c, cudaMemsetAsync(c->d_err(), 0, 4, c->stream));
    {   // untrusted sizes: at least the four states, whole words for plain ANS
        const int grid = int(std::max<uint64_t>(1, std::min<uint64_t>((nb + 255) / 256, uint64_t(c->num_sms) * 4)));
        const bool totals = expected_total != ~uint64_t(0);
        check_block_sizes_kernel<<<grid, 256, 0, c->stream>>>(d_block_bytes, uint32_t(nb), coder_of(c->dec_method) == METHOD_INT && c->block_syms ? 3u : 0u, c->d_err(),
            totals ? c->d_totals() : nullptr, expected_total);
        c->stats.launches++;
        CU(c, cudaGetLastError());
    }
    CU(c, cudaEventRecord(c->ev[0], c->stream));
    RET(launch_decode(c, d_payload, d_offsets, d_block_bytes, n, bs, uint32_t(nb), d_dst));
    CU(c, cudaEventRecord(c->ev[1], c->stream));
    RET(read_small(c));
    c->stats.ms_decode = elapsed(c->ev[0], c->ev[1]);
    c->stats.n_blocks = uint32_t(nb);
    const uint32_t flag = c->h_small()->error_flag;
    if (flag == 4u)
        return c->fail(B200ANS_E_FORMAT, "decompress: block size table does not add up");
    if (flag == 3u)
        return c->fail(B200ANS_E_FORMAT, "decode: a block is shorter than its four states, or a plain-ANS block is not a whole number of words");
    if (flag)
        return c->fail(B200ANS_E_FORMAT, "decode: a block's payload was not consumed exactly (corrupt stream or wrong model)");
    return B200ANS_OK;
}
} // namespace

extern "C" int b200ans_decode_blocks_device(b200ans_ctx* c, const uint8_t* d_payload, const uint64_t* d_offsets,
    const uint32_t* d_block_bytes, size_t n, uint32_t* d_dst)
try {
    if (!c)
        return B200ANS_E_ARG;
    return decode_blocks_impl(c, d_payload, d_offsets, d_block_bytes, n, d_dst, ~uint64_t(0));
}
B200ANS_CATCH_INT

// ==========================================================================================
// decode without any allocation (no counterpart in the reference)
// ==========================================================================================
// b200ans_decompress_device keeps the per-block payload offsets in a grow-only cudaMalloc'ed buffer (ctx->offsets), which
// callers that may not allocate (CUDA graph capture, tight serving loops) cannot use.  Here that scratch -- the offsets the
// size scan writes and the decoder reads, plus the scan's totals -- is a file-scope device array of fixed size.
// Every GPU that loads the library holds these 513 KiB until the process exits.  ALL calls on a device share the array
// and it keeps the last call's contents: the scan rewrites every entry a call reads before the decoder starts (stream
// order), nothing has to start at zero; calls that overlap on the GPU (two contexts / streams on one device) race.
constexpr uint32_t kNoAllocMaxBlocks = B200ANS_NOALLOC_MAX_BLOCKS;
constexpr uint32_t kNoAllocHead = 32; // u64 words in front of the offsets: [0..1] the scan's totals
static __device__ __align__(256) uint64_t g_noalloc_scratch[kNoAllocHead + kNoAllocMaxBlocks];

namespace {

constexpr int kNoAllocMaxDevices = 64;
uint64_t* g_noalloc_addr[kNoAllocMaxDevices]; // device address of g_noalloc_scratch, per device (0: not resolved yet)

// In host code `g_noalloc_scratch` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
int noalloc_scratch(b200ans_ctx* c, uint64_t** out)
{
    if (c->device < 0 || c->device >= kNoAllocMaxDevices)
        return c->fail(B200ANS_E_ARG, "decode_blocks_noalloc: device index beyond the per-device address table");
    uint64_t* p = __atomic_load_n(&g_noalloc_addr[c->device], __ATOMIC_ACQUIRE);
    if (!p) {
        CU(c, cudaSetDevice(c->device));
        void* sym = nullptr;
        CU(c, cudaGetSymbolAddress(&sym, g_noalloc_scratch));
        if (!sym)
            return c->fail(B200ANS_E_CUDA, "decode_blocks_noalloc: cudaGetSymbolAddress returned no address");
        p = static_cast<uint64_t*>(sym);
        __atomic_store_n(&g_noalloc_addr[c->device], p, __ATOMIC_RELEASE); // every thread resolves the same address
    }
    *out = p;
    return B200ANS_OK;
}

} // namespace

extern "C" int b200ans_decode_blocks_noalloc(b200ans_ctx* c, const uint8_t* d_payload, size_t payload_bytes,
    const uint32_t* d_block_bytes, size_t n, uint32_t* d_dst)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_payload || !d_block_bytes || !d_dst || n == 0 || (reinterpret_cast<uintptr_t>(d_payload) & 15)
        || (reinterpret_cast<uintptr_t>(d_block_bytes) & 3))
        return c->fail(B200ANS_E_ARG, "decode_blocks_noalloc: null / unaligned argument (payload: 16 bytes, sizes: 4 bytes)");
    if (!c->have_dec)
        return c->fail(B200ANS_E_STATE, "decode_blocks_noalloc: load a model first (b200ans_model_load is the set-up step that may allocate)");
    if (c->block_syms == 0)
        return c->fail(B200ANS_E_ARG, "decode_blocks_noalloc: needs a blocked layout (block_syms > 0)");
    const uint32_t bs = c->block_syms;
    const uint64_t nb = (uint64_t(n) + bs - 1) / bs;
    if (nb > kNoAllocMaxBlocks)
        return c->fail(B200ANS_E_ARG, "decode_blocks_noalloc: more than B200ANS_NOALLOC_MAX_BLOCKS (65536) blocks; the static scratch holds no more -- use b200ans_decompress_device");
    if (env_int("B200ANS_PURE", -1) == 1 && !c->dec_have_pure && c->dec_pure_possible)
        return c->fail(B200ANS_E_STATE, "decode_blocks_noalloc: B200ANS_PURE=1 would build (and allocate) the whole-group rank records");
    CU(c, cudaSetDevice(c->device));
    uint64_t* scratch = nullptr;
    RET(noalloc_scratch(c, &scratch));
    uint64_t* d_totals = scratch;
    uint64_t* d_offsets = scratch + kNoAllocHead;

    CU(c, cudaMemsetAsync(c->d_err(), 0, 4, c->stream));
    {   // untrusted sizes, as in b200ans_decode_blocks_device; then offsets[b] = sum of the padded sizes in front of b
        const int grid = int(std::max<uint64_t>(1, std::min<uint64_t>((nb + 255) / 256, uint64_t(c->num_sms) * 4)));
        check_block_sizes_kernel<<<grid, 256, 0, c->stream>>>(d_block_bytes, uint32_t(nb), coder_of(c->dec_method) == METHOD_INT ? 3u : 0u, c->d_err());
        c->stats.launches++;
        CU(c, cudaGetLastError());
        scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(d_block_bytes, uint32_t(nb), d_offsets, d_totals, nullptr);
        c->stats.launches++;
        CU(c, cudaGetLastError());
        // the scan's padded total next to the error flag: one small read through the context's pinned block
        CU(c, cudaMemcpyAsync(c->d_totals(), d_totals, 16, cudaMemcpyDeviceToDevice, c->stream));
        RET(read_small(c));
        if (c->h_small()->error_flag)
            return c->fail(B200ANS_E_FORMAT, "decode: a block is shorter than its four states, or a plain-ANS block is not a whole number of words");
        if (c->h_small()->totals[0] > payload_bytes)
            return c->fail(B200ANS_E_FORMAT, "decode_blocks_noalloc: the block sizes add up to more than payload_bytes");
    }
    CU(c, cudaEventRecord(c->ev[0], c->stream));
    c->dec_words_aligned = true; // blocks start 16-byte aligned: the payload base is, the offsets are multiples of 16
    const int rc = launch_decode(c, d_payload, d_offsets, d_block_bytes, n, bs, uint32_t(nb), d_dst);
    c->dec_words_aligned = false;
    RET(rc);
    CU(c, cudaEventRecord(c->ev[1], c->stream));
    RET(read_small(c));
    c->stats.ms_decode = elapsed(c->ev[0], c->ev[1]);
    c->stats.n_blocks = uint32_t(nb);
    if (c->h_small()->error_flag)
        return c->fail(B200ANS_E_FORMAT, "decode: a block's payload was not consumed exactly (corrupt stream or wrong model)");
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// current model + d_src -> container (or reference wire format when block_syms == 0) at d_dst
extern "C" int b200ans_encode_device(b200ans_ctx* c, uint8_t* d_dst, size_t cap, const uint32_t* d_src, size_t n,
    size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_dst || (!d_src && n) || !out_bytes || (reinterpret_cast<uintptr_t>(d_src) & 15)
        || (reinterpret_cast<uintptr_t>(d_dst) & 15))
        return c->fail(B200ANS_E_ARG, "encode: null / unaligned argument");
    if (!c->have_enc)
        return c->fail(B200ANS_E_STATE, "encode: build a model first");
    if (n == 0) // an empty shard of a multi-GPU job: a container without blocks
        return write_empty_container(c, c->enc_method, c->enc_f, d_dst, cap, out_bytes);
    if (c->block_syms == 0 && n >= (uint64_t(1) << 32))
        return c->fail(B200ANS_E_ARG, "single-block mode is limited to 2^32-1 symbols");
    CU(c, cudaSetDevice(c->device));
    const bool raw = c->block_syms == 0;
    const uint32_t bs = raw ? uint32_t(n) : c->block_syms;
    const uint64_t nb64 = (n + bs - 1) / bs;
    if (nb64 > 0xFFFFFFF0ull)
        return c->fail(B200ANS_E_ARG, "too many blocks");
    const uint32_t nb = uint32_t(nb64);
    const size_t stride = block_stride(c->enc_method, bs);
    CU(c, c->scratch.reserve(stride * nb));
    CU(c, c->block_bytes.reserve(size_t(nb) * 4));
    CU(c, c->offsets.reserve(size_t(nb) * 8));

    const size_t pre = c->prelude.size();
    const size_t table_at = B200ANS_HEADER_BYTES + align_up(pre, 4);
    const size_t payload_at = align_up(table_at + size_t(nb) * 4, 16);
    uint64_t* d_header_total = nullptr;
    if (!raw) {
        // The front (header, prelude; the size table is the compaction kernel's) is known before any kernel runs,
        // except the payload total, which the size scan patches in on the device.  Everything after this copy is
        // stream-ordered kernels and one final synchronisation.
        if (cap < payload_at)
            return c->fail(B200ANS_E_CAPACITY, "encode: destination too small");
        c->front.assign(payload_at, 0);
        Header h {};
        h.magic = B200ANS_MAGIC;
        h.version = 1;
        h.method = uint8_t(c->enc_method);
        h.fidelity = uint8_t(c->enc_f);
        h.n_symbols = n;
        h.block_syms = bs;
        h.n_blocks = nb;
        h.prelude_bytes = uint32_t(pre);
        std::memcpy(c->front.data(), &h, sizeof h);
        std::memcpy(c->front.data() + B200ANS_HEADER_BYTES, c->prelude.data(), pre);
        CU(c, cudaMemcpyAsync(d_dst, c->front.data(), payload_at, cudaMemcpyHostToDevice, c->stream));
        d_header_total = reinterpret_cast<uint64_t*>(d_dst + offsetof(Header, payload_bytes));
    }
    CU(c, cudaEventRecord(c->ev[0], c->stream));
    RET(launch_encode(c, d_src, n, bs, nb, c->scratch.as<uint8_t>(), stride, c->block_bytes.as<uint32_t>()));
    CU(c, cudaEventRecord(c->ev[1], c->stream));
    scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(c->block_bytes.as<uint32_t>(), nb, c->offsets.as<uint64_t>(), c->d_totals(),
        d_header_total);
    c->stats.launches++;
    CU(c, cudaGetLastError());
    uint64_t totals[2] = { 0, 0 };
    if (raw) {
        // [prelude][payload], exactly ans_X_compress's output
        CU(c, cudaMemcpyAsync(totals, c->d_totals(), 16, cudaMemcpyDeviceToHost, c->stream));
        RET(check_error_flag(c)); // synchronises
        if (cap < pre + totals[1])
            return c->fail(B200ANS_E_CAPACITY, "encode: destination too small");
        CU(c, cudaMemcpyAsync(d_dst, c->prelude.data(), pre, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpyAsync(d_dst + pre, c->scratch.p, totals[1], cudaMemcpyDeviceToDevice, c->stream));
        CU(c, cudaEventRecord(c->ev[2], c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        *out_bytes = pre + totals[1];
        c->stats.payload_bytes = totals[1];
    } else {
        // compaction is launched without waiting for the total: it skips blocks that would not fit
        const int grid = int(std::max<uint32_t>(1, std::min<uint32_t>((nb + 7) / 8, uint32_t(c->num_sms) * 8)));
        compact_kernel<<<grid, 256, 0, c->stream>>>(c->scratch.as<uint8_t>(), stride, c->block_bytes.as<uint32_t>(),
            c->offsets.as<uint64_t>(), nb, d_dst + payload_at, reinterpret_cast<uint32_t*>(d_dst + table_at),
            uint64_t(cap - payload_at));
        c->stats.launches++;
        CU(c, cudaGetLastError());
        CU(c, cudaEventRecord(c->ev[2], c->stream));
        RET(read_small(c)); // the step's only synchronisation after the model upload
        if (c->h_small()->error_flag)
            return c->fail(B200ANS_E_DOMAIN, "encode: input holds a symbol the model has no frequency for");
        totals[0] = c->h_small()->totals[0];
        totals[1] = c->h_small()->totals[1];
        const size_t total = payload_at + totals[0];
        if (cap < total)
            return c->fail(B200ANS_E_CAPACITY, "encode: destination too small");
        *out_bytes = total;
        c->stats.payload_bytes = totals[1];
    }
    c->stats.ms_encode = elapsed(c->ev[0], c->ev[1]);
    c->stats.ms_compact = elapsed(c->ev[1], c->ev[2]);
    c->stats.n_blocks = nb;
    c->stats.total_bytes = *out_bytes;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// ==========================================================================================
// histogram exchange over NVLink peer memory
// ==========================================================================================
// The semi-static model needs ONE exchange per input: the sum of every rank's bins (8 KiB for ANSfold-1).  At eight
// GPUs an ncclAllReduce of that size costs 0.08-0.18 ms of a 3.1 ms step (profiles/r02m, r02s) -- launch, protocol
// set-up and a ring or tree of hops for a payload that fits one packet per peer.  When the ranks are processes on one
// NVSwitch node each of them exports an exchange buffer by CUDA IPC and maps everybody else's; the exchange is then ONE
// kernel of one CTA: store my bins into my slot of every peer's buffer (remote stores over NVLink), fence, raise my
// flag on every peer to this exchange's number (st.release.sys), wait until every peer's flag in MY buffer has reached
// it (ld.acquire.sys, bounded spin), add the slots up.  Slots are double-buffered by the parity of the exchange number:
// a rank cannot finish exchange e + 1 before every peer has entered it, i.e. has left exchange e, so nobody writes
// into a slot that is still being read.  Payloads beyond a slot (ANSrfold's raw bins, 128 MiB) stay with NCCL.
constexpr uint32_t kIpcFlagWords = 32;       // 64-bit words in front of the slots: one u32 flag per peer (+ padding)
constexpr uint32_t kIpcSlotWords = 8192 + 8; // bins + the running maximum
constexpr size_t kIpcBufBytes = (size_t(kIpcFlagWords) + size_t(2) * b200ans_ctx::kIpcMaxWorld * kIpcSlotWords) * 8;
constexpr uint32_t kIpcSpinLimit = 1u << 24; // x >= 0.2 us: a peer that has not arrived after seconds is gone

struct IpcArgs {
    uint64_t* peer[b200ans_ctx::kIpcMaxWorld];
    int rank, world;
    uint32_t epoch;
    uint64_t* bins;
    uint32_t nbins;
    uint32_t* d_max;
};

__global__ void __launch_bounds__(1024) ipc_allreduce_kernel(const IpcArgs a)
{
    const uint32_t par = a.epoch & 1u;
    const uint32_t words = a.nbins + 1u; // the maximum rides behind the bins
    const uint32_t mine_max = a.d_max ? *a.d_max : 0u;
    for (int p = 0; p < a.world; p++) {
        uint64_t* slot = a.peer[p] + kIpcFlagWords + size_t(par * b200ans_ctx::kIpcMaxWorld + uint32_t(a.rank)) * kIpcSlotWords;
        for (uint32_t i = threadIdx.x; i < words; i += blockDim.x)
            slot[i] = i < a.nbins ? a.bins[i] : uint64_t(mine_max);
    }
    __threadfence_system();
    __syncthreads();
    __shared__ uint32_t s_late;
    if (threadIdx.x == 0)
        s_late = 0u;
    __syncthreads();
    if (threadIdx.x < uint32_t(a.world)) {
        uint32_t* theirs = reinterpret_cast<uint32_t*>(a.peer[threadIdx.x]) + a.rank; // my flag in that peer's buffer
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(a.epoch) : "memory");
        const uint32_t* here = reinterpret_cast<const uint32_t*>(a.peer[a.rank]) + threadIdx.x;
        uint32_t seen, spins = 0;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(here) : "memory");
            if (int32_t(seen - a.epoch) >= 0)
                break;
            if (++spins > kIpcSpinLimit) {
                atomicExch(&s_late, 1u);
                break;
            }
            __nanosleep(200);
        }
    }
    __syncthreads();
    const uint64_t* base = a.peer[a.rank] + kIpcFlagWords + size_t(par * b200ans_ctx::kIpcMaxWorld) * kIpcSlotWords;
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) {
        uint64_t sum = 0, top = 0;
        for (int p = 0; p < a.world; p++) {
            const uint64_t v = __ldcg(base + size_t(p) * kIpcSlotWords + i); // (L2: the peers wrote it there)
            sum += v;
            top = max(top, v);
        }
        if (i < a.nbins)
            a.bins[i] = (s_late && i == 0u) ? ~uint64_t(0) : sum; // a count nobody can have: the host reports the peer that never came
        else if (a.d_max)
            *a.d_max = uint32_t(top);
    }
}

// ==========================================================================================
// one-shot paths
// ==========================================================================================
namespace {

// in place on the context's stream: bins summed, *d_max maximised over the ranks (one NCCL group)
int allreduce_hist(b200ans_ctx* c, uint64_t* d_bins, size_t nbins, uint32_t* d_max)
{
    if (c->ipc_world > 1 && nbins + 1 <= kIpcSlotWords && d_bins && nbins && !env_int("B200ANS_NO_IPC", 0)) {
        IpcArgs a {};
        for (int p = 0; p < c->ipc_world; p++)
            a.peer[p] = static_cast<uint64_t*>(c->ipc_peer[p]);
        a.rank = c->ipc_rank;
        a.world = c->ipc_world;
        a.epoch = ++c->ipc_epoch;
        a.bins = d_bins;
        a.nbins = uint32_t(nbins);
        a.d_max = d_max;
        ipc_allreduce_kernel<<<1, 1024, 0, c->stream>>>(a);
        c->stats.launches++;
        c->stats.peer_exchanges++;
        CU(c, cudaGetLastError());
        return B200ANS_OK;
    }
    if (!c->comm) // (a world of one still goes through NCCL: the same code path as N > 1)
        return B200ANS_OK;
    const NcclApi& n = nccl_api();
    if (!n.ok)
        return c->fail(B200ANS_E_NCCL, "libnccl.so.2 could not be loaded");
    int e = n.GroupStart();
    if (e)
        return c->nccl_fail(e, "ncclGroupStart");
    if (d_bins && nbins) {
        e = n.AllReduce(d_bins, d_bins, nbins, kNcclUint64, kNcclSum, c->comm, c->stream);
        if (e) {
            n.GroupEnd();
            return c->nccl_fail(e, "ncclAllReduce(histogram)");
        }
    }
    if (d_max) {
        e = n.AllReduce(d_max, d_max, 1, kNcclUint32, kNcclMax, c->comm, c->stream);
        if (e) {
            n.GroupEnd();
            return c->nccl_fail(e, "ncclAllReduce(max)");
        }
    }
    e = n.GroupEnd();
    if (e)
        return c->nccl_fail(e, "ncclGroupEnd");
    return B200ANS_OK;
}

// device histogram [-> all-reduce over the communicator] -> host bins
int histogram_to_host(b200ans_ctx* c, int method, uint32_t f, const uint32_t* d_src, size_t n, size_t nbins,
    std::vector<uint64_t>& host_bins, uint32_t* max_value, bool sharded)
{
    CU(c, c->bins.reserve(nbins * 8));
    CU(c, cudaMemsetAsync(c->bins.p, 0, nbins * 8, c->stream));
    if (n)
        RET(histogram_dispatch(c, method, f, d_src, n, c->bins.as<uint64_t>(), nbins, c->d_max()));
    if (sharded) {
        CU(c, cudaEventRecord(c->ev_ar[0], c->stream));
        RET(allreduce_hist(c, c->bins.as<uint64_t>(), nbins, max_value ? c->d_max() : nullptr));
        CU(c, cudaEventRecord(c->ev_ar[1], c->stream));
        c->ar_timed = true;
    }
    host_bins.resize(nbins);
    if (nbins * 8 <= b200ans_ctx::kPinnedBins) {
        // bins and the running maximum through pinned memory: both copies are asynchronous, one synchronisation
        CU(c, cudaMemcpyAsync(c->pinned, c->bins.p, nbins * 8, cudaMemcpyDeviceToHost, c->stream));
        RET(read_small(c));
        std::memcpy(host_bins.data(), c->pinned, nbins * 8);
        if (max_value)
            *max_value = c->h_small()->max_value;
        if (sharded && nbins && host_bins[0] == ~uint64_t(0))
            return c->fail(B200ANS_E_NCCL, "histogram exchange over peer memory: a rank did not arrive");
    } else {
        CU(c, cudaMemcpyAsync(host_bins.data(), c->bins.p, nbins * 8, cudaMemcpyDeviceToHost, c->stream));
        if (max_value)
            CU(c, cudaMemcpyAsync(max_value, c->d_max(), 4, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
    }
    if (sharded && c->ar_timed) {
        c->stats.ms_allreduce += elapsed(c->ev_ar[0], c->ev_ar[1]);
        c->ar_timed = false;
    }
    return B200ANS_OK;
}

// ANSrfold, the fast route (rfold_kernels.cuh, second half): sample -> candidates -> ONE pass that counts the candidates
// exactly and bounds everything else -> exact selection, or *done = false and the caller takes the raw histogram.
int rfold_fast_model(b200ans_ctx* c, uint32_t f, const uint32_t* d_src, size_t n, std::vector<uint64_t>& bins, uint32_t* max_value, bool* done)
{
    *done = false;
    const uint32_t K = 1u << (f + 7);
    const size_t nfolded = alphabet_bins(METHOD_RFOLD, f, 0);
    const size_t nb = size_t(1) << kRfSampleBucketsLog2;
    CU(c, c->rf_sbuckets.reserve(nb * 8));
    CU(c, c->rf_stags.reserve(nb * 4));
    CU(c, c->rf_state.reserve(sizeof(RfState)));
    CU(c, c->rf_hist.reserve(kRfDigits * 4));
    CU(c, c->rf_fast.reserve(sizeof(RfFast)));
    CU(c, c->rf_cand.reserve(kRfCand * 4));
    CU(c, c->rf_ctags.reserve(kRfCandSlots * 4));
    CU(c, c->rf_cdisp.reserve(size_t(2) << kRfCandBucketsLog2));
    CU(c, c->rf_ccount.reserve(kRfCandSlots * 8));
    CU(c, c->rf_sketch.reserve(size_t(8) << kRfSketchLog2));
    CU(c, c->rf_most.reserve(size_t(K) * 4));
    CU(c, c->bins.reserve(nfolded * 8));
    CU(c, cudaMemsetAsync(c->rf_sbuckets.p, 0, nb * 8, c->stream));
    CU(c, cudaMemsetAsync(c->rf_stags.p, 0xFF, nb * 4, c->stream));
    CU(c, cudaMemsetAsync(c->rf_state.p, 0, sizeof(RfState), c->stream));
    CU(c, cudaMemsetAsync(c->rf_hist.p, 0, kRfDigits * 4, c->stream));
    CU(c, cudaMemsetAsync(c->rf_fast.p, 0, sizeof(RfFast), c->stream));
    CU(c, cudaMemsetAsync(c->rf_cand.p, 0xFF, kRfCand * 4, c->stream));
    CU(c, cudaMemsetAsync(c->rf_ccount.p, 0, kRfCandSlots * 8, c->stream));
    CU(c, cudaMemsetAsync(c->rf_sketch.p, 0, size_t(8) << kRfSketchLog2, c->stream));
    CU(c, cudaMemsetAsync(c->bins.p, 0, nfolded * 8, c->stream));
    CU(c, cudaMemsetAsync(c->d_max(), 0, 4, c->stream));
    RfState* st = c->rf_state.as<RfState>();
    RfFast* fs = c->rf_fast.as<RfFast>();
    auto* buckets = c->rf_sbuckets.as<unsigned long long>();
    // 1. sample, 2. the fullest buckets
    const uint32_t n_sample_vec = uint32_t(std::min<uint64_t>(uint64_t(n) >> 2, uint64_t(1) << 17)); // 2^19 symbols
    const int sgrid = int(std::max<uint32_t>(1, std::min<uint32_t>((n_sample_vec + 255) / 256, uint32_t(c->num_sms) * 8)));
    rf_sample_kernel<<<sgrid, 256, 0, c->stream>>>(d_src, n, n_sample_vec, buckets, c->rf_stags.as<uint32_t>());
    const int bgrid = int(std::min<uint32_t>(uint32_t(nb / kRfThreads), uint32_t(c->num_sms) * 8));
    for (uint32_t p = 0; p < 2; p++) { // sample counts stay below 2^24
        const uint32_t shift = (1 - p) * kRfDigitBits;
        rf_digit_hist_kernel<<<bgrid, kRfThreads, 0, c->stream>>>(buckets, uint32_t(nb), shift, p == 0, st, c->rf_hist.as<unsigned int>(), st);
        rf_digit_pick_kernel<<<1, kRfDigits / 4, 0, c->stream>>>(c->rf_hist.as<unsigned int>(), shift, p == 0, kRfCand, st);
    }
    rf_cand_collect_kernel<<<bgrid, 256, 0, c->stream>>>(buckets, c->rf_stags.as<uint32_t>(), st, false, fs, c->rf_cand.as<uint32_t>());
    rf_cand_collect_kernel<<<bgrid, 256, 0, c->stream>>>(buckets, c->rf_stags.as<uint32_t>(), st, true, fs, c->rf_cand.as<uint32_t>());
    c->stats.launches += 7;
    std::vector<uint32_t> cand(kRfCand);
    CU(c, cudaMemcpyAsync(cand.data(), c->rf_cand.p, kRfCand * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    std::sort(cand.begin(), cand.end());
    cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
    while (!cand.empty() && cand.back() == 0xFFFFFFFFu)
        cand.pop_back();
    if (cand.size() < K)
        return B200ANS_OK; // too few distinct values in the sample: the raw histogram decides
    // perfect hash of the candidates
    std::vector<uint32_t> tags, slot_of;
    std::vector<uint16_t> disp;
    uint32_t mul = 0x9E3779B1u;
    bool ok = false;
    for (int attempt = 0; attempt < 16 && !ok; attempt++, mul = mul * 0x01000193u + 0x7F4A7C15u) {
        mul |= 1u;
        ok = rfold_build_hash(cand, kRfCandSlots, mul, 32u - kRfCandBucketsLog2, tags, disp, slot_of);
        if (ok)
            break;
    }
    if (!ok)
        return B200ANS_OK;
    CU(c, cudaMemcpyAsync(c->rf_ctags.p, tags.data(), kRfCandSlots * 4, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaMemcpyAsync(c->rf_cdisp.p, disp.data(), disp.size() * 2, cudaMemcpyHostToDevice, c->stream));
    std::vector<uint16_t> cslot(slot_of.begin(), slot_of.end()); // candidate -> slot, for the ranking
    CU(c, c->rf_cslot.reserve(kRfCand * 2));
    CU(c, cudaMemcpyAsync(c->rf_cslot.p, cslot.data(), cslot.size() * 2, cudaMemcpyHostToDevice, c->stream));
    // 3. the pass, 4. ids and the proof
    const size_t smem = (size_t(2) * kRfCandSlots + (size_t(1) << kRfSketchLog2) + nfolded) * 4 + (size_t(2) << kRfCandBucketsLog2);
    CU(c, cudaFuncSetAttribute(rf_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    rf_count_kernel<<<c->num_sms, kRfCountThreads, smem, c->stream>>>(d_src, n, f, c->rf_ctags.as<uint32_t>(), c->rf_cdisp.as<uint16_t>(), mul,
        uint32_t(nfolded), c->rf_ccount.as<unsigned long long>(), c->rf_sketch.as<unsigned long long>(), c->bins.as<unsigned long long>(), c->d_max());
    rf_cand_rank_kernel<<<(uint32_t(cand.size()) * 32 + kRfThreads - 1) / kRfThreads, kRfThreads, 0, c->stream>>>(c->rf_ctags.as<uint32_t>(),
        c->rf_ccount.as<unsigned long long>(), c->rf_cslot.as<uint16_t>(), uint32_t(cand.size()), K, f, uint32_t(nfolded), c->rf_most.as<uint32_t>(),
        c->bins.as<unsigned long long>(), fs);
    rf_verify_kernel<<<1, 1024, 0, c->stream>>>(c->rf_sketch.as<unsigned long long>(), K, fs);
    c->stats.launches += 3;
    CU(c, cudaGetLastError());
    RfFast h_fs;
    std::vector<uint32_t> most(K);
    bins.resize(nfolded);
    CU(c, cudaMemcpyAsync(&h_fs, fs, sizeof h_fs, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaMemcpyAsync(most.data(), c->rf_most.p, size_t(K) * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaMemcpyAsync(bins.data(), c->bins.p, nfolded * 8, cudaMemcpyDeviceToHost, c->stream));
    RET(read_small(c));
    *max_value = c->h_small()->max_value;
    if (!h_fs.ok)
        return B200ANS_OK; // a value outside the candidates could reach the K-th count (flat input, ties): not proven
    if (uint64_t(*max_value) + K >= (1u << 30))
        return c->fail(B200ANS_E_DOMAIN, "ANSrfold needs value + 2^(f+7) < 2^30");
    c->most_frequent = std::move(most);
    c->rfold_sigma = h_fs.n_valid; // (a lower bound: the fast route does not count the distinct values)
    c->n_values = *max_value + 1;
    CU(c, c->mapping.reserve(size_t(c->n_values) * 4));
    const int grid = int(std::max<uint32_t>(1, std::min<uint32_t>((c->n_values + kRfThreads - 1) / kRfThreads, uint32_t(c->num_sms) * 8)));
    rf_mapping_fill_kernel<<<grid, kRfThreads, 0, c->stream>>>(nullptr, K, c->n_values, c->mapping.as<uint32_t>());
    rf_mapping_hot_kernel<<<(K + kRfThreads - 1) / kRfThreads, kRfThreads, 0, c->stream>>>(nullptr, K, c->rf_most.as<uint32_t>(), c->mapping.as<uint32_t>());
    c->stats.launches += 2;
    c->map_f = f;
    c->have_mapping = true;
    *done = true;
    return B200ANS_OK;
}

// ans_X_encode::create (ans_fold.hpp:68-92, ans_int.hpp:38-68, ans_msb.hpp:77-101, ans_reorder_fold.hpp:70-130) with the
// input on the device: histogram (K1) [-> all-reduce] -> host normalise -> encoder tables.  *empty: nothing to model
// (every shard of the job is empty).
int model_from_device(b200ans_ctx* c, int method, uint32_t f, const uint32_t* d_src, size_t n, bool sharded, bool* empty)
{
    *empty = false;
    CU(c, cudaEventRecord(c->ev[3], c->stream));
    CU(c, cudaMemsetAsync(c->d_max(), 0, 4, c->stream));
    uint32_t max_value = 0;
    std::vector<uint64_t> bins;
    if (method == METHOD_INT) {
        // The reference takes the maximum first, then counts (ans_int.hpp:42-49).  K1 keeps the running maximum
        // while it counts, so alphabets that fit four shared-memory replicas need one pass: count into kOptBins
        // bins (larger values are dropped by the kernel's bounds check) and look at the maximum afterwards.
        constexpr size_t kOptBins = kHistSmemWords / 4 - 1;
        RET(histogram_to_host(c, METHOD_INT, f, d_src, n, kOptBins, bins, &max_value, sharded));
        // one limit on both sides: b200ans_model_load refuses alphabets beyond 2^28 values too
        if (max_value >= (1u << 28))
            return c->fail(B200ANS_E_DOMAIN, "plain ANS: values of 2^28 and more are not supported (one table entry per value up to the maximum)");
        if (max_value >= kOptBins)
            RET(histogram_to_host(c, METHOD_INT, f, d_src, n, size_t(max_value) + 1, bins, nullptr, sharded));
        else
            bins.resize(size_t(max_value) + 1);
    } else if (method == METHOD_BYTE) {
        RET(histogram_to_host(c, METHOD_BYTE, f, d_src, n, kByteSigma, bins, &max_value, sharded));
        if (max_value >= kByteSigma)
            return c->fail(B200ANS_E_DOMAIN, "ANSbyte codes bytes: values must be below 256");
    } else if (method == METHOD_RFOLD) {
        // K4 on the device: ONE optimistic pass counts the raw values into `guess` bins while it keeps the running
        // maximum (a second pass only when a value lies beyond the guess); selection of the 2^(f+7) most frequent
        // values, mapping[] and the folded histogram follow on the device from the raw bins (rfold_kernels.cuh).
        const uint32_t K = 1u << (f + 7);
        bool rf_done = false;
        {   // the fast route for large inputs whose folded alphabet fits shared memory (it proves its own exactness or
            // declines); B200ANS_RF_FAST = 0 / 1 switches it off / forces it for any size
            const int mode = env_int("B200ANS_RF_FAST", -1);
            bool done = false;
            if (!sharded && f <= 3 && n >= 4 && mode != 0 && (mode == 1 || n >= (size_t(1) << 24)))
                RET(rfold_fast_model(c, f, d_src, n, bins, &max_value, &done));
            c->rf_fast_used = done;
            rf_done = done;
        }
        if (!rf_done) {
        size_t guess = std::max<size_t>(c->rf_guess, std::min<size_t>(size_t(1) << 24, std::max<size_t>(65536, size_t(next_power_of_two(uint64_t(n) * 4)))));
        for (int attempt = 0;; attempt++) {
            CU(c, c->rf_raw.reserve(guess * 8));
            CU(c, cudaMemsetAsync(c->d_max(), 0, 4, c->stream));
            if (n && uint64_t(n) < (uint64_t(1) << 32) && guess > kHistSmemWords) {
                // 32-bit counters for the pass (they stay in L2), widened afterwards
                CU(c, c->rf_raw32.reserve(guess * 4));
                CU(c, cudaMemsetAsync(c->rf_raw32.p, 0, guess * 4, c->stream));
                RET(launch_histogram<METHOD_INT>(c, f, d_src, n, nullptr, guess, c->d_max(), c->rf_raw32.as<unsigned int>()));
                widen_bins_kernel<<<c->num_sms * 8, 256, 0, c->stream>>>(c->rf_raw32.as<unsigned int>(), c->rf_raw.as<unsigned long long>(), uint32_t(guess));
                c->stats.launches++;
            } else {
                CU(c, cudaMemsetAsync(c->rf_raw.p, 0, guess * 8, c->stream));
                if (n)
                    RET(histogram_dispatch(c, METHOD_INT, f, d_src, n, c->rf_raw.as<uint64_t>(), guess, c->d_max()));
            }
            if (sharded) {
                CU(c, cudaEventRecord(c->ev_ar[0], c->stream));
                RET(allreduce_hist(c, c->rf_raw.as<uint64_t>(), guess, c->d_max()));
                CU(c, cudaEventRecord(c->ev_ar[1], c->stream));
            }
            RET(read_small(c));
            if (sharded)
                c->stats.ms_allreduce += elapsed(c->ev_ar[0], c->ev_ar[1]);
            max_value = c->h_small()->max_value;
            if (max_value < guess || attempt == 1)
                break;
            guess = align_up(size_t(max_value) + 1, 4096);
        }
        c->rf_guess = guess;
        if (uint64_t(max_value) + K >= (1u << 30))
            return c->fail(B200ANS_E_DOMAIN, "ANSrfold needs value + 2^(f+7) < 2^30");
        const size_t nfolded = alphabet_bins(method, f, 0);
        CU(c, c->bins.reserve(nfolded * 8));
        const uint32_t count_bits = sharded ? 64u : uint32_t(64 - __builtin_clzll(uint64_t(n) | 1));
        c->have_mapping = false;
        RET(rfold_select_launch(c, f, c->rf_raw.as<uint64_t>(), max_value + 1, count_bits, c->bins.as<uint64_t>()));
        RfState st;
        c->most_frequent.resize(K);
        bins.resize(nfolded);
        CU(c, cudaMemcpyAsync(&st, c->rf_state.p, sizeof st, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaMemcpyAsync(c->most_frequent.data(), c->rf_most.p, size_t(K) * 4, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaMemcpyAsync(bins.data(), c->bins.p, nfolded * 8, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        c->rfold_sigma = st.sigma;
        if (!st.reorder) {
            c->most_frequent.clear();
            if (st.sigma != 0 && max_value >= K)
                return c->fail(B200ANS_E_DOMAIN,
                    "ANSrfold: fewer than 2^(f+7) distinct values and a value >= 2^(f+7): the reference's decoder returns wrong values for this input");
        }
        c->n_values = max_value + 1;
        c->map_f = f;
        c->have_mapping = true;
        }
    } else {
        RET(histogram_to_host(c, method, f, d_src, n, alphabet_bins(method, f, 0), bins, &max_value, sharded));
        if (max_value >= (1u << 30)) // the decoder keeps the byte count in bits 30-31 (ans_fold.hpp:198-200)
            return c->fail(B200ANS_E_DOMAIN, "folded coders need values < 2^30");
    }
    CU(c, cudaEventRecord(c->ev[4], c->stream));
    CU(c, cudaEventSynchronize(c->ev[4]));
    const float ms_hist = elapsed(c->ev[3], c->ev[4]) - c->stats.ms_allreduce;
    {
        bool any = false;
        for (uint64_t v : bins)
            any = any || v != 0;
        if (!any) { // every shard of the job is empty
            *empty = true;
            return B200ANS_OK;
        }
    }
    RET(b200ans_model_build(c, method, f, bins.data(), bins.size()));
    c->stats.ms_histogram = ms_hist;
    return B200ANS_OK;
}

// ans_X_compress with device-resident buffers: histogram (K1) [-> all-reduce] -> host normalise -> encode (K2)
int compress_device_impl(b200ans_ctx* c, int method, uint32_t f, uint8_t* d_dst, size_t cap, const uint32_t* d_src, size_t n,
    size_t* out_bytes, bool sharded)
{
    if (!d_dst || !out_bytes || (!d_src && n) || !valid_method(method, f)
        || (reinterpret_cast<uintptr_t>(d_src) & 15) || (reinterpret_cast<uintptr_t>(d_dst) & 15))
        return c->fail(B200ANS_E_ARG, "compress: bad argument (device buffers must be 16-byte aligned)");
    CU(c, cudaSetDevice(c->device));
    sharded = sharded && c->comm;
    c->stats.ms_histogram = c->stats.ms_encode = c->stats.ms_compact = c->stats.ms_model = c->stats.ms_allreduce = 0.f;
    if (n == 0 && !sharded)
        return write_empty_container(c, method, f, d_dst, cap, out_bytes);
    if (sharded && c->block_syms == 0)
        return c->fail(B200ANS_E_ARG, "sharded compression writes blocked containers (block_syms > 0)");
    bool empty = false;
    RET(model_from_device(c, method, f, d_src, n, sharded, &empty));
    if (empty)
        return write_empty_container(c, method, f, d_dst, cap, out_bytes);
    const float ms_hist = c->stats.ms_histogram;
    if (n == 0) { // this rank's shard is empty: it took part in the reduction, its container holds no blocks
        c->stats.ms_histogram = ms_hist;
        return write_empty_container(c, method, f, d_dst, cap, out_bytes);
    }
    RET(b200ans_encode_device(c, d_dst, cap, d_src, n, out_bytes));
    c->stats.ms_histogram = ms_hist;
    return B200ANS_OK;
}

} // namespace

extern "C" int b200ans_compress_device(b200ans_ctx* c, int method, uint32_t f, uint8_t* d_dst, size_t cap,
    const uint32_t* d_src, size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    return compress_device_impl(c, method, f, d_dst, cap, d_src, n, out_bytes, false);
}
B200ANS_CATCH_INT

// ans_X_encode::create(in_u32, n) (ans_fold.hpp:68-92, ans_int.hpp:38-68, ans_msb.hpp:77-101, ans_reorder_fold.hpp:70-130)
// as a call of its own: device histogram (ANSrfold: + selection of the most frequent values) -> host normalisation ->
// prelude + device encoder tables.  The model becomes the context's current encoder model (b200ans_model_prelude /
// _nfreqs / _most_frequent read it, b200ans_encode_device codes with it).
extern "C" int b200ans_model_create_device(b200ans_ctx* c, int method, uint32_t f, const uint32_t* d_src, size_t n)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_src || n == 0 || !valid_method(method, f) || (reinterpret_cast<uintptr_t>(d_src) & 15))
        return c->fail(B200ANS_E_ARG, "model_create: bad argument (d_src must be 16-byte aligned, n > 0)");
    CU(c, cudaSetDevice(c->device));
    c->stats.ms_histogram = c->stats.ms_model = c->stats.ms_allreduce = 0.f;
    bool empty = false;
    return model_from_device(c, method, f, d_src, n, false, &empty);
}
B200ANS_CATCH_INT

// the same for a host buffer: the symbols are copied to the device (they stay in the context's input buffer)
extern "C" int b200ans_model_create(b200ans_ctx* c, int method, uint32_t f, const uint32_t* src, size_t n)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!src || n == 0 || !valid_method(method, f))
        return c->fail(B200ANS_E_ARG, "model_create: bad argument");
    CU(c, cudaSetDevice(c->device));
    CU(c, c->io_in.reserve(std::max<size_t>(n * 4, 16)));
    CU(c, cudaMemcpyAsync(c->io_in.p, src, n * 4, cudaMemcpyHostToDevice, c->stream));
    return b200ans_model_create_device(c, method, f, c->io_in.as<uint32_t>(), n);
}
B200ANS_CATCH_INT

// ANSrfold: most_frequent[] and sigma of the current encoder model (ans_reorder_fold.hpp:180-184); count 0 when the
// model keeps the identity mapping
extern "C" int b200ans_model_most_frequent(b200ans_ctx* c, uint32_t* dst, size_t capacity, size_t* out_count, uint64_t* sigma)
try {
    if (!c || !out_count)
        return B200ANS_E_ARG;
    if (!c->have_enc || c->enc_method != METHOD_RFOLD)
        return c->fail(B200ANS_E_STATE, "no ANSrfold encoder model");
    *out_count = c->most_frequent.size();
    if (sigma)
        *sigma = c->rfold_sigma;
    if (!dst)
        return B200ANS_OK;
    if (capacity < c->most_frequent.size())
        return c->fail(B200ANS_E_CAPACITY, "most_frequent buffer too small");
    std::copy(c->most_frequent.begin(), c->most_frequent.end(), dst);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// ==========================================================================================
// ans_byte (ans_byte.hpp): the byte coder of the vbyte+ANS baselines (methods.hpp:432-482)
// ==========================================================================================
// Same 4-state coder as plain ANS over 256 symbols, with its own normaliser (frame <= 4096) and prelude.  METHOD_BYTE
// runs through every uint32 entry point (values below 256); these take and return BYTES: the symbols are widened to
// uint32 on the device in front of the block coders and narrowed behind them.
namespace {

__global__ void __launch_bounds__(256) widen_bytes_kernel(const uint8_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t n)
{
    const uint64_t n4 = n >> 2; // (in is 4-byte aligned: our own buffer)
    const uint32_t* in4 = reinterpret_cast<const uint32_t*>(in);
    uint4* out4 = reinterpret_cast<uint4*>(out);
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += uint64_t(gridDim.x) * blockDim.x) {
        const uint32_t w = __ldg(in4 + i);
        out4[i] = make_uint4(w & 0xFFu, (w >> 8) & 0xFFu, (w >> 16) & 0xFFu, w >> 24);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3u))
        out[(n4 << 2) + threadIdx.x] = in[(n4 << 2) + threadIdx.x];
}

__global__ void __launch_bounds__(256) narrow_bytes_kernel(const uint32_t* __restrict__ in, uint8_t* __restrict__ out, uint64_t n)
{
    const uint64_t n4 = n >> 2;
    const uint4* in4 = reinterpret_cast<const uint4*>(in);
    uint32_t* out4 = reinterpret_cast<uint32_t*>(out);
    for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += uint64_t(gridDim.x) * blockDim.x) {
        const uint4 q = __ldg(in4 + i);
        out4[i] = (q.x & 0xFFu) | ((q.y & 0xFFu) << 8) | ((q.z & 0xFFu) << 16) | (q.w << 24);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3u))
        out[(n4 << 2) + threadIdx.x] = uint8_t(in[(n4 << 2) + threadIdx.x]);
}

int byte_grid(const b200ans_ctx* c, size_t n) { return int(std::max<size_t>(1, std::min<size_t>(((n >> 2) + 255) / 256, size_t(c->num_sms) * 8))); }

} // namespace

// ans_byte_compress (ans_byte.hpp:236-277) with device buffers: d_src holds n bytes (4-byte aligned), d_dst the container
extern "C" int b200ans_byte_compress_device(b200ans_ctx* c, uint8_t* d_dst, size_t cap, const uint8_t* d_src, size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_dst || !out_bytes || (!d_src && n) || (reinterpret_cast<uintptr_t>(d_src) & 3) || (reinterpret_cast<uintptr_t>(d_dst) & 15))
        return c->fail(B200ANS_E_ARG, "byte_compress: bad argument (source 4-byte, destination 16-byte aligned)");
    CU(c, cudaSetDevice(c->device));
    CU(c, c->io_in.reserve(std::max<size_t>(n * 4, 16)));
    if (n) {
        widen_bytes_kernel<<<byte_grid(c, n), 256, 0, c->stream>>>(d_src, c->io_in.as<uint32_t>(), n);
        c->stats.launches++;
        CU(c, cudaGetLastError());
    }
    return compress_device_impl(c, METHOD_BYTE, 0, d_dst, cap, c->io_in.as<uint32_t>(), n, out_bytes, false);
}
B200ANS_CATCH_INT

// ans_byte_decompress (ans_byte.hpp:279-305): n bytes to d_dst (4-byte aligned)
extern "C" int b200ans_byte_decompress_device(b200ans_ctx* c, uint8_t* d_dst, size_t n, const uint8_t* d_src, size_t src_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!d_dst && n) || !d_src || (reinterpret_cast<uintptr_t>(d_dst) & 3))
        return c->fail(B200ANS_E_ARG, "byte_decompress: bad argument (destination 4-byte aligned)");
    CU(c, cudaSetDevice(c->device));
    CU(c, c->io_in.reserve(std::max<size_t>(align_up(n * 4, 32), 32)));
    RET(b200ans_decompress_device(c, METHOD_BYTE, 0, c->io_in.as<uint32_t>(), n, d_src, src_bytes));
    if (n) {
        narrow_bytes_kernel<<<byte_grid(c, n), 256, 0, c->stream>>>(c->io_in.as<uint32_t>(), d_dst, n);
        c->stats.launches++;
        CU(c, cudaGetLastError());
    }
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// host buffers
extern "C" int b200ans_byte_compress(b200ans_ctx* c, uint8_t* dst, size_t cap, const uint8_t* src, size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!dst || !out_bytes || (!src && n))
        return c->fail(B200ANS_E_ARG, "byte_compress: bad argument");
    CU(c, cudaSetDevice(c->device));
    const size_t bound = std::min(b200ans_compress_bound(c, METHOD_BYTE, n), std::max<size_t>(cap + 64, 4096));
    CU(c, c->io_bytes.reserve(std::max<size_t>(align_up(n, 16), 16)));
    CU(c, c->io_out.reserve(bound));
    if (n)
        CU(c, cudaMemcpyAsync(c->io_bytes.p, src, n, cudaMemcpyHostToDevice, c->stream));
    size_t bytes = 0;
    RET(b200ans_byte_compress_device(c, c->io_out.as<uint8_t>(), std::min(bound, c->io_out.cap), c->io_bytes.as<uint8_t>(), n, &bytes));
    if (bytes > cap)
        return c->fail(B200ANS_E_CAPACITY, "byte_compress: destination too small");
    if (bytes)
        CU(c, cudaMemcpyAsync(dst, c->io_out.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    *out_bytes = bytes;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_byte_decompress(b200ans_ctx* c, uint8_t* dst, size_t n, const uint8_t* src, size_t src_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!dst && n) || !src)
        return c->fail(B200ANS_E_ARG, "byte_decompress: bad argument");
    CU(c, cudaSetDevice(c->device));
    if (n == 0 && c->block_syms == 0)
        return B200ANS_OK;
    CU(c, c->io_out.reserve(std::max<size_t>(align_up(src_bytes, 16), 64)));
    CU(c, c->io_bytes.reserve(std::max<size_t>(align_up(n, 16), 16)));
    CU(c, cudaMemcpyAsync(c->io_out.p, src, src_bytes, cudaMemcpyHostToDevice, c->stream));
    RET(b200ans_byte_decompress_device(c, c->io_bytes.as<uint8_t>(), n, c->io_out.as<uint8_t>(), src_bytes));
    if (n)
        CU(c, cudaMemcpyAsync(dst, c->io_bytes.p, n, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// the byte normaliser on its own (ans_byte.hpp:41-97); host only
extern "C" int b200ans_byte_adjust_freqs(const uint64_t* freqs, uint16_t* out)
try {
    if (!freqs || !out)
        return B200ANS_E_ARG;
    uint64_t total = 0;
    for (int i = 0; i < 256; i++)
        total += freqs[i];
    if (total == 0)
        return B200ANS_E_DOMAIN;
    const std::array<uint16_t, 256> nf = byte_adjust_freqs(freqs);
    std::copy(nf.begin(), nf.end(), out);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// ==========================================================================================
// multi-GPU: communicator + sharded compression
// ==========================================================================================
extern "C" int b200ans_comm_unique_id(uint8_t id[B200ANS_COMM_ID_BYTES])
try {
    if (!id)
        return B200ANS_E_ARG;
    const NcclApi& n = nccl_api();
    if (!n.ok)
        return B200ANS_E_NCCL;
    NcclUniqueId u;
    if (n.GetUniqueId(&u) != 0)
        return B200ANS_E_NCCL;
    std::memcpy(id, u.internal, B200ANS_COMM_ID_BYTES);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_comm_release(b200ans_ctx* c)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (c->comm && c->own_comm && nccl_api().ok) {
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
        nccl_api().CommDestroy(c->comm);
    }
    c->comm = nullptr;
    c->own_comm = false;
    c->comm_rank = 0;
    c->comm_world = 1;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_comm_init(b200ans_ctx* c, const uint8_t id[B200ANS_COMM_ID_BYTES], int rank, int world)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!id || world < 1 || rank < 0 || rank >= world)
        return c->fail(B200ANS_E_ARG, "comm_init: bad rank / world");
    const NcclApi& n = nccl_api();
    if (!n.ok)
        return c->fail(B200ANS_E_NCCL, "libnccl.so.2 could not be loaded");
    b200ans_comm_release(c);
    CU(c, cudaSetDevice(c->device));
    NcclUniqueId u;
    std::memcpy(u.internal, id, B200ANS_COMM_ID_BYTES);
    void* comm = nullptr;
    const int e = n.CommInitRank(&comm, world, u, rank);
    if (e)
        return c->nccl_fail(e, "ncclCommInitRank");
    c->comm = comm;
    c->own_comm = true;
    c->comm_rank = rank;
    c->comm_world = world;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_comm_attach(b200ans_ctx* c, void* nccl_comm, int rank, int world)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!nccl_comm || world < 1 || rank < 0 || rank >= world)
        return c->fail(B200ANS_E_ARG, "comm_attach: bad communicator / rank / world");
    if (!nccl_api().ok)
        return c->fail(B200ANS_E_NCCL, "libnccl.so.2 could not be loaded");
    b200ans_comm_release(c);
    c->comm = nccl_comm;
    c->own_comm = false;
    c->comm_rank = rank;
    c->comm_world = world;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

namespace {
void ipc_release(b200ans_ctx* c)
{
    if (!c->ipc_buf)
        return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    for (int p = 0; p < c->ipc_world; p++)
        if (p != c->ipc_rank && c->ipc_peer[p])
            cudaIpcCloseMemHandle(c->ipc_peer[p]);
    cudaFree(c->ipc_buf);
    c->ipc_buf = nullptr;
    for (auto& p : c->ipc_peer)
        p = nullptr;
    c->ipc_world = 0;
    c->ipc_rank = 0;
    c->ipc_epoch = 0;
}
} // namespace

extern "C" int b200ans_comm_ipc_export(b200ans_ctx* c, uint8_t handle[B200ANS_IPC_HANDLE_BYTES])
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!handle)
        return c->fail(B200ANS_E_ARG, "comm_ipc_export: no room for the handle");
    static_assert(sizeof(cudaIpcMemHandle_t) == B200ANS_IPC_HANDLE_BYTES, "CUDA IPC handle size");
    CU(c, cudaSetDevice(c->device));
    ipc_release(c);
    CU(c, cudaMalloc(&c->ipc_buf, kIpcBufBytes));
    CU(c, cudaMemset(c->ipc_buf, 0, kIpcBufBytes));
    CU(c, cudaDeviceSynchronize());
    cudaIpcMemHandle_t h;
    CU(c, cudaIpcGetMemHandle(&h, c->ipc_buf));
    std::memcpy(handle, &h, sizeof h);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_comm_ipc_attach(b200ans_ctx* c, int rank, int world, const uint8_t* handles)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!handles || world < 1 || world > b200ans_ctx::kIpcMaxWorld || rank < 0 || rank >= world)
        return c->fail(B200ANS_E_ARG, "comm_ipc_attach: bad rank / world (at most 16 ranks) / handles");
    if (!c->ipc_buf)
        return c->fail(B200ANS_E_ARG, "comm_ipc_attach: b200ans_comm_ipc_export comes first");
    CU(c, cudaSetDevice(c->device));
    for (int p = 0; p < c->ipc_world; p++) // attached before: those mappings go first
        if (p != c->ipc_rank && c->ipc_peer[p]) {
            cudaIpcCloseMemHandle(c->ipc_peer[p]);
            c->ipc_peer[p] = nullptr;
        }
    c->ipc_world = 0;
    for (int p = 0; p < world; p++) {
        if (p == rank) {
            c->ipc_peer[p] = c->ipc_buf;
            continue;
        }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + size_t(p) * B200ANS_IPC_HANDLE_BYTES, sizeof h);
        void* ptr = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            for (int q = 0; q < p; q++)
                if (q != rank && c->ipc_peer[q]) {
                    cudaIpcCloseMemHandle(c->ipc_peer[q]);
                    c->ipc_peer[q] = nullptr;
                }
            return c->fail(B200ANS_E_NCCL, "comm_ipc_attach: cudaIpcOpenMemHandle failed (ranks must be processes on one NVLink node)");
        }
        c->ipc_peer[p] = ptr;
    }
    c->ipc_rank = rank;
    c->ipc_world = world;
    c->ipc_epoch = 0;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_comm_ipc_release(b200ans_ctx* c)
try {
    if (!c)
        return B200ANS_E_ARG;
    ipc_release(c);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_allreduce_hist(b200ans_ctx* c, uint64_t* d_bins, size_t nbins, uint32_t* d_max)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!d_bins && nbins) || (!d_bins && !d_max))
        return c->fail(B200ANS_E_ARG, "allreduce_hist: nothing to reduce");
    CU(c, cudaSetDevice(c->device));
    return allreduce_hist(c, d_bins, nbins, d_max);
}
B200ANS_CATCH_INT

extern "C" int b200ans_compress_sharded_device(b200ans_ctx* c, int method, uint32_t f, uint8_t* d_dst, size_t cap,
    const uint32_t* d_src, size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    return compress_device_impl(c, method, f, d_dst, cap, d_src, n, out_bytes, true);
}
B200ANS_CATCH_INT

extern "C" int b200ans_decompress_device(b200ans_ctx* c, int method, uint32_t f, uint32_t* d_dst, size_t n,
    const uint8_t* d_src, size_t src_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!d_dst && n) || !d_src || !valid_method(method, f) || (reinterpret_cast<uintptr_t>(d_src) & 15))
        return c->fail(B200ANS_E_ARG, "decompress: bad argument (d_src must be 16-byte aligned)");
    CU(c, cudaSetDevice(c->device));
    c->stats.ms_decode = c->stats.ms_model = 0.f;
    if (c->block_syms == 0) {
        // reference wire format: [prelude][payload]; n and src_bytes come from the caller (methods.hpp:541-546)
        if (n == 0)
            return B200ANS_OK;
        const size_t head = std::min<size_t>(src_bytes, size_t(64) << 20);
        std::vector<uint8_t> front(head);
        CU(c, cudaMemcpyAsync(front.data(), d_src, head, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        size_t used = 0;
        RET(b200ans_model_load(c, method, f, front.data(), head, &used));
        if (src_bytes < used + 32)
            return c->fail(B200ANS_E_FORMAT, "decompress: stream shorter than prelude + states");
        CU(c, c->offsets.reserve(8));
        CU(c, c->block_bytes.reserve(4));
        const uint64_t off = used;
        const uint32_t sz = uint32_t(src_bytes - used);
        if (src_bytes - used > 0x7FFFFF00ull)
            return c->fail(B200ANS_E_ARG, "single-block payloads are limited to 2 GiB");
        CU(c, cudaMemcpyAsync(c->offsets.p, &off, 8, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpyAsync(c->block_bytes.p, &sz, 4, cudaMemcpyHostToDevice, c->stream));
        return b200ans_decode_blocks_device(c, d_src, c->offsets.as<uint64_t>(), c->block_bytes.as<uint32_t>(), n, d_dst);
    }
    if (src_bytes < B200ANS_HEADER_BYTES)
        return c->fail(B200ANS_E_FORMAT, "decompress: shorter than a header");
    // header and (normally) the whole prelude in one copy
    const size_t head = std::min<size_t>(src_bytes, B200ANS_HEADER_BYTES + 4032);
    c->front.resize(std::max(c->front.size(), head));
    CU(c, cudaMemcpyAsync(c->front.data(), d_src, head, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    Header h;
    std::memcpy(&h, c->front.data(), sizeof h);
    if (h.magic != B200ANS_MAGIC || h.version != 1)
        return c->fail(B200ANS_E_FORMAT, "decompress: bad magic / version");
    if (h.flags & kFlagPseudoAdaptive)
        return c->fail(B200ANS_E_FORMAT, "decompress: pseudo-adaptive container (use b200ans_pa_decompress)");
    if (h.method != method || (h.fidelity != f && (method == METHOD_FOLD || method == METHOD_RFOLD)))
        return c->fail(B200ANS_E_FORMAT, "decompress: container was written by another method");
    if (h.n_symbols != n)
        return c->fail(B200ANS_E_ARG, "decompress: symbol count does not match the container");
    if (n == 0)
        return B200ANS_OK;
    if (h.block_syms == 0 || h.block_syms % 4 != 0 || h.n_blocks != (n + h.block_syms - 1) / h.block_syms)
        return c->fail(B200ANS_E_FORMAT, "decompress: inconsistent block geometry");
    const size_t table_at = B200ANS_HEADER_BYTES + align_up(h.prelude_bytes, 4);
    const size_t payload_at = align_up(table_at + size_t(h.n_blocks) * 4, 16);
    if (payload_at > src_bytes || h.payload_bytes > src_bytes - payload_at || h.prelude_bytes < 3)
        return c->fail(B200ANS_E_FORMAT, "decompress: container truncated");
    std::vector<uint8_t> prelude(h.prelude_bytes);
    const size_t have = std::min<size_t>(h.prelude_bytes, head - B200ANS_HEADER_BYTES);
    std::memcpy(prelude.data(), c->front.data() + B200ANS_HEADER_BYTES, have);
    if (have < h.prelude_bytes) {
        CU(c, cudaMemcpyAsync(prelude.data() + have, d_src + B200ANS_HEADER_BYTES + have, h.prelude_bytes - have,
            cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
    }
    RET(b200ans_model_load(c, method, f, prelude.data(), prelude.size(), nullptr));
    CU(c, c->offsets.reserve(size_t(h.n_blocks) * 8));
    const uint32_t* d_sizes = reinterpret_cast<const uint32_t*>(d_src + table_at);
    scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(d_sizes, h.n_blocks, c->offsets.as<uint64_t>(), c->d_totals(), nullptr);
    c->stats.launches++;
    CU(c, cudaGetLastError());
    const uint32_t saved = c->block_syms;
    c->block_syms = h.block_syms;
    c->dec_words_aligned = true; // container blocks start 16-byte aligned (offsets come from our own scan)
    // (the scan's total is compared with the header on the device, in front of the decoder)
    const int rc = decode_blocks_impl(c, d_src + payload_at, c->offsets.as<uint64_t>(), d_sizes, n, d_dst, h.payload_bytes);
    c->dec_words_aligned = false;
    c->block_syms = saved;
    return rc;
}
B200ANS_CATCH_INT

// ---- host buffers: what methods.hpp's encode()/decode() hand over ----------------------------
namespace {

// ---- pageable host buffers -----------------------------------------------------------------------------------------
// cudaMemcpyAsync on pageable memory is staged by the driver at about 11 GB/s (profiles/r02aq: 11.2 GB/s end to end
// against 46.5 GB/s with page-locked buffers) -- and that is the memory the reference's callers hold (std::vector,
// methods.hpp:535-546).  Copies of 256 MiB and more on unregistered memory therefore go through the library's own
// page-locked staging buffers: kStageThreads host threads each move every kStageThreads-th 16 MiB chunk, double-buffered,
// on a stream of their own, so that the host-side memcpy of one chunk overlaps the PCIe transfer of the others.
// B200ANS_NO_STAGING=1 keeps the plain copies.
bool host_pageable(const void* p)
{
    cudaPointerAttributes a {};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
}

bool use_staging(const void* host, size_t bytes)
{
    // (from 256 MiB: setting the staging buffers up once costs about what a plain copy of that size does)
    const size_t min_bytes = size_t(std::max(1, env_int("B200ANS_STAGE_MIN_MB", 256))) << 20; // (testing: a lower threshold)
    return bytes >= min_bytes && !env_int("B200ANS_NO_STAGING", 0) && host_pageable(host);
}

int ensure_stager(b200ans_ctx* c)
{
    if (c->stage_ready)
        return B200ANS_OK;
    for (int w = 0; w < b200ans_ctx::kStageThreads; w++) {
        CU(c, cudaStreamCreateWithFlags(&c->stage_st[w], cudaStreamNonBlocking));
        CU(c, cudaEventCreateWithFlags(&c->stage_done[w], cudaEventDisableTiming));
        for (int b = 0; b < 2; b++) {
            CU(c, cudaHostAlloc(reinterpret_cast<void**>(&c->stage_buf[w][b]), b200ans_ctx::kStageChunk, cudaHostAllocDefault));
            CU(c, cudaEventCreateWithFlags(&c->stage_ev[w][b], cudaEventDisableTiming));
        }
    }
    CU(c, cudaEventCreateWithFlags(&c->stage_fork, cudaEventDisableTiming));
    c->stage_ready = true;
    return B200ANS_OK;
}

// host (pageable) -> device, ordered behind everything queued on `stream` so far; `stream` waits for the copies
int staged_h2d(b200ans_ctx* c, void* d_dst, const void* h_src, size_t bytes, cudaStream_t stream)
{
    RET(ensure_stager(c));
    const int T = std::max(1, std::min(int(b200ans_ctx::kStageThreads), env_int("B200ANS_STAGE_THREADS", b200ans_ctx::kStageThreads)));
    constexpr size_t CH = b200ans_ctx::kStageChunk;
    const size_t n_chunks = (bytes + CH - 1) / CH;
    CU(c, cudaEventRecord(c->stage_fork, stream));
    std::atomic<int> bad { 0 };
    std::thread workers[b200ans_ctx::kStageThreads];
    int started = 0;
    try {
    for (int w = 0; w < T; w++, started++)
        workers[w] = std::thread([=, &bad]() {
            if (cudaSetDevice(c->device) != cudaSuccess || cudaStreamWaitEvent(c->stage_st[w], c->stage_fork, 0) != cudaSuccess) {
                bad = 1;
                return;
            }
            size_t round = 0;
            for (size_t k = size_t(w); k < n_chunks; k += T, round++) {
                const int b = int(round & 1);
                const size_t at = k * CH, len = std::min(CH, bytes - at);
                if (round >= 2 && cudaEventSynchronize(c->stage_ev[w][b]) != cudaSuccess) // the copy that last read this buffer
                    bad = 1;
                std::memcpy(c->stage_buf[w][b], static_cast<const uint8_t*>(h_src) + at, len);
                if (cudaMemcpyAsync(static_cast<uint8_t*>(d_dst) + at, c->stage_buf[w][b], len, cudaMemcpyHostToDevice, c->stage_st[w]) != cudaSuccess
                    || cudaEventRecord(c->stage_ev[w][b], c->stage_st[w]) != cudaSuccess)
                    bad = 1;
            }
            if (cudaEventRecord(c->stage_done[w], c->stage_st[w]) != cudaSuccess)
                bad = 1;
        });
    } catch (...) { // no thread to be had: the chunks of the missing workers stay uncopied, the call fails
        bad = 1;
    }
    for (int w = 0; w < started; w++)
        workers[w].join();
    if (bad) {
        cudaGetLastError();
        return c->fail(B200ANS_E_CUDA, "staged host-to-device copy failed");
    }
    for (int w = 0; w < T; w++)
        CU(c, cudaStreamWaitEvent(stream, c->stage_done[w], 0));
    return B200ANS_OK;
}

// device -> host (pageable), ordered behind everything queued on `stream` so far; when `ready` is given, the chunk that
// ends at byte offset `end` also waits for ready(end) (an event recorded on some stream, or null).  Returns with the
// bytes in place.
template <class Ready> int staged_d2h(b200ans_ctx* c, void* h_dst, const void* d_src, size_t bytes, cudaStream_t stream, Ready ready)
{
    RET(ensure_stager(c));
    const int T = std::max(1, std::min(int(b200ans_ctx::kStageThreads), env_int("B200ANS_STAGE_THREADS", b200ans_ctx::kStageThreads)));
    constexpr size_t CH = b200ans_ctx::kStageChunk;
    const size_t n_chunks = (bytes + CH - 1) / CH;
    CU(c, cudaEventRecord(c->stage_fork, stream));
    std::atomic<int> bad { 0 };
    std::thread workers[b200ans_ctx::kStageThreads];
    int started = 0;
    try {
    for (int w = 0; w < T; w++, started++)
        workers[w] = std::thread([=, &bad]() {
            if (cudaSetDevice(c->device) != cudaSuccess || cudaStreamWaitEvent(c->stage_st[w], c->stage_fork, 0) != cudaSuccess) {
                bad = 1;
                return;
            }
            size_t round = 0, prev_at = 0, prev_len = 0;
            int prev_b = -1;
            auto drain = [&]() { // the chunk before this one: wait for its copy, hand it to the caller's buffer
                if (prev_b < 0)
                    return;
                if (cudaEventSynchronize(c->stage_ev[w][prev_b]) != cudaSuccess)
                    bad = 1;
                std::memcpy(static_cast<uint8_t*>(h_dst) + prev_at, c->stage_buf[w][prev_b], prev_len);
            };
            for (size_t k = size_t(w); k < n_chunks; k += T, round++) {
                const int b = int(round & 1);
                const size_t at = k * CH, len = std::min(CH, bytes - at);
                cudaEvent_t dep = ready(at + len);
                if (dep && cudaStreamWaitEvent(c->stage_st[w], dep, 0) != cudaSuccess)
                    bad = 1;
                if (cudaMemcpyAsync(c->stage_buf[w][b], static_cast<const uint8_t*>(d_src) + at, len, cudaMemcpyDeviceToHost, c->stage_st[w]) != cudaSuccess
                    || cudaEventRecord(c->stage_ev[w][b], c->stage_st[w]) != cudaSuccess)
                    bad = 1;
                drain(); // (buffer 1 - b is free again after this)
                prev_b = b, prev_at = at, prev_len = len;
            }
            drain();
        });
    } catch (...) { // no thread to be had: the chunks of the missing workers stay uncopied, the call fails
        bad = 1;
    }
    for (int w = 0; w < started; w++)
        workers[w].join();
    if (bad) {
        cudaGetLastError();
        return c->fail(B200ANS_E_CUDA, "staged device-to-host copy failed");
    }
    return B200ANS_OK;
}

int compress_host_impl(b200ans_ctx* c, int method, uint32_t f, uint8_t* dst, size_t cap, const uint32_t* src, size_t n,
    size_t* out_bytes, bool sharded)
{
    if (!dst || !out_bytes || (!src && n) || !valid_method(method, f))
        return c->fail(B200ANS_E_ARG, "compress: bad argument");
    CU(c, cudaSetDevice(c->device));
    const size_t bound = std::min(b200ans_compress_bound(c, method, n), std::max<size_t>(cap + 64, 4096));
    CU(c, c->io_in.reserve(std::max<size_t>(n * 4, 16)));
    CU(c, c->io_out.reserve(bound));
    CU(c, cudaEventRecord(c->ev[5], c->stream));
    // (a staged copy that fails -- no thread, no page-locked memory -- is redone as a plain one)
    if (n && !(use_staging(src, n * 4) && staged_h2d(c, c->io_in.p, src, n * 4, c->stream) == B200ANS_OK))
        CU(c, cudaMemcpyAsync(c->io_in.p, src, n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaEventRecord(c->ev[6], c->stream));
    size_t bytes = 0;
    RET(compress_device_impl(c, method, f, c->io_out.as<uint8_t>(), std::min(bound, c->io_out.cap), c->io_in.as<uint32_t>(), n, &bytes, sharded));
    if (bytes > cap)
        return c->fail(B200ANS_E_CAPACITY, "compress: destination too small");
    CU(c, cudaEventRecord(c->ev[7], c->stream));
    if (bytes && !(use_staging(dst, bytes) && staged_d2h(c, dst, c->io_out.p, bytes, c->stream, [](size_t) { return cudaEvent_t(nullptr); }) == B200ANS_OK))
        CU(c, cudaMemcpyAsync(dst, c->io_out.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaEventRecord(c->ev[3], c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->stats.ms_h2d = elapsed(c->ev[5], c->ev[6]);
    c->stats.ms_d2h = elapsed(c->ev[7], c->ev[3]);
    *out_bytes = bytes;
    return B200ANS_OK;
}

} // namespace

extern "C" int b200ans_compress(b200ans_ctx* c, int method, uint32_t f, uint8_t* dst, size_t cap, const uint32_t* src,
    size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    return compress_host_impl(c, method, f, dst, cap, src, n, out_bytes, false);
}
B200ANS_CATCH_INT

extern "C" int b200ans_compress_sharded(b200ans_ctx* c, int method, uint32_t f, uint8_t* dst, size_t cap,
    const uint32_t* src, size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    return compress_host_impl(c, method, f, dst, cap, src, n, out_bytes, true);
}
B200ANS_CATCH_INT

namespace {

// Host-buffer decode of a blocked container as a three-stage pipeline over up to kPipeChunks runs of blocks:
//   copy_in : container front (header, prelude, size table), then the payload run by run            (H2D engine)
//   stream  : model tables, size scan, then one decoder launch per run as its bytes arrive          (SMs)
//   copy_out: the decoded symbols of a run, as soon as its launch is done                           (D2H engine)
// PCIe is full duplex, so the (4 n)-byte copy out overlaps the copy in and the decoding instead of following them.
// Header, prelude and block sizes are read from the caller's buffer on the host: no device round trip before
// the first launch.  Blocks are independent given the model, so a run needs only its own payload bytes.
int decompress_pipelined(b200ans_ctx* c, int method, uint32_t f, uint32_t* dst, size_t n, const uint8_t* src, size_t src_bytes)
{
    if (src_bytes < B200ANS_HEADER_BYTES)
        return c->fail(B200ANS_E_FORMAT, "decompress: shorter than a header");
    Header h;
    std::memcpy(&h, src, sizeof h);
    if (h.magic != B200ANS_MAGIC || h.version != 1)
        return c->fail(B200ANS_E_FORMAT, "decompress: bad magic / version");
    if (h.flags & kFlagPseudoAdaptive)
        return c->fail(B200ANS_E_FORMAT, "decompress: pseudo-adaptive container (use b200ans_pa_decompress)");
    if (h.method != method || (h.fidelity != f && (method == METHOD_FOLD || method == METHOD_RFOLD)))
        return c->fail(B200ANS_E_FORMAT, "decompress: container was written by another method");
    if (h.n_symbols != n)
        return c->fail(B200ANS_E_ARG, "decompress: symbol count does not match the container");
    if (n == 0)
        return B200ANS_OK;
    if (h.block_syms == 0 || h.block_syms % 4 != 0 || h.n_blocks != (n + h.block_syms - 1) / h.block_syms)
        return c->fail(B200ANS_E_FORMAT, "decompress: inconsistent block geometry");
    const size_t table_at = B200ANS_HEADER_BYTES + align_up(h.prelude_bytes, 4);
    const size_t payload_at = align_up(table_at + size_t(h.n_blocks) * 4, 16);
    if (payload_at < table_at || payload_at + h.payload_bytes > src_bytes || payload_at + h.payload_bytes < payload_at)
        return c->fail(B200ANS_E_FORMAT, "decompress: container truncated");
    const uint32_t nb = h.n_blocks;

    // runs of blocks with about equal payload; the size table must add up to the header's total
    const int want = int(std::max<uint64_t>(1, std::min<uint64_t>(b200ans_ctx::kPipeChunks, h.payload_bytes / (uint64_t(4) << 20))));
    uint32_t run_end[b200ans_ctx::kPipeChunks];
    uint64_t run_bytes_end[b200ans_ctx::kPipeChunks];
    int runs = 0;
    {
        const uint8_t* tp = src + table_at;
        uint64_t total = 0;
        const uint64_t per = (h.payload_bytes + uint64_t(want) - 1) / uint64_t(want);
        for (uint32_t b = 0; b < nb; b++) {
            uint32_t v;
            std::memcpy(&v, tp + size_t(b) * 4, 4);
            if (v < 32u || v > 0x7FFFFF00u || (coder_of(method) == METHOD_INT && (v & 3u)))
                return c->fail(B200ANS_E_FORMAT, "decompress: a block is shorter than its four states (or, plain ANS, not a whole number of words)");
            total += (uint64_t(v) + 15u) & ~uint64_t(15);
            if (total > h.payload_bytes)
                return c->fail(B200ANS_E_FORMAT, "decompress: block size table does not add up");
            if ((total >= per * uint64_t(runs + 1) && runs < want - 1) || b == nb - 1) {
                run_end[runs] = b + 1;
                run_bytes_end[runs] = total;
                runs++;
            }
        }
        if (total != h.payload_bytes)
            return c->fail(B200ANS_E_FORMAT, "decompress: block size table does not add up");
    }

    RET(ensure_side_streams(c));
    CU(c, c->io_out.reserve(src_bytes + 64));
    CU(c, c->io_in.reserve(std::max<size_t>(n * 4, 16)));
    CU(c, c->offsets.reserve(size_t(nb) * 8));
    uint8_t* d_src = c->io_out.as<uint8_t>();
    uint32_t* d_dst = c->io_in.as<uint32_t>();

    // everything queued on `stream` so far (e.g. the caller's work on a borrowed stream) comes first
    CU(c, cudaEventRecord(c->ev[5], c->stream));
    CU(c, cudaStreamWaitEvent(c->copy_in, c->ev[5], 0));
    CU(c, cudaStreamWaitEvent(c->copy_out, c->ev[5], 0));
    CU(c, cudaMemcpyAsync(d_src, src, payload_at, cudaMemcpyHostToDevice, c->copy_in));
    CU(c, cudaMemsetAsync(d_src + src_bytes, 0, 64, c->copy_in));
    CU(c, cudaEventRecord(c->ev_in[b200ans_ctx::kPipeChunks], c->copy_in));
    for (int r = 0; r < runs; r++) {
        const uint64_t lo = r ? run_bytes_end[r - 1] : 0;
        CU(c, cudaMemcpyAsync(d_src + payload_at + lo, src + payload_at + lo, size_t(run_bytes_end[r] - lo), cudaMemcpyHostToDevice, c->copy_in));
        CU(c, cudaEventRecord(c->ev_in[r], c->copy_in));
    }

    // the model: prelude -> tables (host), uploaded on `stream` while the payload is on its way
    RET(b200ans_model_load(c, method, f, src + B200ANS_HEADER_BYTES, h.prelude_bytes, nullptr));
    CU(c, cudaStreamWaitEvent(c->stream, c->ev_in[b200ans_ctx::kPipeChunks], 0));
    const uint32_t* d_sizes = reinterpret_cast<const uint32_t*>(d_src + table_at);
    scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(d_sizes, nb, c->offsets.as<uint64_t>(), c->d_totals(), nullptr);
    c->stats.launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaMemsetAsync(c->d_err(), 0, 4, c->stream));
    CU(c, cudaEventRecord(c->ev[0], c->stream));
    const bool staged_out = use_staging(dst, n * 4); // pageable destination: the symbols leave through the staging threads
    uint64_t run_out_end[b200ans_ctx::kPipeChunks] {};
    for (int r = 0; r < runs; r++) {
        const uint32_t b0 = r ? run_end[r - 1] : 0, b1 = run_end[r];
        const uint64_t s0 = uint64_t(b0) * h.block_syms, s1 = std::min<uint64_t>(n, uint64_t(b1) * h.block_syms);
        CU(c, cudaStreamWaitEvent(c->stream, c->ev_in[r], 0));
        c->dec_words_aligned = true;
        const int rc = launch_decode(c, d_src + payload_at, c->offsets.as<uint64_t>() + b0, d_sizes + b0, size_t(s1 - s0), h.block_syms, b1 - b0, d_dst + s0);
        c->dec_words_aligned = false;
        RET(rc);
        CU(c, cudaEventRecord(c->ev_dec[r], c->stream));
        run_out_end[r] = s1 * 4;
        if (!staged_out) {
            CU(c, cudaStreamWaitEvent(c->copy_out, c->ev_dec[r], 0));
            CU(c, cudaMemcpyAsync(dst + s0, d_dst + s0, size_t(s1 - s0) * 4, cudaMemcpyDeviceToHost, c->copy_out));
        }
    }
    if (staged_out // a chunk of the output leaves once the run that decodes its last byte is done
        && staged_d2h(c, dst, d_dst, n * 4, c->copy_out, [&](size_t end) {
               int r = 0;
               while (r < runs - 1 && run_out_end[r] < end)
                   r++;
               return c->ev_dec[r];
           }) != B200ANS_OK) {
        CU(c, cudaStreamWaitEvent(c->copy_out, c->ev_dec[runs - 1], 0)); // (redone as one plain copy behind the last run)
        CU(c, cudaMemcpyAsync(dst, d_dst, n * 4, cudaMemcpyDeviceToHost, c->copy_out));
    }
    CU(c, cudaEventRecord(c->ev[1], c->stream));
    CU(c, cudaEventRecord(c->ev[6], c->copy_out));
    CU(c, cudaStreamWaitEvent(c->stream, c->ev[6], 0)); // the call's work ends on `stream`, as every other entry point's
    CU(c, cudaEventRecord(c->ev[7], c->stream));
    RET(read_small(c)); // the call's one synchronisation: error flag + everything above
    c->stats.ms_decode = elapsed(c->ev[0], c->ev[1]);
    c->stats.ms_h2d = 0.f; // overlapped: the stages are not separable
    c->stats.ms_d2h = elapsed(c->ev[1], c->ev[7]);
    c->stats.n_blocks = nb;
    if (c->h_small()->error_flag)
        return c->fail(B200ANS_E_FORMAT, "decompress: a block's payload was not consumed exactly (corrupt stream)");
    return B200ANS_OK;
}

} // namespace

extern "C" int b200ans_decompress(b200ans_ctx* c, int method, uint32_t f, uint32_t* dst, size_t n, const uint8_t* src,
    size_t src_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!dst && n) || !src || !valid_method(method, f))
        return c->fail(B200ANS_E_ARG, "decompress: bad argument");
    CU(c, cudaSetDevice(c->device));
    if (c->block_syms != 0 && !env_int("B200ANS_NO_PIPELINE", 0)) {
        c->stats.ms_decode = c->stats.ms_model = 0.f;
        const int rc = decompress_pipelined(c, method, f, dst, n, src, src_bytes);
        if (rc != B200ANS_OK && c->copy_in) { // nothing of a failed call may still be reading or writing the caller's buffers
            cudaStreamSynchronize(c->copy_in);
            cudaStreamSynchronize(c->stream);
            cudaStreamSynchronize(c->copy_out);
            cudaGetLastError();
        }
        return rc;
    }
    CU(c, c->io_out.reserve(src_bytes + 64));
    CU(c, c->io_in.reserve(std::max<size_t>(n * 4, 16)));
    CU(c, cudaEventRecord(c->ev[5], c->stream));
    CU(c, cudaMemcpyAsync(c->io_out.p, src, src_bytes, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaMemsetAsync(c->io_out.as<uint8_t>() + src_bytes, 0, 64, c->stream));
    CU(c, cudaEventRecord(c->ev[6], c->stream));
    RET(b200ans_decompress_device(c, method, f, c->io_in.as<uint32_t>(), n, c->io_out.as<uint8_t>(), src_bytes));
    CU(c, cudaEventRecord(c->ev[7], c->stream));
    if (n)
        CU(c, cudaMemcpyAsync(dst, c->io_in.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaEventRecord(c->ev[3], c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->stats.ms_h2d = elapsed(c->ev[5], c->ev[6]);
    c->stats.ms_d2h = elapsed(c->ev[7], c->ev[3]);
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// ==========================================================================================
// pseudo-adaptive mode: one semi-static model per block of `model_syms` symbols
// (src/pseudo_adaptive.cpp:64-177 -- every block goes through T::encode on its own; no shared histogram)
// ==========================================================================================
// Container: [64-byte header, flags bit 0 set, block_syms = model_syms, n_blocks = models]
//            [u32 prelude bytes of model m][u32 payload bytes of model m][pad to 16]
//            [preludes, each 16-byte aligned][payloads, each 16-byte aligned]
// prelude(m) followed by payload(m) is exactly what the reference's ans_X_compress() writes for block m.
// Everything per block runs on the device: K1 (pa_histogram_kernel) -> adjust_freqs restated (pa_sort_kernel + pa_fit_kernel; the
// rare model whose accept/reject margin is within 1e-9 is redone by the host's adjust_freqs) -> prelude
// (pa_prelude_encode_kernel) -> encoder tables (pa_enc_tables_kernel) -> one reference-format stream per block
// (encode_kernel<.., PA>) -> size scans + compaction of preludes and payloads.  Decode mirrors it
// (pa_prelude_decode_kernel, pa_dec_tables_kernel, decode_kernel<.., PA>).  The host only lays the container out.
namespace {

constexpr uint32_t kPaIntMaxModelSyms = 16256; // bins 0 .. model_syms in 16 384 (prelude decoder depth, 16-bit symbols in the sort keys)
// bins per model: the folded alphabet, or 0 .. model_syms for plain ANS (rounded up to a multiple of 16)
uint32_t pa_bins(int method, uint32_t f, uint32_t model_syms)
{
    return method == METHOD_INT ? uint32_t(align_up(size_t(model_syms) + 1, 16)) : uint32_t(alphabet_bins(method, f, 0));
}

int pa_check(b200ans_ctx* c, int method, uint32_t f, uint32_t model_syms)
{
    if (!(method == METHOD_MSB || method == METHOD_INT || (method == METHOD_FOLD && f >= 1 && f <= 2)))
        return c->fail(B200ANS_E_ARG, "pseudo-adaptive mode: ANSfold-1, ANSfold-2, ANSmsb, and plain ANS over dense ranks");
    if (model_syms < 1024 || model_syms % 128 != 0 || model_syms > (1u << 24))
        return c->fail(B200ANS_E_ARG, "pseudo-adaptive mode: symbols per model must be a multiple of 128 in [1024, 2^24]");
    // plain ANS: the alphabet of a block is 0 .. model_syms (ranks behind the driver's remap, pseudo_adaptive.cpp:83-99)
    if (method == METHOD_INT && model_syms > kPaIntMaxModelSyms)
        return c->fail(B200ANS_E_ARG, "pseudo-adaptive plain ANS: at most 16256 symbols per model");
    return B200ANS_OK;
}

// Launch shape of the shared-memory-table coders (pa_kernels.cuh): one CTA per SM, as many warps (8 models each) as
// fit next to `fixed` bytes, spread in balanced waves: tasks / (SMs * waves) warps per CTA, rounded up.
struct PaLaunch {
    uint32_t max_warps, warps, grid;
    size_t smem;
};

PaLaunch pa_launch_shape(const b200ans_ctx* c, size_t fixed, size_t warp_bytes, uint32_t n_tasks)
{
    PaLaunch p {};
    const size_t room = c->max_smem > fixed + 1024 ? c->max_smem - fixed - 1024 : 0;
    p.max_warps = uint32_t(std::min<size_t>(size_t(kPaCodeMaxWarps), room / warp_bytes));
    const int cap = env_int("B200ANS_PA_WARPS", 0);
    if (cap > 0)
        p.max_warps = std::min<uint32_t>(p.max_warps, uint32_t(cap));
    if (p.max_warps == 0)
        return p;
    const uint32_t sms = uint32_t(c->num_sms);
    const uint32_t waves = (n_tasks + sms * p.max_warps - 1) / (sms * p.max_warps);
    p.warps = std::max<uint32_t>(1, std::min<uint32_t>(p.max_warps, (n_tasks + sms * waves - 1) / (sms * waves)));
    if (n_tasks >= 2u * sms)
        p.warps = std::max<uint32_t>(p.warps, std::min<uint32_t>(p.max_warps, 2u));
    p.grid = std::max<uint32_t>(1, std::min<uint32_t>(sms, (n_tasks + p.warps - 1) / p.warps));
    p.smem = fixed + size_t(p.warps) * warp_bytes;
    return p;
}

// pinned staging that grows on demand
int pa_pinned(b200ans_ctx* c, size_t bytes)
{
    if (bytes <= c->pa_pinned_cap)
        return B200ANS_OK;
    if (c->pa_pinned)
        cudaFreeHost(c->pa_pinned);
    c->pa_pinned = nullptr;
    c->pa_pinned_cap = 0;
    CU(c, cudaHostAlloc(reinterpret_cast<void**>(&c->pa_pinned), bytes + bytes / 8, cudaHostAllocDefault));
    c->pa_pinned_cap = bytes + bytes / 8;
    return B200ANS_OK;
}

} // namespace

extern "C" size_t b200ans_pa_compress_bound(const b200ans_ctx* c, int method, uint32_t model_syms, size_t n)
{
    (void)c;
    if (model_syms == 0)
        return 0;
    const size_t models = (n + model_syms - 1) / model_syms;
    // per model: prelude (<= 8 bytes per bin + header), sizes, padding; per symbol at most 7 bytes + states
    const size_t per_model = (method == METHOD_INT ? size_t(model_syms) + 16 : size_t(2048)) * 8 + 128 + 8 + 32 + 64;
    return B200ANS_HEADER_BYTES + models * per_model + n * (method == METHOD_INT ? 4 : 7) + 64;
}

extern "C" int b200ans_pa_compress_device(b200ans_ctx* c, int method, uint32_t f, uint32_t model_syms, uint8_t* d_dst,
    size_t cap, const uint32_t* d_src, size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_dst || !out_bytes || (!d_src && n) || (reinterpret_cast<uintptr_t>(d_src) & 15) || (reinterpret_cast<uintptr_t>(d_dst) & 15))
        return c->fail(B200ANS_E_ARG, "pa_compress: bad argument (device buffers must be 16-byte aligned)");
    RET(pa_check(c, method, f, model_syms));
    CU(c, cudaSetDevice(c->device));
    c->stats.ms_histogram = c->stats.ms_encode = c->stats.ms_compact = c->stats.ms_model = 0.f;
    const uint64_t nm64 = (n + model_syms - 1) / model_syms;
    if (nm64 > (uint64_t(1) << 24))
        return c->fail(B200ANS_E_ARG, "pa_compress: more than 2^24 models");
    const uint32_t nm = uint32_t(nm64);
    Header h {};
    h.magic = B200ANS_MAGIC;
    h.version = 1;
    h.method = uint8_t(method);
    h.fidelity = uint8_t(f);
    h.n_symbols = n;
    h.block_syms = model_syms;
    h.n_blocks = nm;
    h.flags = kFlagPseudoAdaptive;
    if (n == 0) {
        if (cap < B200ANS_HEADER_BYTES)
            return c->fail(B200ANS_E_CAPACITY, "pa_compress: destination too small");
        CU(c, cudaMemcpyAsync(d_dst, &h, sizeof h, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        *out_bytes = B200ANS_HEADER_BYTES;
        return B200ANS_OK;
    }
    const uint32_t nbins = pa_bins(method, f, model_syms);
    const size_t bins_bytes = size_t(nm) * nbins * 4;
    const size_t sizes_at = B200ANS_HEADER_BYTES;
    const size_t pre_at = align_up(sizes_at + size_t(nm) * 8, 16);
    if (cap < pre_at)
        return c->fail(B200ANS_E_CAPACITY, "pa_compress: destination too small");
    const int mgrid = int(std::min<uint32_t>(nm, uint32_t(c->num_sms) * 8));

    // K1 per model block
    CU(c, cudaEventRecord(c->ev[3], c->stream));
    CU(c, c->bins.reserve(bins_bytes));
    CU(c, cudaMemsetAsync(c->small.p, 0, 64, c->stream));
    {
        const size_t smem = size_t(4) * (nbins | 1u) * 4;
        auto hk = method == METHOD_MSB ? pa_histogram_kernel<METHOD_MSB> : method == METHOD_INT ? pa_histogram_kernel<METHOD_INT> : pa_histogram_kernel<METHOD_FOLD>;
        CU(c, cudaFuncSetAttribute(hk, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        hk<<<mgrid, kPaThreads, smem, c->stream>>>(d_src, n, model_syms, nm, f, nbins, c->bins.as<uint32_t>(), c->d_max());
        c->stats.launches++;
        CU(c, cudaGetLastError());
    }
    CU(c, cudaEventRecord(c->ev[4], c->stream));

    // device normalisation (pa_sort_kernel + pa_fit_kernel): counts -> frequencies, log2 M and a status byte per model
    CU(c, c->pa_nfreqs.reserve(bins_bytes + 2 * size_t(nm)));
    uint8_t* d_lg = c->pa_nfreqs.as<uint8_t>() + bins_bytes;
    uint8_t* d_status = d_lg + nm;
    {
        uint32_t n_keys = 1;
        while (n_keys < nbins)
            n_keys <<= 1;
        const size_t smem = size_t(n_keys) * 8;
        // as many CTAs (one model at a time each) as the SMs hold: 64 threads + the keys in shared memory
        const int grid = int(std::min<uint32_t>(nm, uint32_t(c->num_sms) * uint32_t(std::max<size_t>(1, std::min<size_t>(32, (c->max_smem_sm - 4096) / (smem + 2048))))));
        if (!c->pa_log2_ready) { // once per context (kernels.cuh pa_log2_table_kernel)
            CU(c, c->pa_log2tab.reserve(size_t(kPaLog2Tab) * 8));
            pa_log2_table_kernel<<<kPaLog2Tab / 256, 256, 0, c->stream>>>(c->pa_log2tab.as<double>());
            c->stats.launches++;
            c->pa_log2_ready = true;
        }
        // B200ANS_PA_MARGIN (testing): a huge margin sends every model to the host's adjust_freqs
        const char* mv = std::getenv("B200ANS_PA_MARGIN");
        const double margin = mv && *mv ? std::atof(mv) : kPaMargin;
        const size_t keys_bytes = size_t(nm) * n_keys * 8;
        CU(c, c->pa_keys.reserve(keys_bytes + size_t(nm) * sizeof(PaMeta)));
        uint64_t* d_keys = c->pa_keys.as<uint64_t>();
        PaMeta* d_meta = reinterpret_cast<PaMeta*>(c->pa_keys.as<uint8_t>() + keys_bytes);
        // counting sort (one warp per model) whenever the symbols with large counts fit its side list; B200ANS_PA_BITONIC=1
        // (testing) keeps the bitonic sort of the first version
        if (uint64_t(model_syms) <= uint64_t(kPaSortSmall) * kPaSortBig && !env_int("B200ANS_PA_BITONIC", 0)) {
            const size_t csmem = pa_sort_smem_bytes();
            CU(c, cudaFuncSetAttribute(pa_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(csmem)));
            const uint32_t per_sm = uint32_t(std::max<size_t>(1, (c->max_smem_sm - 4096) / (csmem + 1024)));
            const uint32_t cgrid = std::min<uint32_t>((nm + kPaSortWarps - 1) / kPaSortWarps, uint32_t(c->num_sms) * per_sm);
            pa_sort_kernel<<<cgrid, kPaSortWarps * 32, csmem, c->stream>>>(c->bins.as<uint32_t>(), nbins, n_keys, nm, c->h_approx, d_keys, d_meta,
                c->pa_nfreqs.as<uint32_t>(), c->pa_log2tab.as<double>());
        } else {
            pa_sort_bitonic_kernel<<<grid, kPaNormThreads, smem, c->stream>>>(c->bins.as<uint32_t>(), nbins, nm, c->h_approx, d_keys, d_meta,
                c->pa_nfreqs.as<uint32_t>());
        }
        const uint32_t fit_threads = nm * uint32_t(kPaFitGroup);
        pa_fit_kernel<<<(fit_threads + kPaFitThreads - 1) / kPaFitThreads, kPaFitThreads, 0, c->stream>>>(d_keys, n_keys, d_meta, nm, margin, d_lg,
            d_status, c->d_max_log2(), c->pa_log2tab.as<double>(), method != METHOD_INT);
        pa_fit_store_kernel<<<(nm + 31) / 32, 32, 0, c->stream>>>(d_keys, n_keys, d_meta, nbins, nm, d_lg, d_status, c->pa_nfreqs.as<uint32_t>());
        c->stats.launches++;
        c->stats.launches += 2;
        CU(c, cudaGetLastError());
    }
    CU(c, cudaEventRecord(c->ev[5], c->stream));
    RET(pa_pinned(c, size_t(nm) + 64));
    CU(c, cudaMemcpyAsync(c->pa_pinned, d_status, nm, cudaMemcpyDeviceToHost, c->stream));
    RET(read_small(c));
    c->stats.ms_histogram = elapsed(c->ev[3], c->ev[4]);
    const float ms_normalise = elapsed(c->ev[4], c->ev[5]);
    if (method == METHOD_INT && c->h_small()->max_value > model_syms)
        return c->fail(B200ANS_E_DOMAIN, "pseudo-adaptive plain ANS codes dense ranks: values up to the symbols per model");
    if (c->h_small()->max_value >= (1u << 30))
        return c->fail(B200ANS_E_DOMAIN, "folded coders need values < 2^30");

    // host: only the models the device left undecided (decision margin below kPaMargin) go through adjust_freqs here
    const double t0 = now_ms();
    const uint8_t* status = c->pa_pinned;
    uint32_t undecided = 0;
    uint32_t max_lg = c->h_small()->max_log2;
    for (uint32_t m = 0; m < nm; m++) {
        if (status[m] == 1)
            return c->fail(B200ANS_E_DOMAIN, "pa_compress: normalisation produced an unusable frame size");
        if (status[m] != 2)
            continue;
        undecided++;
        std::vector<uint32_t> cnt(nbins);
        CU(c, cudaMemcpyAsync(cnt.data(), c->bins.as<uint32_t>() + size_t(m) * nbins, size_t(nbins) * 4, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        std::vector<uint64_t> freqs(cnt.begin(), cnt.end());
        uint32_t max_sym = 0;
        for (uint32_t b = 0; b < nbins; b++)
            if (cnt[b])
                max_sym = b;
        freqs.resize(size_t(max_sym) + 1);
        // plain ANS has no 16-bit guard (ans_int.hpp:50) -- and without it adjust_freqs never returns for a block of ONE
        // distinct symbol (entropy 0: no frame beats the limit).  The driver never codes such a block (pseudo_adaptive.cpp:114);
        // here it gets the model the guarded normaliser gives it.
        uint32_t distinct = 0;
        for (uint32_t b = 0; b <= max_sym; b++)
            distinct += cnt[b] != 0;
        const std::vector<uint32_t> nf = adjust_freqs(freqs, max_sym, method != METHOD_INT || distinct < 2, c->h_approx);
        uint64_t frame = 0;
        for (uint32_t v : nf)
            frame += v;
        if (!is_power_of_two(frame) || frame > (uint64_t(1) << 20))
            return c->fail(B200ANS_E_DOMAIN, "pa_compress: normalisation produced an unusable frame size");
        std::fill(cnt.begin(), cnt.end(), 0u);
        std::copy(nf.begin(), nf.end(), cnt.begin());
        const uint8_t lgm = uint8_t(63 - __builtin_clzll(frame));
        max_lg = std::max<uint32_t>(max_lg, lgm);
        CU(c, cudaMemcpyAsync(c->pa_nfreqs.as<uint32_t>() + size_t(m) * nbins, cnt.data(), size_t(nbins) * 4, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpyAsync(d_lg + m, &lgm, 1, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
    }
    c->pa_undecided = undecided;
    if (method == METHOD_INT && max_lg > 15u)
        return c->fail(B200ANS_E_DOMAIN, "pseudo-adaptive plain ANS: a block's frame exceeds 2^15 slots (16-bit tables)");
    const float ms_host = float(now_ms() - t0);

    // Preludes on the device (the counts are no longer needed: their buffer is the coder's scratch), then their sizes
    // are scanned.  The prelude coder (one warp per model) needs only the frequencies, as the encoder does: it runs
    // NEXT TO the encoder on a side stream, and its total comes back while the encoder runs.
    // table entries per model: beyond the largest symbol of the input, a multiple of 16
    uint32_t pa_syms;
    {
        uint32_t kk;
        const uint32_t top_sym = method == METHOD_INT ? c->h_small()->max_value
            : method == METHOD_MSB                    ? msb_map(c->h_small()->max_value, kk)
                                                      : fold_map(c->h_small()->max_value, f, kk);
        pa_syms = std::min<uint32_t>(nbins, (top_sym + 1u + 15u) & ~15u);
    }
    RET(ensure_side_streams(c));
    CU(c, cudaEventRecord(c->ev[6], c->stream));
    CU(c, cudaStreamWaitEvent(c->copy_in, c->ev[6], 0));
    const uint32_t pre_stride = uint32_t(align_up(size_t(nbins) * 4 + 64, 16));
    CU(c, c->pa_pre.reserve(size_t(nm) * pre_stride + size_t(nm) * 12));
    uint8_t* d_pre = c->pa_pre.as<uint8_t>();
    uint64_t* d_pre_off = reinterpret_cast<uint64_t*>(d_pre + size_t(nm) * pre_stride);
    uint32_t* d_pre_bytes = reinterpret_cast<uint32_t*>(d_pre_off + nm);
    {
        // (the arrays in shared memory stop behind the largest symbol of the input: half of ANSfold-1's 1024 bins on C4)
        uint32_t pw = kPaPreEncWarps; // warps per CTA: as many as fit (plain ANS: up to 16 384 bins per model)
        while (pw > 1 && pa_prelude_enc_smem(pa_syms, pw) + 2048 > c->max_smem)
            pw >>= 1;
        const size_t smem = pa_prelude_enc_smem(pa_syms, pw);
        CU(c, cudaFuncSetAttribute(pa_prelude_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
        const uint32_t per_sm = uint32_t(std::max<size_t>(1, std::min<size_t>(8, (c->max_smem_sm - 4096) / (smem + 1024))));
        const uint32_t grid = std::min<uint32_t>((nm + pw - 1) / pw, uint32_t(c->num_sms) * per_sm);
        pa_prelude_encode_kernel<<<grid, pw * 32, smem, c->copy_in>>>(c->pa_nfreqs.as<uint32_t>(), nbins, pa_syms, nm, d_lg, d_pre, pre_stride, d_pre_bytes);
    }
    scan_block_sizes_kernel<<<1, 1024, 0, c->copy_in>>>(d_pre_bytes, nm, d_pre_off, c->d_totals_b(), nullptr);
    c->stats.launches += 2;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(c->h_small(), c->small.p, sizeof(b200ans_ctx::SmallHost), cudaMemcpyDeviceToHost, c->copy_in));
    CU(c, cudaEventRecord(c->ev[7], c->copy_in));

    // frequencies -> one reference-format stream per model block.  Tables: built per block in shared memory by the
    // coder itself (pa_kernels.cuh) when every frame is at most 2^16, else one 16-byte entry per symbol and model in
    // global memory (pa_enc_tables_kernel + encode_kernel<.., PA>, the first generation; B200ANS_PATH_GROUP forces it)
    const size_t stride = block_stride(method, model_syms);
    CU(c, c->scratch.reserve(stride * nm));
    CU(c, c->block_bytes.reserve(size_t(nm) * 4));
    CU(c, c->offsets.reserve(size_t(nm) * 8));
    c->have_enc = false; // the single-model encoder table is gone
    EncArgs a {};
    a.in = d_src;
    a.n = n;
    a.block_syms = model_syms;
    a.n_blocks = nm;
    a.n_entries = nbins;
    a.fidelity = f;
    a.scratch = c->scratch.as<uint8_t>();
    a.stride = stride;
    a.block_bytes = c->block_bytes.as<uint32_t>();
    a.error_flag = c->d_err();
    a.pa_log2 = d_lg;
    a.pa_table_stride = nbins;
    a.pa_nfreqs = c->pa_nfreqs.as<uint32_t>();
    a.pa_nbins = nbins;
    a.pa_syms = pa_syms;
    const uint32_t n_tasks = (nm + 7) / 8;
    PaLaunch pl = pa_launch_shape(c, size_t(kPaRcpN) * 8, size_t(8) * pa_enc_group_bytes(a.pa_syms), n_tasks);
    const bool smem_tables = c->kernel_path != B200ANS_PATH_GROUP && max_lg <= 16u && pl.warps >= 2u;
    CU(c, cudaEventRecord(c->ev[0], c->stream));
    // first choice: a lane per block (the L3 encoder with per-lane models, lpb3_kernels.cuh)
    int rc3 = 1;
    if (c->kernel_path != B200ANS_PATH_GROUP)
        rc3 = method == METHOD_MSB ? launch_pa3_encode<METHOD_MSB>(c, a, max_lg)
            : method == METHOD_INT ? launch_pa3_encode<METHOD_INT>(c, a, max_lg)
                                   : launch_pa3_encode<METHOD_FOLD>(c, a, max_lg);
    if (rc3 < 0)
        return rc3;
    if (rc3 != B200ANS_OK && !smem_tables) {
        CU(c, c->enc_table.reserve(size_t(nm) * nbins * sizeof(EncEntry)));
        a.table = c->enc_table.as<EncEntry>();
        pa_enc_tables_kernel<<<mgrid, kPaThreads, (size_t(nbins) + 8) * 4, c->stream>>>(c->pa_nfreqs.as<uint32_t>(), nbins, nm, c->enc_table.as<EncEntry>());
        c->stats.launches++;
        CU(c, cudaGetLastError());
        CU(c, cudaEventRecord(c->ev[0], c->stream));
    }
    if (rc3 == B200ANS_OK) {
    } else if (smem_tables) {
        auto k = method == METHOD_MSB ? pa_smem_encode_kernel<METHOD_MSB>
            : method == METHOD_INT    ? pa_smem_encode_kernel<METHOD_INT>
                                      : pa_smem_encode_kernel<METHOD_FOLD>;
        CU(c, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(pl.smem)));
        k<<<pl.grid, pl.warps * 32, pl.smem, c->stream>>>(a);
        set_name(c->stats.encode_kernel, "pa_smem_encode_kernel");
        c->stats.launches++;
        CU(c, cudaGetLastError());
    } else {
        set_name(c->stats.encode_kernel, "encode_kernel");
        const size_t rows = size_t(kEncWarps) * 8 * kEncRowBytes;
        if (method == METHOD_MSB) {
            auto k = encode_kernel<METHOD_MSB, false, true>;
            k<<<persistent_grid(c, k, kEncWarps * 32, rows, n_tasks, kEncWarps), kEncWarps * 32, rows, c->stream>>>(a);
        } else if (method == METHOD_INT) {
            auto k = encode_kernel<METHOD_INT, false, true>;
            k<<<persistent_grid(c, k, kEncWarps * 32, rows, n_tasks, kEncWarps), kEncWarps * 32, rows, c->stream>>>(a);
        } else {
            auto k = encode_kernel<METHOD_FOLD, false, true>;
            k<<<persistent_grid(c, k, kEncWarps * 32, rows, n_tasks, kEncWarps), kEncWarps * 32, rows, c->stream>>>(a);
        }
        c->stats.launches++;
        CU(c, cudaGetLastError());
    }
    CU(c, cudaEventRecord(c->ev[1], c->stream));

    // the prelude total is on the host by now (the encoder is still running): lay the container out
    CU(c, cudaEventSynchronize(c->ev[7]));
    CU(c, cudaStreamWaitEvent(c->stream, c->ev[7], 0)); // the compaction below reads what the side stream wrote
    const size_t pre_region = c->h_small()->totals_b[0];
    const size_t payload_at = pre_at + pre_region;
    if (pre_region > 0xFFFFFFF0ull)
        return c->fail(B200ANS_E_ARG, "pa_compress: preludes exceed 4 GiB");
    if (cap < payload_at)
        return c->fail(B200ANS_E_CAPACITY, "pa_compress: destination too small");
    h.prelude_bytes = uint32_t(pre_region);
    std::memcpy(c->pinned, &h, sizeof h); // pinned: the copy below is asynchronous
    CU(c, cudaMemcpyAsync(d_dst, c->pinned, sizeof h, cudaMemcpyHostToDevice, c->stream));
    const int cgrid = int(std::max<uint32_t>(1, std::min<uint32_t>((nm + 7) / 8, uint32_t(c->num_sms) * 8)));
    compact_kernel<<<cgrid, 256, 0, c->stream>>>(d_pre, pre_stride, d_pre_bytes, d_pre_off, nm, d_dst + pre_at,
        reinterpret_cast<uint32_t*>(d_dst + sizes_at), uint64_t(pre_region));
    scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(c->block_bytes.as<uint32_t>(), nm, c->offsets.as<uint64_t>(), c->d_totals(),
        reinterpret_cast<uint64_t*>(d_dst + offsetof(Header, payload_bytes)));
    compact_kernel<<<cgrid, 256, 0, c->stream>>>(c->scratch.as<uint8_t>(), stride, c->block_bytes.as<uint32_t>(), c->offsets.as<uint64_t>(),
        nm, d_dst + payload_at, reinterpret_cast<uint32_t*>(d_dst + sizes_at + size_t(nm) * 4), uint64_t(cap - payload_at));
    c->stats.launches += 3;
    CU(c, cudaGetLastError());
    CU(c, cudaEventRecord(c->ev[2], c->stream));
    RET(read_small(c));
    if ((c->h_small()->error_flag & 2u) && !c->pa3_off) {
        // a model the lane-per-block encoder does not take (two different frequencies beyond its reciprocal table): the
        // whole input again through the four-lanes-per-block encoder
        struct Once {
            bool& flag;
            explicit Once(bool& f) : flag(f) { flag = true; }
            ~Once() { flag = false; }
        } once(c->pa3_off);
        return b200ans_pa_compress_device(c, method, f, model_syms, d_dst, cap, d_src, n, out_bytes);
    }
    if (c->h_small()->error_flag)
        return c->fail(B200ANS_E_DOMAIN, "pa_compress: a symbol had no frequency in its block's model");
    const size_t total = payload_at + c->h_small()->totals[0];
    if (cap < total)
        return c->fail(B200ANS_E_CAPACITY, "pa_compress: destination too small");
    *out_bytes = total;
    c->stats.ms_model = ms_normalise + ms_host + elapsed(c->ev[6], c->ev[7]);
    c->stats.ms_encode = elapsed(c->ev[0], c->ev[1]);
    c->stats.ms_compact = elapsed(c->ev[1], c->ev[2]);
    c->stats.n_blocks = nm;
    c->stats.prelude_bytes = pre_region;
    c->stats.payload_bytes = c->h_small()->totals[1];
    c->stats.total_bytes = total;
    if (std::getenv("B200ANS_PA_TRACE"))
        std::fprintf(stderr, "pa_compress: %u models, normalise %.3f ms, host %.3f ms (%u undecided), preludes %.3f ms\n", nm, ms_normalise,
            ms_host, undecided, elapsed(c->ev[6], c->ev[7]));
    return B200ANS_OK;
}
B200ANS_CATCH_INT

namespace {

// Decoder of a pseudo-adaptive container at d_src (device).  With h_src / h_dst set, the container still sits in host
// memory and the symbols are wanted there: header and size tables are read from h_src, the payload travels in runs
// of models on `copy_in` behind the front part (sizes + preludes), every run is decoded as soon as it has arrived and
// its symbols leave on `copy_out` while the next run is decoded (the same three-stream pipeline as
// decompress_pipelined).  Without them the container is already in HBM and nothing is copied.
int pa_decompress_core(b200ans_ctx* c, int method, uint32_t f, uint32_t* d_dst, size_t n, const uint8_t* d_src,
    size_t src_bytes, const uint8_t* h_src, uint32_t* h_dst)
{
    c->stats.ms_decode = c->stats.ms_model = 0.f;
    if (src_bytes < B200ANS_HEADER_BYTES)
        return c->fail(B200ANS_E_FORMAT, "pa_decompress: shorter than a header");
    Header h;
    if (h_src) {
        std::memcpy(&h, h_src, sizeof h);
    } else {
        CU(c, cudaMemcpyAsync(&h, d_src, sizeof h, cudaMemcpyDeviceToHost, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
    }
    if (h.magic != B200ANS_MAGIC || h.version != 1 || !(h.flags & kFlagPseudoAdaptive))
        return c->fail(B200ANS_E_FORMAT, "pa_decompress: not a pseudo-adaptive b200ans container");
    if (h.method != method || h.fidelity != f)
        return c->fail(B200ANS_E_FORMAT, "pa_decompress: container was written by another method");
    if (h.n_symbols != n)
        return c->fail(B200ANS_E_ARG, "pa_decompress: symbol count does not match the container");
    if (n == 0)
        return B200ANS_OK;
    RET(pa_check(c, method, f, h.block_syms));
    const uint32_t nm = h.n_blocks;
    if (nm != (n + h.block_syms - 1) / h.block_syms)
        return c->fail(B200ANS_E_FORMAT, "pa_decompress: inconsistent block geometry");
    const uint32_t nbins = pa_bins(method, f, h.block_syms);
    const size_t bins_bytes = size_t(nm) * nbins * 4;
    const size_t sizes_at = B200ANS_HEADER_BYTES;
    const size_t pre_at = align_up(sizes_at + size_t(nm) * 8, 16);
    const size_t payload_at = pre_at + h.prelude_bytes;
    if (payload_at + h.payload_bytes > src_bytes || payload_at < pre_at)
        return c->fail(B200ANS_E_FORMAT, "pa_decompress: container truncated");

    // host buffers: runs of models with about equal payload, queued on copy_in behind the front part
    constexpr int kMaxRuns = b200ans_ctx::kPipeChunks;
    uint32_t run_end[kMaxRuns];
    uint64_t run_bytes_end[kMaxRuns];
    int runs = 0;
    if (h_src) {
        const int cap = std::max(1, std::min(kMaxRuns, env_int("B200ANS_PA_RUNS", kMaxRuns)));
        const int want = int(std::max<uint64_t>(1, std::min<uint64_t>(uint64_t(cap), h.payload_bytes / (uint64_t(8) << 20))));
        const uint8_t* tp = h_src + sizes_at + size_t(nm) * 4;
        uint64_t total = 0;
        const uint64_t per = (h.payload_bytes + uint64_t(want) - 1) / uint64_t(want);
        for (uint32_t b = 0; b < nm; b++) {
            uint32_t v;
            std::memcpy(&v, tp + size_t(b) * 4, 4);
            total += (uint64_t(v) + 15u) & ~uint64_t(15);
            if (total > h.payload_bytes)
                return c->fail(B200ANS_E_FORMAT, "pa_decompress: block size table does not add up");
            // runs end on a multiple of 8 models (one warp's task) so that no task straddles two runs
            if ((total >= per * uint64_t(runs + 1) && runs < want - 1 && (b + 1) % 8 == 0) || b == nm - 1) {
                run_end[runs] = b + 1;
                run_bytes_end[runs] = total;
                runs++;
            }
        }
        RET(ensure_side_streams(c));
        CU(c, cudaEventRecord(c->ev[5], c->stream));
        CU(c, cudaStreamWaitEvent(c->copy_in, c->ev[5], 0));
        CU(c, cudaStreamWaitEvent(c->copy_out, c->ev[5], 0));
        uint8_t* d_in = const_cast<uint8_t*>(d_src);
        CU(c, cudaMemcpyAsync(d_in, h_src, payload_at, cudaMemcpyHostToDevice, c->copy_in));
        CU(c, cudaMemsetAsync(d_in + src_bytes, 0, 64, c->copy_in));
        CU(c, cudaEventRecord(c->ev_in[kMaxRuns], c->copy_in));
        for (int r = 0; r < runs; r++) {
            const uint64_t lo = r ? run_bytes_end[r - 1] : 0;
            if (run_bytes_end[r] > lo)
                CU(c, cudaMemcpyAsync(d_in + payload_at + lo, h_src + payload_at + lo, size_t(run_bytes_end[r] - lo), cudaMemcpyHostToDevice, c->copy_in));
            CU(c, cudaEventRecord(c->ev_in[r], c->copy_in));
        }
        CU(c, cudaStreamWaitEvent(c->stream, c->ev_in[kMaxRuns], 0));
    }

    // device: size scans, preludes -> frequencies (pa_prelude_decode_kernel), one synchronisation for the verdicts
    CU(c, cudaEventRecord(c->ev[3], c->stream));
    const uint32_t* d_pre_bytes = reinterpret_cast<const uint32_t*>(d_src + sizes_at);
    const uint32_t* d_pay_bytes = d_pre_bytes + nm;
    CU(c, c->pa_pre.reserve(size_t(nm) * 8));
    CU(c, c->offsets.reserve(size_t(nm) * 8));
    CU(c, c->bins.reserve(bins_bytes));
    CU(c, c->pa_nfreqs.reserve(bins_bytes + 2 * size_t(nm)));
    uint8_t* d_lg = c->pa_nfreqs.as<uint8_t>() + bins_bytes;
    uint64_t* d_pre_off = c->pa_pre.as<uint64_t>();
    CU(c, cudaMemsetAsync(c->small.p, 0, 64, c->stream));
    scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(d_pre_bytes, nm, d_pre_off, c->d_totals_b(), nullptr);
    scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(d_pay_bytes, nm, c->offsets.as<uint64_t>(), c->d_totals(), nullptr);
    pa_prelude_decode_kernel<<<(nm + kPaPreThreads - 1) / kPaPreThreads, kPaPreThreads, 0, c->stream>>>(d_src + pre_at, d_pre_off, d_pre_bytes,
        uint64_t(h.prelude_bytes), nbins, nm, c->bins.as<uint32_t>(), c->pa_nfreqs.as<uint32_t>(), d_lg);
    pa_prelude_finish_kernel<<<(nm + kPaPreFinishWarps - 1) / kPaPreFinishWarps, kPaPreFinishWarps * 32, 0, c->stream>>>(c->bins.as<uint32_t>(), nbins, nm,
        c->pa_nfreqs.as<uint32_t>(), d_lg, c->d_err(), c->d_max_log2(), c->d_max_symbol());
    c->stats.launches += 4;
    CU(c, cudaGetLastError());
    CU(c, cudaEventRecord(c->ev[4], c->stream));
    RET(read_small(c));
    if (c->h_small()->totals_b[0] != h.prelude_bytes)
        return c->fail(B200ANS_E_FORMAT, "pa_decompress: prelude size table does not add up");
    if (c->h_small()->totals[0] != h.payload_bytes)
        return c->fail(B200ANS_E_FORMAT, "pa_decompress: block size table does not add up");
    if (c->h_small()->error_flag)
        return c->fail(B200ANS_E_FORMAT, "pa_decompress: a block's prelude is not a valid model");
    c->stats.ms_model = elapsed(c->ev[3], c->ev[4]);
    c->have_dec = false;

    DecArgs a {};
    a.payload = d_src + payload_at;
    a.offsets = c->offsets.as<uint64_t>();
    a.block_bytes = d_pay_bytes;
    a.out = d_dst;
    a.n = n;
    a.block_syms = h.block_syms;
    a.n_blocks = nm;
    a.fidelity = f;
    a.pa_log2 = d_lg;
    a.pa_nfreqs = c->pa_nfreqs.as<uint32_t>();
    a.pa_nbins = nbins;
    const uint32_t n_tasks = (nm + 7) / 8;
    // Tables: prefix sums + bucket index per block in shared memory, built by the decoder (pa_kernels.cuh), when every
    // frame is at most 2^16; else M-entry slot tables in global memory (pa_dec_tables_kernel + decode_kernel<.., PA>).
    // Bucket index: the finest (2^10 .. 2^8 entries) that still lets one round of tasks be resident.
    a.pa_syms = std::min<uint32_t>(nbins, (c->h_small()->max_symbol + 1u + 15u) & ~15u);
    PaLaunch pl {};
    const int blog_env = env_int("B200ANS_PA_BLOG", 0);
    const uint32_t want_warps = std::min<uint32_t>(uint32_t(kPaCodeMaxWarps), (n_tasks + uint32_t(c->num_sms) - 1) / uint32_t(c->num_sms));
    for (uint32_t blog = blog_env ? uint32_t(blog_env) : 10u; blog >= 8u; blog--) {
        pl = pa_launch_shape(c, 0, size_t(8) * pa_dec_group_bytes(a.pa_syms, blog), n_tasks);
        a.pa_bucket_log2 = blog;
        if (pl.max_warps >= want_warps || blog_env)
            break;
    }
    // (frames of 2^16 slots go to the slot tables: cum[] is 16 bits wide and the searches compare it)
    // ... and only when at most two rounds of tasks result: with a handful of resident warps per SM (plain ANS over 6000
    // ranks: two) the latency-bound shared-memory decoder loses to the slot tables in global memory (12.4 against 7.0 ms)
    const bool smem_tables = c->kernel_path != B200ANS_PATH_GROUP && c->h_small()->max_log2 <= 15u && pl.warps >= 2u && 2u * pl.max_warps >= want_warps;
    set_name(c->stats.decode_kernel, smem_tables ? "pa_smem_decode_kernel" : "decode_kernel");
    if (smem_tables && h_src && runs > 1) {
        // one launch per run of models; the pointers of the argument block move to the run's first model
        CU(c, cudaEventRecord(c->ev[0], c->stream));
        auto k = method == METHOD_MSB ? pa_smem_decode_kernel<METHOD_MSB> : method == METHOD_INT ? pa_smem_decode_kernel<METHOD_INT> : pa_smem_decode_kernel<METHOD_FOLD>;
        const size_t warp_bytes = size_t(8) * pa_dec_group_bytes(a.pa_syms, a.pa_bucket_log2);
        CU(c, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(size_t(pl.max_warps) * warp_bytes)));
        for (int r = 0; r < runs; r++) {
            const uint32_t b0 = r ? run_end[r - 1] : 0, b1 = run_end[r];
            const uint64_t s0 = uint64_t(b0) * h.block_syms, s1 = std::min<uint64_t>(n, uint64_t(b1) * h.block_syms);
            DecArgs ar = a;
            ar.offsets = a.offsets + b0;
            ar.block_bytes = a.block_bytes + b0;
            ar.out = d_dst + s0;
            ar.n = size_t(s1 - s0);
            ar.n_blocks = b1 - b0;
            ar.pa_log2 = d_lg + b0;
            ar.pa_nfreqs = a.pa_nfreqs + size_t(b0) * nbins;
            const PaLaunch pr = pa_launch_shape(c, 0, warp_bytes, (b1 - b0 + 7) / 8);
            CU(c, cudaStreamWaitEvent(c->stream, c->ev_in[r], 0));
            k<<<pr.grid, pr.warps * 32, pr.smem, c->stream>>>(ar);
            c->stats.launches++;
            CU(c, cudaGetLastError());
            CU(c, cudaEventRecord(c->ev_dec[r], c->stream));
            CU(c, cudaStreamWaitEvent(c->copy_out, c->ev_dec[r], 0));
            CU(c, cudaMemcpyAsync(h_dst + s0, d_dst + s0, size_t(s1 - s0) * 4, cudaMemcpyDeviceToHost, c->copy_out));
        }
        CU(c, cudaEventRecord(c->ev[1], c->stream));
        CU(c, cudaEventRecord(c->ev[6], c->copy_out));
        CU(c, cudaStreamWaitEvent(c->stream, c->ev[6], 0)); // the call's work ends on `stream`
        CU(c, cudaStreamSynchronize(c->stream));
        c->stats.ms_decode = elapsed(c->ev[0], c->ev[1]);
        c->stats.n_blocks = nm;
        return B200ANS_OK;
    }
    if (h_src)
        for (int r = 0; r < runs; r++)
            CU(c, cudaStreamWaitEvent(c->stream, c->ev_in[r], 0));
    if (smem_tables) {
        CU(c, cudaEventRecord(c->ev[0], c->stream));
        auto k = method == METHOD_MSB ? pa_smem_decode_kernel<METHOD_MSB> : method == METHOD_INT ? pa_smem_decode_kernel<METHOD_INT> : pa_smem_decode_kernel<METHOD_FOLD>;
        CU(c, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, int(pl.smem)));
        k<<<pl.grid, pl.warps * 32, pl.smem, c->stream>>>(a);
        c->stats.launches++;
        CU(c, cudaGetLastError());
    } else {
        const uint64_t slot_stride = uint64_t(1) << c->h_small()->max_log2;
        if (slot_stride * nm * sizeof(DecEntry) > (uint64_t(48) << 30))
            return c->fail(B200ANS_E_ARG, "pa_decompress: slot tables of all models exceed 48 GiB; decode in pieces");
        CU(c, c->dec_slots.reserve(slot_stride * nm * sizeof(DecEntry)));
        const int tgrid = int(std::min<uint32_t>(nm, uint32_t(c->num_sms) * 8));
        auto tk = method == METHOD_MSB ? pa_dec_tables_kernel<METHOD_MSB> : method == METHOD_INT ? pa_dec_tables_kernel<METHOD_INT> : pa_dec_tables_kernel<METHOD_FOLD>;
        tk<<<tgrid, kPaThreads, (size_t(nbins) + 8) * 4, c->stream>>>(c->pa_nfreqs.as<uint32_t>(), nbins, nm, f, d_lg, c->dec_slots.as<DecEntry>(), slot_stride);
        c->stats.launches++;
        CU(c, cudaGetLastError());
        a.slots = c->dec_slots.as<DecEntry>();
        a.pa_slot_stride = slot_stride;
        CU(c, cudaEventRecord(c->ev[0], c->stream));
        const size_t smem = size_t(kDecWarps) * 8 * kDecRowBytes;
        if (method == METHOD_MSB) {
            auto k = decode_kernel<METHOD_MSB, DEC_FOLD_GLOBAL, true>;
            k<<<persistent_grid(c, k, kDecWarps * 32, smem, n_tasks, kDecWarps), kDecWarps * 32, smem, c->stream>>>(a);
        } else if (method == METHOD_INT) {
            auto k = decode_kernel<METHOD_INT, DEC_FOLD_GLOBAL, true>;
            k<<<persistent_grid(c, k, kDecWarps * 32, smem, n_tasks, kDecWarps), kDecWarps * 32, smem, c->stream>>>(a);
        } else {
            auto k = decode_kernel<METHOD_FOLD, DEC_FOLD_GLOBAL, true>;
            k<<<persistent_grid(c, k, kDecWarps * 32, smem, n_tasks, kDecWarps), kDecWarps * 32, smem, c->stream>>>(a);
        }
        c->stats.launches++;
        CU(c, cudaGetLastError());
    }
    CU(c, cudaEventRecord(c->ev[1], c->stream));
    if (h_dst)
        CU(c, cudaMemcpyAsync(h_dst, d_dst, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->stats.ms_decode = elapsed(c->ev[0], c->ev[1]);
    c->stats.n_blocks = nm;
    return B200ANS_OK;
}

} // namespace

extern "C" int b200ans_pa_decompress_device(b200ans_ctx* c, int method, uint32_t f, uint32_t* d_dst, size_t n,
    const uint8_t* d_src, size_t src_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!d_dst && n) || !d_src || (reinterpret_cast<uintptr_t>(d_src) & 15))
        return c->fail(B200ANS_E_ARG, "pa_decompress: bad argument (d_src must be 16-byte aligned)");
    CU(c, cudaSetDevice(c->device));
    return pa_decompress_core(c, method, f, d_dst, n, d_src, src_bytes, nullptr, nullptr);
}
B200ANS_CATCH_INT

extern "C" int b200ans_pa_compress(b200ans_ctx* c, int method, uint32_t f, uint32_t model_syms, uint8_t* dst, size_t cap,
    const uint32_t* src, size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!dst || !out_bytes || (!src && n) || model_syms == 0)
        return c->fail(B200ANS_E_ARG, "pa_compress: bad argument");
    CU(c, cudaSetDevice(c->device));
    const size_t bound = std::min(b200ans_pa_compress_bound(c, method, model_syms, n), std::max<size_t>(cap + 64, 4096));
    CU(c, c->io_in.reserve(std::max<size_t>(n * 4, 16)));
    CU(c, c->io_out.reserve(bound));
    // (pageable buffers: large copies through the staging threads, as in compress_host_impl)
    if (n && !(use_staging(src, n * 4) && staged_h2d(c, c->io_in.p, src, n * 4, c->stream) == B200ANS_OK))
        CU(c, cudaMemcpyAsync(c->io_in.p, src, n * 4, cudaMemcpyHostToDevice, c->stream));
    size_t bytes = 0;
    RET(b200ans_pa_compress_device(c, method, f, model_syms, c->io_out.as<uint8_t>(), std::min(bound, c->io_out.cap),
        c->io_in.as<uint32_t>(), n, &bytes));
    if (bytes > cap)
        return c->fail(B200ANS_E_CAPACITY, "pa_compress: destination too small");
    if (!(use_staging(dst, bytes) && staged_d2h(c, dst, c->io_out.p, bytes, c->stream, [](size_t) { return cudaEvent_t(nullptr); }) == B200ANS_OK))
        CU(c, cudaMemcpyAsync(dst, c->io_out.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    *out_bytes = bytes;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_pa_decompress(b200ans_ctx* c, int method, uint32_t f, uint32_t* dst, size_t n, const uint8_t* src,
    size_t src_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!dst && n) || !src)
        return c->fail(B200ANS_E_ARG, "pa_decompress: bad argument");
    CU(c, cudaSetDevice(c->device));
    CU(c, c->io_out.reserve(src_bytes + 64));
    CU(c, c->io_in.reserve(std::max<size_t>(n * 4, 16)));
    const int rc = pa_decompress_core(c, method, f, c->io_in.as<uint32_t>(), n, c->io_out.as<uint8_t>(), src_bytes, src, dst);
    if (rc != B200ANS_OK && c->copy_in) { // nothing of a failed call may still be reading or writing the caller's buffers
        cudaStreamSynchronize(c->copy_in);
        cudaStreamSynchronize(c->stream);
        cudaStreamSynchronize(c->copy_out);
        cudaGetLastError();
    }
    return rc;
}
B200ANS_CATCH_INT

// ==========================================================================================
// pseudo-adaptive mode with the driver's dense-rank alphabet remap (src/pseudo_adaptive.cpp:83-125; remap_kernels.cuh)
// ==========================================================================================
// Container: [header, flags = pseudo-adaptive | remap; reserved[0..7] = bytes of the alphabet section; payload_bytes =
// bytes of the inner container][u32 alphabet bytes of block b][pad to 16][alphabets, each 16-byte aligned:
// [sigma][universe][interp of the running sums], byte for byte what the driver writes in front of block b]
// [the pseudo-adaptive container of the RANKS (b200ans_pa_compress_device): prelude(b) + payload(b) = T::encode(ranks of b)]
extern "C" int b200ans_pa_remap_compress_device(b200ans_ctx* c, int method, uint32_t f, uint32_t model_syms, uint8_t* d_dst, size_t cap,
    const uint32_t* d_src, size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!d_dst || !out_bytes || (!d_src && n) || (reinterpret_cast<uintptr_t>(d_src) & 15) || (reinterpret_cast<uintptr_t>(d_dst) & 15))
        return c->fail(B200ANS_E_ARG, "pa_remap_compress: bad argument (device buffers must be 16-byte aligned)");
    if (method != METHOD_MSB && method != METHOD_INT)
        return c->fail(B200ANS_E_ARG, "pa_remap_compress: the remapped blocks are coded with ANSmsb or plain ANS (run<ANSmsb> / run<ANSint>, pseudo_adaptive.cpp:235-236)");
    RET(pa_check(c, method, f, model_syms));
    CU(c, cudaSetDevice(c->device));
    const uint64_t nm64 = (n + model_syms - 1) / model_syms;
    if (nm64 > (uint64_t(1) << 24))
        return c->fail(B200ANS_E_ARG, "pa_remap_compress: more than 2^24 blocks");
    const uint32_t nm = uint32_t(nm64);
    Header h {};
    h.magic = B200ANS_MAGIC;
    h.version = 1;
    h.method = uint8_t(method);
    h.fidelity = uint8_t(f);
    h.n_symbols = n;
    h.block_syms = model_syms;
    h.n_blocks = nm;
    h.flags = kFlagPseudoAdaptive | kFlagRemap;
    const size_t table_at = B200ANS_HEADER_BYTES;
    const size_t alpha_at = align_up(table_at + size_t(nm) * 4, 16);
    if (cap < alpha_at)
        return c->fail(B200ANS_E_CAPACITY, "pa_remap_compress: destination too small");
    if (n == 0) {
        CU(c, cudaMemcpyAsync(d_dst, &h, sizeof h, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaStreamSynchronize(c->stream));
        *out_bytes = B200ANS_HEADER_BYTES;
        return B200ANS_OK;
    }
    // the value range decides the shared-memory bitmap
    CU(c, cudaMemsetAsync(c->small.p, 0, 64, c->stream));
    RET(b200ans_max_device(c, d_src, n, c->d_max()));
    RET(read_small(c));
    const uint32_t max_value = c->h_small()->max_value;
    if (max_value > kRemapMaxValue || remap_smem_bytes(max_value) > c->max_smem)
        return c->fail(B200ANS_E_DOMAIN, "pa_remap_compress: the block alphabets are ranked in shared memory: values up to 2^20 + 2^17");
    const uint64_t alpha_stride = align_up(size_t(8) + size_t(4) * std::min<size_t>(model_syms, size_t(max_value) + 1) + 16, 16);
    CU(c, c->rm_ranks.reserve(align_up(n * 4, 16)));
    CU(c, c->rm_vals.reserve(size_t(nm) * model_syms * 4));
    CU(c, c->rm_sigma.reserve(size_t(nm) * 4));
    CU(c, c->rm_alpha.reserve(size_t(nm) * alpha_stride));
    CU(c, c->rm_bytes.reserve(size_t(nm) * 4));
    CU(c, c->rm_offsets.reserve(size_t(nm) * 8));
    CU(c, c->rm_err.reserve(16));
    CU(c, cudaMemsetAsync(c->rm_err.p, 0, 16, c->stream));
    const size_t smem = remap_smem_bytes(max_value);
    CU(c, cudaFuncSetAttribute(remap_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    const int per_sm = int(std::max<size_t>(1, std::min<size_t>(2, c->max_smem_sm / (smem + 4096))));
    remap_block_kernel<<<int(std::min<uint32_t>(nm, uint32_t(c->num_sms * per_sm))), kRemapThreads, smem, c->stream>>>(d_src, n, model_syms, nm,
        max_value, c->rm_ranks.as<uint32_t>(), c->rm_vals.as<uint32_t>(), c->rm_sigma.as<uint32_t>());
    remap_alpha_kernel<<<(nm + kRemapEncThreads / 32 - 1) / (kRemapEncThreads / 32), kRemapEncThreads, 0, c->stream>>>(c->rm_vals.as<uint32_t>(),
        c->rm_sigma.as<uint32_t>(), model_syms, nm, c->rm_alpha.as<uint8_t>(), alpha_stride, c->rm_bytes.as<uint32_t>(), c->rm_err.as<uint32_t>());
    scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(c->rm_bytes.as<uint32_t>(), nm, c->rm_offsets.as<uint64_t>(), c->d_totals(), nullptr);
    const int cgrid = int(std::max<uint32_t>(1, std::min<uint32_t>((nm + 7) / 8, uint32_t(c->num_sms) * 8)));
    compact_kernel<<<cgrid, 256, 0, c->stream>>>(c->rm_alpha.as<uint8_t>(), alpha_stride, c->rm_bytes.as<uint32_t>(), c->rm_offsets.as<uint64_t>(), nm,
        d_dst + alpha_at, reinterpret_cast<uint32_t*>(d_dst + table_at), uint64_t(cap - alpha_at));
    c->stats.launches += 4;
    CU(c, cudaGetLastError());
    uint32_t err = 0;
    CU(c, cudaMemcpyAsync(&err, c->rm_err.p, 4, cudaMemcpyDeviceToHost, c->stream));
    RET(read_small(c));
    if (err)
        return c->fail(B200ANS_E_DOMAIN, "pa_remap_compress: the running sums of a block's alphabet exceed 32 bits (outside the reference driver's working domain)");
    const uint64_t alpha_total = c->h_small()->totals[0];
    const size_t inner_at = alpha_at + alpha_total;
    if (cap < inner_at + B200ANS_HEADER_BYTES)
        return c->fail(B200ANS_E_CAPACITY, "pa_remap_compress: destination too small");
    size_t inner_bytes = 0;
    RET(b200ans_pa_compress_device(c, method, f, model_syms, d_dst + inner_at, cap - inner_at, c->rm_ranks.as<uint32_t>(), n, &inner_bytes));
    h.payload_bytes = inner_bytes;
    std::memcpy(h.reserved, &alpha_total, 8);
    CU(c, cudaMemcpyAsync(d_dst, &h, sizeof h, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    *out_bytes = inner_at + inner_bytes;
    c->stats.total_bytes = *out_bytes;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_pa_remap_decompress_device(b200ans_ctx* c, int method, uint32_t f, uint32_t* d_dst, size_t n, const uint8_t* d_src,
    size_t src_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!d_dst && n) || !d_src || (reinterpret_cast<uintptr_t>(d_src) & 15))
        return c->fail(B200ANS_E_ARG, "pa_remap_decompress: bad argument (d_src must be 16-byte aligned)");
    CU(c, cudaSetDevice(c->device));
    if (src_bytes < B200ANS_HEADER_BYTES)
        return c->fail(B200ANS_E_FORMAT, "pa_remap_decompress: shorter than a header");
    Header h;
    CU(c, cudaMemcpyAsync(&h, d_src, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    if (h.magic != B200ANS_MAGIC || h.version != 1 || h.flags != (kFlagPseudoAdaptive | kFlagRemap))
        return c->fail(B200ANS_E_FORMAT, "pa_remap_decompress: not a remapped pseudo-adaptive container");
    if (h.method != method || (method != METHOD_MSB && method != METHOD_INT))
        return c->fail(B200ANS_E_FORMAT, "pa_remap_decompress: container was written by another method");
    if (h.n_symbols != n)
        return c->fail(B200ANS_E_ARG, "pa_remap_decompress: symbol count does not match the container");
    if (n == 0)
        return B200ANS_OK;
    const uint32_t model_syms = h.block_syms;
    if (pa_check(c, method, f, model_syms) != B200ANS_OK || h.n_blocks != (n + model_syms - 1) / model_syms)
        return c->fail(B200ANS_E_FORMAT, "pa_remap_decompress: inconsistent block geometry");
    const uint32_t nm = h.n_blocks;
    uint64_t alpha_total = 0;
    std::memcpy(&alpha_total, h.reserved, 8);
    const size_t table_at = B200ANS_HEADER_BYTES;
    const size_t alpha_at = align_up(table_at + size_t(nm) * 4, 16);
    if (alpha_at > src_bytes || alpha_total > src_bytes - alpha_at || (alpha_total & 15) || h.payload_bytes > src_bytes - alpha_at - alpha_total
        || h.payload_bytes < B200ANS_HEADER_BYTES)
        return c->fail(B200ANS_E_FORMAT, "pa_remap_decompress: container truncated");
    const size_t inner_at = alpha_at + alpha_total;
    CU(c, c->rm_ranks.reserve(align_up(n * 4, 32)));
    CU(c, c->rm_vals.reserve(size_t(nm) * model_syms * 4));
    CU(c, c->rm_sigma.reserve(size_t(nm) * 4));
    CU(c, c->rm_offsets.reserve(size_t(nm) * 8));
    CU(c, c->rm_err.reserve(16));
    CU(c, cudaMemsetAsync(c->rm_err.p, 0, 16, c->stream));
    CU(c, cudaMemsetAsync(c->small.p, 0, 64, c->stream));
    const uint32_t* d_sizes = reinterpret_cast<const uint32_t*>(d_src + table_at);
    scan_block_sizes_kernel<<<1, 1024, 0, c->stream>>>(d_sizes, nm, c->rm_offsets.as<uint64_t>(), c->d_totals(), nullptr);
    c->stats.launches++;
    CU(c, cudaGetLastError());
    RET(read_small(c));
    if (c->h_small()->totals[0] != alpha_total)
        return c->fail(B200ANS_E_FORMAT, "pa_remap_decompress: alphabet size table does not add up");
    // The alphabets (a serial bit stream per block, one thread each: 3.5 warps per SM for 16 384 blocks) are decoded on a
    // side stream NEXT TO the ranks: both are chains of dependent steps that leave most issue slots free.
    RET(ensure_side_streams(c));
    CU(c, cudaEventRecord(c->ev_side[0], c->stream));
    CU(c, cudaStreamWaitEvent(c->copy_out, c->ev_side[0], 0));
    remap_alpha_decode_kernel<<<(nm + kRemapAlphaThreads - 1) / kRemapAlphaThreads, kRemapAlphaThreads, 0, c->copy_out>>>(d_src + alpha_at,
        c->rm_offsets.as<uint64_t>(), d_sizes, alpha_total, model_syms, nm, c->rm_vals.as<uint32_t>(), c->rm_sigma.as<uint32_t>());
    remap_alpha_finish_kernel<<<(nm + kRemapEncThreads / 32 - 1) / (kRemapEncThreads / 32), kRemapEncThreads, 0, c->copy_out>>>(c->rm_vals.as<uint32_t>(),
        c->rm_sigma.as<uint32_t>(), model_syms, nm, c->rm_err.as<uint32_t>());
    c->stats.launches += 2;
    CU(c, cudaGetLastError());
    CU(c, cudaEventRecord(c->ev_side[1], c->copy_out));
    const int inner = b200ans_pa_decompress_device(c, method, f, c->rm_ranks.as<uint32_t>(), n, d_src + inner_at, h.payload_bytes);
    CU(c, cudaStreamWaitEvent(c->stream, c->ev_side[1], 0)); // (also on failure: nothing of this call may still be running when it returns)
    if (inner != B200ANS_OK) {
        cudaStreamSynchronize(c->copy_out);
        return inner;
    }
    const int grid = int(std::max<size_t>(1, std::min<size_t>((n + 255) / 256, size_t(c->num_sms) * 16)));
    unremap_kernel<<<grid, 256, 0, c->stream>>>(c->rm_ranks.as<uint32_t>(), n, model_syms, c->rm_vals.as<uint32_t>(), c->rm_sigma.as<uint32_t>(), d_dst,
        c->rm_err.as<uint32_t>());
    c->stats.launches++;
    CU(c, cudaGetLastError());
    uint32_t err = 0;
    CU(c, cudaMemcpyAsync(&err, c->rm_err.p, 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    if (err)
        return c->fail(B200ANS_E_FORMAT, "pa_remap_decompress: a block alphabet or a rank is corrupt");
    return B200ANS_OK;
}
B200ANS_CATCH_INT

// host buffers
extern "C" int b200ans_pa_remap_compress(b200ans_ctx* c, int method, uint32_t f, uint32_t model_syms, uint8_t* dst, size_t cap, const uint32_t* src,
    size_t n, size_t* out_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if (!dst || !out_bytes || (!src && n))
        return c->fail(B200ANS_E_ARG, "pa_remap_compress: bad argument");
    CU(c, cudaSetDevice(c->device));
    // the alphabets can take as much as the block itself (4 bytes per distinct value) on top of the coded ranks
    const size_t bound = b200ans_pa_compress_bound(c, method, model_syms, n) + n * 4 + (n / std::max<uint32_t>(model_syms, 1u) + 2) * 48 + 256;
    CU(c, c->io_in.reserve(std::max<size_t>(n * 4, 16)));
    CU(c, c->io_out.reserve(bound));
    if (n)
        CU(c, cudaMemcpyAsync(c->io_in.p, src, n * 4, cudaMemcpyHostToDevice, c->stream));
    size_t bytes = 0;
    RET(b200ans_pa_remap_compress_device(c, method, f, model_syms, c->io_out.as<uint8_t>(), c->io_out.cap, c->io_in.as<uint32_t>(), n, &bytes));
    if (bytes > cap)
        return c->fail(B200ANS_E_CAPACITY, "pa_remap_compress: destination too small");
    CU(c, cudaMemcpyAsync(dst, c->io_out.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    *out_bytes = bytes;
    return B200ANS_OK;
}
B200ANS_CATCH_INT

extern "C" int b200ans_pa_remap_decompress(b200ans_ctx* c, int method, uint32_t f, uint32_t* dst, size_t n, const uint8_t* src, size_t src_bytes)
try {
    if (!c)
        return B200ANS_E_ARG;
    if ((!dst && n) || !src)
        return c->fail(B200ANS_E_ARG, "pa_remap_decompress: bad argument");
    CU(c, cudaSetDevice(c->device));
    CU(c, c->io_out.reserve(std::max<size_t>(align_up(src_bytes, 16), 64)));
    CU(c, c->io_in.reserve(std::max<size_t>(align_up(n * 4, 32), 32)));
    CU(c, cudaMemcpyAsync(c->io_out.p, src, src_bytes, cudaMemcpyHostToDevice, c->stream));
    RET(b200ans_pa_remap_decompress_device(c, method, f, c->io_in.as<uint32_t>(), n, c->io_out.as<uint8_t>(), src_bytes));
    if (n)
        CU(c, cudaMemcpyAsync(dst, c->io_in.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return B200ANS_OK;
}
B200ANS_CATCH_INT

