/* b200ans -- C ABI of the B200-native semi-static large-alphabet ANS coders.
 *
 * This is the drop-in boundary.  The reference
 * (mpetri/ans-large-alphabet) is header-only C++ with no FFI layer; the entry points a
 * binding would call are its L2 free functions, and each function below names the one it
 * replaces (paths under /root/reference/include):
 *
 *   b200ans_compress / _decompress        ans_int_compress / ans_int_decompress    ans_int.hpp:197,257
 *                                         ans_fold_compress<f> / _decompress<f>    ans_fold.hpp:238,283
 *                                         ans_msb_compress / _decompress           ans_msb.hpp:235,294
 *                                         ans_reorder_fold_compress<f> / _de..<f>  ans_reorder_fold.hpp:312,357
 *                                         (what methods.hpp:484-567 ANSint/ANSmsb/ANSfold<f>/ANSrfold<f> call)
 *   b200ans_histogram_device              the counting loops of ans_fold.hpp:74-78, ans_msb.hpp:81-85,
 *                                         ans_int.hpp:42-49, ans_reorder_fold.hpp:75-84,110-115
 *   b200ans_adjust_freqs                  adjust_freqs            ans_util.hpp:100-157
 *   b200ans_serialize_interp              ans_serialize_interp    ans_util.hpp:46-63
 *   b200ans_load_interp                   ans_load_interp         ans_util.hpp:25-42
 *   b200ans_interp_encode / _decode       interpolative_internal::encode / decode   interp.hpp:100-118
 *   b200ans_model_* / _encode_device /    ans_X_encode::create / serialize / encode_symbol / flush_state and
 *   _decode_device                        ans_X_decode::load / init_state / decode_sym (ans_fold.hpp:68-125,177-235 ...)
 *
 * Conventions: plain pointers and sizes only; every function returns 0 (B200ANS_OK) or a
 * negative error code; nothing throws or calls exit(); the library never frees caller memory.
 * "d_" parameters are device pointers on the context's GPU, everything else is host memory.
 * There is NO CPU fallback: without a usable CUDA device b200ans_create fails.
 *
 * Stream format.  block_syms > 0 (default 8192): a blocked container
 *     [64-byte header][reference-format prelude][u32 bytes-per-block x n_blocks][pad to 16]
 *     [block payloads, each 16-byte aligned]
 * where every block payload is byte-identical to what the reference coder emits for that
 * sub-range with the same (global) model: raw low bytes + renormalisation words in reverse
 * input order, then the four 64-bit states (ans_fold.hpp:250-278).
 * block_syms == 0: "reference wire format" -- prelude + one payload, no header; the output
 * equals the reference's ans_X_compress() byte for byte (modulo the reference's own
 * uninitialised prelude padding bits) and reference-produced streams decode. One block means
 * four chains: correct, but slow on a GPU; it exists for interoperability and tests.
 */
#ifndef B200ANS_H
#define B200ANS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ANS_OK 0
#define B200ANS_E_ARG (-1)      /* bad argument (null pointer, unknown method, fidelity outside 1..8, alignment) */
#define B200ANS_E_CUDA (-2)     /* a CUDA call failed; see b200ans_last_error */
#define B200ANS_E_CAPACITY (-3) /* destination too small */
#define B200ANS_E_DOMAIN (-4)   /* input outside the reference's valid domain */
#define B200ANS_E_FORMAT (-5)   /* not a b200ans container / corrupt header */
#define B200ANS_E_STATE (-6)    /* call sequence error (e.g. encode before a model exists) */
#define B200ANS_E_NCCL (-7)     /* NCCL missing (libnccl.so.2 is loaded on first use) or a collective failed */
#define B200ANS_E_NOMEM (-8)    /* host allocation failed */
#define B200ANS_E_INTERNAL (-9) /* an unexpected C++ exception was stopped at the ABI boundary */

#define B200ANS_METHOD_INT 0   /* "ANS"        plain large-alphabet ANS, ans_int.hpp          */
#define B200ANS_METHOD_FOLD 1  /* "ANSfold-f"  ans_fold.hpp                                   */
#define B200ANS_METHOD_MSB 2   /* "ANSmsb"     ans_msb.hpp                                    */
#define B200ANS_METHOD_RFOLD 3 /* "ANSrfold-f" ans_reorder_fold.hpp                           */
#define B200ANS_METHOD_BYTE 4  /* "ANSbyte"    ans_byte.hpp (bytes: values below 256)            */

/* kernel families (both produce identical bytes; AUTO picks by block count) */
#define B200ANS_PATH_AUTO 0
#define B200ANS_PATH_GROUP 1 /* 4 lanes per block, one lane per rANS chain: small / medium inputs */
#define B200ANS_PATH_LPB 2   /* one lane per block, four chains per lane: large inputs          */

#define B200ANS_DEFAULT_BLOCK_SYMS 8192u
#define B200ANS_HEADER_BYTES 64u
#define B200ANS_MAGIC 0x4E413242u /* "B2AN" */

typedef struct b200ans_ctx b200ans_ctx;

/* Timings of the most recent call, measured with CUDA events on the context's stream. */
typedef struct b200ans_stats {
    float ms_histogram;   /* K1 (all histogram / max launches)                 */
    float ms_model;       /* host: normalise + prelude + table build + upload  */
    float ms_encode;      /* K2 block encoder                                  */
    float ms_compact;     /* size scan + payload compaction                    */
    float ms_decode;      /* K3 block decoder                                  */
    float ms_h2d;         /* host->device copies inside the call               */
    float ms_d2h;         /* device->host copies inside the call               */
    uint64_t launches;    /* kernels launched by this context since creation   */
    uint64_t frame_size;  /* M of the current model                            */
    uint32_t n_symbols_model; /* max_sym + 1                                   */
    uint32_t n_blocks;
    uint64_t prelude_bytes;
    uint64_t payload_bytes; /* sum of block payload bytes (unpadded)           */
    uint64_t total_bytes;   /* whole container                                 */
    char encode_kernel[32]; /* block encoder / decoder the most recent call launched   */
    char decode_kernel[32]; /* (as ncu names them: "lpb3_encode_kernel", ...)          */
    float ms_allreduce;     /* histogram / maximum all-reduce (sharded calls)          */
    uint32_t model_flags;   /* B200ANS_MODEL_* of the current encoder model            */
    uint64_t model_cache_hits; /* b200ans_model_load calls (decompress included) answered by the decoder model already
                                * on the device: same method, fidelity and prelude bytes (since creation)           */
    uint64_t peer_exchanges;   /* histogram exchanges that ran over NVLink peer memory (b200ans_comm_ipc_*; since
                                * creation) -- the others went through ncclAllReduce                                 */
} b200ans_stats;
/* The model was normalised from 2^31 symbols or more: the reference's cross_entropy() sums the counts in an `int`
 * (util.hpp:288-289), so adjust_freqs' stopping rule changes -- in [2^31, 2^32) the first frame that fits is taken
 * (a tiny frame, more bits per symbol), from 2^32 on the frame grows until a frequency nears 16 bits (2^18 .. 2^19
 * slots).  b200ans reproduces this bit for bit (the frame is part of the stream format); the flag tells the caller. */
#define B200ANS_MODEL_INT_ACCUMULATOR_QUIRK 1u
/* ANSrfold: the value reordering came from the fast route -- a sample proposes candidates, ONE pass counts them exactly
 * and bounds every other value, and the route only answers when that bound proves the selection exact (otherwise the
 * raw histogram of the whole value range decides, as in ans_reorder_fold.hpp:75-106).  Same model either way. */
#define B200ANS_MODEL_RFOLD_FAST 2u

/* What a container says about itself (the reference stores neither n nor the byte count, methods.hpp:541-546). */
typedef struct b200ans_container {
    int method;            /* B200ANS_METHOD_*                                   */
    uint32_t fidelity;
    uint64_t n_symbols;
    uint32_t block_syms;
    uint32_t n_blocks;
    uint64_t prelude_bytes;
    uint64_t payload_bytes; /* padded sum of the block payloads                  */
    uint64_t total_bytes;   /* header + prelude + size table + payloads          */
} b200ans_container;

/* ---- context ------------------------------------------------------------------------- */
/* stream: a cudaStream_t (as void*) to run on, or NULL to let the context create its own. */
int b200ans_create(b200ans_ctx** out, int device, void* stream);
void b200ans_destroy(b200ans_ctx* ctx);
int b200ans_set_block_syms(b200ans_ctx* ctx, uint32_t block_syms); /* 0 or a multiple of 128 */
int b200ans_set_kernel_path(b200ans_ctx* ctx, int path); /* B200ANS_PATH_* ; default AUTO */
/* Block size for an input of n symbols (per GPU): the multiple of 128 in [lo, hi] whose blocks fill whole rounds of the
 * lane-per-block coders (32 blocks per warp, 16 warps on each SM; a round lasts as long as one block), larger sizes
 * winning ties. 2^30 symbols, [8192, 16384] -> 14208: one round of 16 warps per SM instead of two of 14 at 8192.
 * The reference has no blocks (one stream per call, ans_fold.hpp:238-281); the block size is this container's own
 * parameter and is stored in its header. ctx may be NULL (148 SMs assumed). Returns 0 for lo > hi or hi < 128. */
uint32_t b200ans_suggest_block_syms(const b200ans_ctx* ctx, size_t n, uint32_t lo, uint32_t hi);
/* H_approx of adjust_freqs (ans_util.hpp:100-101; ANSsint<H> / ANSsmsb<H>, methods.hpp:584-620): the frame stops
 * growing once the cross entropy is within H/1000 of the entropy. Default 1 (ANS / ANSmsb / ANSfold / ANSrfold). */
int b200ans_set_h_approx(b200ans_ctx* ctx, uint32_t h_approx);
int b200ans_get_stats(const b200ans_ctx* ctx, b200ans_stats* out);
const char* b200ans_last_error(const b200ans_ctx* ctx);
const char* b200ans_strerror(int code);
const char* b200ans_method_name(int method, uint32_t fidelity, char* buf, size_t buflen);
/* upper bound on the container size for n symbols */
/* Host-only: parse the 64-byte header at `src` (host memory). B200ANS_E_FORMAT if it is not a b200ans container. */
int b200ans_container_info(const uint8_t* src, size_t src_bytes, b200ans_container* out);
size_t b200ans_compress_bound(const b200ans_ctx* ctx, int method, size_t n);

/* ---- one-shot, host buffers (the methods.hpp encode/decode path) ------------------------ */
int b200ans_compress(b200ans_ctx* ctx, int method, uint32_t fidelity, uint8_t* dst, size_t dst_capacity,
    const uint32_t* src, size_t n, size_t* out_bytes);
int b200ans_decompress(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t* dst, size_t n,
    const uint8_t* src, size_t src_bytes);

/* ---- one-shot, device-resident buffers (d_src / d_dst 16-byte aligned) ------------------ */
int b200ans_compress_device(b200ans_ctx* ctx, int method, uint32_t fidelity, uint8_t* d_dst, size_t dst_capacity,
    const uint32_t* d_src, size_t n, size_t* out_bytes);
int b200ans_decompress_device(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t* d_dst, size_t n,
    const uint8_t* d_src, size_t src_bytes);

/* ---- staged path (multi-GPU: histogram -> all-reduce by the caller -> model -> encode) -- */
/* max over d_src, atomically max-combined into *d_max (device u32, caller-initialised). */
int b200ans_max_device(b200ans_ctx* ctx, const uint32_t* d_src, size_t n, uint32_t* d_max);
/* adds the (mapped) symbol counts of d_src into d_bins[nbins] (device u64, caller-zeroed).
 * nbins: INT -> max value + 1; FOLD / RFOLD -> 2^(f+9); MSB -> 1280.
 * RFOLD uses the context's current value mapping (b200ans_rfold_prepare) -- call with
 * method INT on the raw values first to obtain the full-alphabet histogram. */
int b200ans_histogram_device(b200ans_ctx* ctx, int method, uint32_t fidelity, const uint32_t* d_src, size_t n,
    uint64_t* d_bins, size_t nbins);
/* the same, plus the largest raw value of the input to *d_max (device); values whose bin is beyond nbins are not
 * counted, so plain ANS (ans_int.hpp:42-49: maximum first, then counts) can count optimistically in one pass */
int b200ans_histogram_max_device(b200ans_ctx* ctx, int method, uint32_t fidelity, const uint32_t* d_src, size_t n,
    uint64_t* d_bins, size_t nbins, uint32_t* d_max);
/* RFOLD only: from the full-alphabet histogram (host u64[nvalues]) pick the 2^(f+7) most
 * frequent values, ties to the smaller value (ans_reorder_fold.hpp:79-106), and upload the
 * value -> reordered-value table. */
int b200ans_rfold_prepare(b200ans_ctx* ctx, uint32_t fidelity, const uint64_t* raw_bins, size_t nvalues);
/* the same on the device (ans_reorder_fold.hpp:75-115, the whole of create() before adjust_freqs): d_raw_bins is the
 * full-alphabet histogram in device memory (u64[nvalues], e.g. from b200ans_histogram_max_device with method INT,
 * all-reduced by the caller when the job is sharded); selects the 2^(f+7) most frequent values (radix select on the
 * counts, ties to the smaller value), fills the value -> reordered-value table, and writes the FOLDED histogram over
 * the reordered values to d_folded_bins (device u64[2^(f+9)]) -- derived from the raw bins, no second pass over the
 * input.  most_frequent[] comes back for the prelude; one stream synchronisation. */
int b200ans_rfold_prepare_device(b200ans_ctx* ctx, uint32_t fidelity, const uint64_t* d_raw_bins, size_t nvalues,
    uint64_t* d_folded_bins);
/* host histogram -> normalised model (adjust_freqs) + prelude + device encoder tables. */
int b200ans_model_build(b200ans_ctx* ctx, int method, uint32_t fidelity, const uint64_t* bins, size_t nbins);
/* ans_X_encode::create(in_u32, n) as one call (ans_fold.hpp:68-92, ans_int.hpp:38-68, ans_msb.hpp:77-101,
 * ans_reorder_fold.hpp:70-130): device histogram (+ ANSrfold's selection of the most frequent values) -> normalised
 * model + prelude + device encoder tables; becomes the context's current encoder model.  Host buffer (copied to the
 * device) or device buffer (16-byte aligned). */
int b200ans_model_create(b200ans_ctx* ctx, int method, uint32_t fidelity, const uint32_t* src, size_t n);
int b200ans_model_create_device(b200ans_ctx* ctx, int method, uint32_t fidelity, const uint32_t* d_src, size_t n);
/* ANSrfold: most_frequent[] (2^(f+7) values, or none for the identity mapping) and sigma of the current encoder model
 * (ans_reorder_fold.hpp:180-184) */
int b200ans_model_most_frequent(b200ans_ctx* ctx, uint32_t* dst, size_t capacity, size_t* out_count, uint64_t* sigma);
/* prelude of the current model; returns its size through out_bytes */
int b200ans_model_prelude(b200ans_ctx* ctx, uint8_t* dst, size_t capacity, size_t* out_bytes);
/* normalised frequencies of the current model (count = max_sym + 1) */
int b200ans_model_nfreqs(b200ans_ctx* ctx, uint32_t* dst, size_t capacity, size_t* out_count, uint64_t* frame_size);
/* decoder side: parse a reference-format prelude, build + upload the slot tables */
int b200ans_model_load(b200ans_ctx* ctx, int method, uint32_t fidelity, const uint8_t* prelude, size_t prelude_bytes,
    size_t* consumed);
/* encode d_src with the current model into a container at d_dst */
int b200ans_encode_device(b200ans_ctx* ctx, uint8_t* d_dst, size_t dst_capacity, const uint32_t* d_src, size_t n,
    size_t* out_bytes);
/* per-block payload access for parity tests: encode, leaving payload b at d_scratch + b*stride */
int b200ans_encode_blocks_device(b200ans_ctx* ctx, const uint32_t* d_src, size_t n, uint8_t* d_scratch,
    size_t stride, uint32_t* d_block_bytes);
int b200ans_block_stride(const b200ans_ctx* ctx, int method, size_t* stride);
/* decode payload blocks that sit at d_payload + d_offsets[b] with sizes d_block_bytes[b] using
 * the currently loaded decoder model */
int b200ans_decode_blocks_device(b200ans_ctx* ctx, const uint8_t* d_payload, const uint64_t* d_offsets,
    const uint32_t* d_block_bytes, size_t n, uint32_t* d_dst);

/* ---- decoding without any allocation ---------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * b200ans_decode_blocks_device's work for blocks laid out as in the container's payload area -- block b starts at
 * d_payload + sum over b' < b of roundup16(d_block_bytes[b']) -- with the current decoder model (b200ans_model_load is
 * the set-up step, and the one that may allocate).  The per-block offsets, which b200ans_decompress_device keeps in a
 * cudaMalloc'ed grow-only buffer, live in a file-scope device array of B200ANS_NOALLOC_MAX_BLOCKS entries (513 KiB, held
 * by every GPU that loads the library until the process exits); this call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine.  More blocks than that (n / block_syms, rounded up): B200ANS_E_ARG.
 * d_payload 16-byte aligned, payload_bytes = bytes readable behind it (the sizes must not add up to more);
 * block size = the context's (b200ans_set_block_syms); symbols go to d_dst[0..n).
 * ALL calls on one GPU share the array and it keeps the last call's contents (everything a call reads is rewritten by
 * the same call first, in stream order): calls that overlap on the GPU -- two contexts or streams -- race. */
#define B200ANS_NOALLOC_MAX_BLOCKS 65536u
int b200ans_decode_blocks_noalloc(b200ans_ctx* ctx, const uint8_t* d_payload, size_t payload_bytes,
    const uint32_t* d_block_bytes, size_t n, uint32_t* d_dst);

/* ---- multi-GPU: one context per GPU (one process per GPU, or one thread per GPU) ---------- */
/* the input is split into per-GPU block ranges; the single semi-static model comes from ONE
 * all-reduce (sum) of the histogram -- plus the maximum for the coders whose alphabet is the value range -- after
 * which every rank runs the same deterministic normalisation on the same counts (identical model, no broadcast) and
 * codes its own shard.  Decoding needs no communication: b200ans_decompress[_device] per shard.
 * The reference has no counterpart (single process, ans_fold.hpp:68-92 counts one array); NCCL is opened with
 * dlopen("libnccl.so.2") on first use, so a process that already holds NCCL (torch) shares that copy. */
#define B200ANS_COMM_ID_BYTES 128
/* rank 0: a fresh NCCL unique id (ncclGetUniqueId); hand the bytes to every rank by any means */
int b200ans_comm_unique_id(uint8_t id[B200ANS_COMM_ID_BYTES]);
/* collective over all ranks: ncclCommInitRank on the context's device; the context owns the communicator */
int b200ans_comm_init(b200ans_ctx* ctx, const uint8_t id[B200ANS_COMM_ID_BYTES], int rank, int world);
/* or borrow a communicator the caller already has (an ncclComm_t as void*); never destroyed by the library */
int b200ans_comm_attach(b200ans_ctx* ctx, void* nccl_comm, int rank, int world);
int b200ans_comm_release(b200ans_ctx* ctx);
/* Ranks that are processes on ONE NVLink / NVSwitch node can run the exchange over peer memory instead (CUDA IPC):
 * every rank exports its exchange buffer (_export: 64 opaque bytes), hands the bytes to every other rank by any means,
 * and maps the others' (_attach: `handles` = world x 64 bytes in rank order; every rank must have exported before any
 * rank attaches).  b200ans_allreduce_hist and b200ans_compress_sharded[_device] then exchange payloads of up to 8192
 * bins in ONE kernel -- remote stores into every peer's buffer, a release / acquire flag per peer, a local sum -- and
 * keep NCCL for anything larger (ANSrfold's raw bins), so a communicator is still needed for those.  A rank that does
 * not arrive within seconds makes the call fail with B200ANS_E_NCCL (the staged b200ans_allreduce_hist leaves ~0 in
 * d_bins[0]).  _release: collective in effect -- no rank may still be exchanging; b200ans_destroy calls it.
 * No counterpart in the reference. */
#define B200ANS_IPC_HANDLE_BYTES 64
int b200ans_comm_ipc_export(b200ans_ctx* ctx, uint8_t handle[B200ANS_IPC_HANDLE_BYTES]);
int b200ans_comm_ipc_attach(b200ans_ctx* ctx, int rank, int world, const uint8_t* handles);
int b200ans_comm_ipc_release(b200ans_ctx* ctx);
/* in place, on the context's stream: d_bins[nbins] (device u64) summed over the ranks and, when d_max is not NULL,
 * *d_max (device u32) replaced by the maximum over the ranks -- one kernel over peer memory after b200ans_comm_ipc_attach,
 * else one NCCL group. Without either: no-op. */
int b200ans_allreduce_hist(b200ans_ctx* ctx, uint64_t* d_bins, size_t nbins, uint32_t* d_max);
/* b200ans_compress_device / b200ans_compress on this rank's shard with the GLOBAL model: histogram (K1) -> all-reduce
 * -> normalise -> encode (K2).  Collective: every rank of the communicator must call it, a rank whose shard is empty
 * (n == 0) included -- it takes part in the reduction and writes an empty container.  Each rank gets its own
 * container (the same prelude in all of them). */
int b200ans_compress_sharded_device(b200ans_ctx* ctx, int method, uint32_t fidelity, uint8_t* d_dst, size_t dst_capacity,
    const uint32_t* d_src, size_t n, size_t* out_bytes);
int b200ans_compress_sharded(b200ans_ctx* ctx, int method, uint32_t fidelity, uint8_t* dst, size_t dst_capacity,
    const uint32_t* src, size_t n, size_t* out_bytes);

/* ---- pseudo-adaptive mode (src/pseudo_adaptive.cpp:64-177) ------------------------------ */
/* Every `model_syms` consecutive symbols (a multiple of 128 in [1024, 2^24]) are coded with their own semi-static
 * model -- own histogram, own adjust_freqs, own prelude -- as the reference's driver does by calling T::encode block by
 * block; nothing is shared between blocks, so shards need no collective. Methods: ANSfold-1, ANSfold-2, ANSmsb, and
 * plain ANS (B200ANS_METHOD_INT) for inputs whose values do not exceed model_syms <= 16256 -- the dense ranks the
 * driver's remap produces (below); a block's alphabet is then 0 .. model_syms and its frame at most 2^15 slots.
 * Container: header (flags bit 0), u32 prelude bytes and u32 payload bytes per block, the preludes, the payloads;
 * prelude(b) followed by payload(b) is byte for byte what ans_X_compress() writes for block b (ans_fold.hpp:238-281).
 * The reference driver's dense-rank alphabet remap (pseudo_adaptive.cpp:83-105) is not applied: values are coded as
 * they are (the folded coders do not need a dense alphabet). */
size_t b200ans_pa_compress_bound(const b200ans_ctx* ctx, int method, uint32_t model_syms, size_t n);
int b200ans_pa_compress(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t model_syms, uint8_t* dst,
                        size_t dst_capacity, const uint32_t* src, size_t n, size_t* out_bytes);
int b200ans_pa_decompress(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t* dst, size_t n, const uint8_t* src,
                          size_t src_bytes);
int b200ans_pa_compress_device(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t model_syms, uint8_t* d_dst,
                               size_t dst_capacity, const uint32_t* d_src, size_t n, size_t* out_bytes);
int b200ans_pa_decompress_device(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t* d_dst, size_t n,
                                 const uint8_t* d_src, size_t src_bytes);

/* ---- pseudo-adaptive mode with the driver's dense-rank alphabet remap (src/pseudo_adaptive.cpp:83-125) ----
 * Per block of model_syms symbols: the values that occur become the ranks 1..sigma, the block is coded as those ranks
 * with ANSmsb or plain ANS (its own model, as b200ans_pa_compress), and the block alphabet goes in front as the driver writes it:
 * [u32 sigma][u32 universe][interpolative code of the running sums of the sorted values].  Alphabet bytes and
 * prelude + payload of every block are byte-identical to the driver's output for that block (a block with one distinct
 * value still gets a coded body here; the driver writes the alphabet only).  Values up to 2^20 + 2^17 (the block's value
 * range is ranked in shared memory); an alphabet whose running sums pass 32 bits is B200ANS_E_DOMAIN, as it is outside
 * the driver's own working domain.  method: B200ANS_METHOD_MSB or B200ANS_METHOD_INT (run<ANSmsb> / run<ANSint>,
 * pseudo_adaptive.cpp:235-236; plain ANS: model_syms <= 16256, and a block with one distinct value -- which the driver
 * never codes and for which ans_int.hpp's unguarded adjust_freqs would not terminate -- gets the guarded normaliser's
 * model). */
int b200ans_pa_remap_compress(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t model_syms, uint8_t* dst, size_t dst_capacity,
    const uint32_t* src, size_t n, size_t* out_bytes);
int b200ans_pa_remap_decompress(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t* dst, size_t n, const uint8_t* src, size_t src_bytes);
int b200ans_pa_remap_compress_device(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t model_syms, uint8_t* d_dst, size_t dst_capacity,
    const uint32_t* d_src, size_t n, size_t* out_bytes);
int b200ans_pa_remap_decompress_device(b200ans_ctx* ctx, int method, uint32_t fidelity, uint32_t* d_dst, size_t n, const uint8_t* d_src,
    size_t src_bytes);

/* ---- ans_byte (ans_byte.hpp): the byte coder behind the vbyteANS / streamvbyteANS baselines (methods.hpp:432-482) ----
 * B200ANS_METHOD_BYTE: 256 symbols, frame <= 4096, the reference's byte normaliser (ans_byte.hpp:41-97) and prelude
 * (:122-137), the same 4-state coder as plain ANS (:236-305).  The method also runs through every uint32 entry point
 * above (values below 256); these take and return bytes.  Block payloads and, with block_syms = 0, the whole stream
 * are byte-identical to ans_byte_compress. */
int b200ans_byte_compress(b200ans_ctx* ctx, uint8_t* dst, size_t dst_capacity, const uint8_t* src, size_t n, size_t* out_bytes);
int b200ans_byte_decompress(b200ans_ctx* ctx, uint8_t* dst, size_t n, const uint8_t* src, size_t src_bytes);
/* device buffers: bytes 4-byte aligned, the container 16-byte aligned */
int b200ans_byte_compress_device(b200ans_ctx* ctx, uint8_t* d_dst, size_t dst_capacity, const uint8_t* d_src, size_t n, size_t* out_bytes);
int b200ans_byte_decompress_device(b200ans_ctx* ctx, uint8_t* d_dst, size_t n, const uint8_t* d_src, size_t src_bytes);
/* adjust_freqs of ans_byte.hpp:41-97 (host only): freqs[256] -> out[256], summing to min(4096, 64 x distinct) rounded up to 2^k */
int b200ans_byte_adjust_freqs(const uint64_t* freqs, uint16_t* out);

/* ---- host-only model API (ans_util.hpp / interp.hpp); usable without a GPU ------------- */
/* out: largest_sym + 1 entries */
int b200ans_adjust_freqs(const uint64_t* freqs, size_t nbins, uint32_t largest_sym, int require_u16,
    uint32_t H_approx, uint32_t* out);
/* scale_freqs (ans_util.hpp:77-95): one fit of the counts F[nbins] to the frame M -- symbols in `mapping` order
 * (sigma entries, rarest first), running ratio, (uint32_t)(0.5 + ratio * F), floor at 1.  S[nbins] receives the
 * shares; *fits = 1 iff the frame came out exactly (the reference's return value). */
int b200ans_scale_freqs(uint32_t* S, const uint64_t* F, size_t nbins, const uint32_t* mapping, size_t sigma, int64_t M,
    uint64_t freq_sum, int* fits);
int b200ans_serialize_interp(const uint32_t* nfreqs, size_t count, uint64_t frame_size, uint8_t* dst, size_t capacity,
    size_t* out_bytes);
/* dst may be NULL to query the count; consumed may be NULL */
int b200ans_load_interp(const uint8_t* src, uint32_t* dst, size_t capacity, size_t* out_count, size_t* consumed);
int b200ans_interp_encode(uint32_t* dst, const uint32_t* in, size_t n, size_t universe, size_t* out_bytes);
int b200ans_interp_decode(const uint32_t* src, uint32_t* out, size_t n, size_t universe, size_t* words_read);
/* mapping rules (ans_fold.hpp:38-65,150-175; ans_msb.hpp:41-73,156-176) */
uint32_t b200ans_fold_map(uint32_t fidelity, uint32_t x, uint32_t* nbytes);
uint32_t b200ans_fold_unmap(uint32_t fidelity, uint32_t sym, uint32_t* nbytes);
uint32_t b200ans_msb_map(uint32_t x, uint32_t* nbytes);
uint32_t b200ans_msb_unmap(uint32_t sym, uint32_t* nbytes);

#ifdef __cplusplus
}
#endif
#endif /* B200ANS_H */
