/* TEST INFRASTRUCTURE ONLY -- never linked into, or loaded by, the product library.
 *
 * Plain-C restatement ("port") of the semi-static large-alphabet ANS coders of
 * mpetri/ans-large-alphabet: ans_int, ans_fold<f>, ans_msb, ans_reorder_fold<f>, their
 * ans_util.hpp model utilities and the interp.hpp/vbyte.hpp prelude.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks every function here
 * against the unmodified reference headers compiled into oracle/_ref/libansref.so
 * (oracle/ref_shim.cpp), and tests/golden/ holds vectors generated from that library.
 */
#ifndef ANS_ORACLE_H
#define ANS_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_INT = 0, ORC_FOLD = 1, ORC_MSB = 2, ORC_RFOLD = 3, ORC_BYTE = 4 }; /* ORC_BYTE: ans_byte.hpp, symbols < 256 handed over as uint32_t */

typedef struct orc_model orc_model;

/* mappings: ans_fold.hpp:38-65,150-175 ; ans_msb.hpp:41-73,156-176 */
uint32_t orc_fold_mapping(uint32_t f, uint32_t x, uint32_t* nbytes);
void orc_fold_undo(uint32_t f, uint32_t s, uint32_t* value, uint32_t* nbytes);
uint32_t orc_msb_mapping(uint32_t x, uint32_t* nbytes);
void orc_msb_undo(uint32_t s, uint32_t* value, uint32_t* nbytes);

/* ans_util.hpp:77-157 (+ util.hpp:271-298) */
int orc_adjust_freqs(const uint64_t* freqs, size_t nbins, uint32_t largest_sym, int require_u16,
    uint32_t H_approx, uint32_t* out);
/* ans_util.hpp:46-63 / :25-42 ; returns bytes written / number of entries */
long orc_serialize_interp(const uint32_t* nfreqs, size_t count, uint64_t frame_size, uint8_t* out);
long orc_load_interp(const uint8_t* in, uint32_t* out);
/* interp.hpp:100-118 */
/* ans_byte.hpp:41-97: freqs[256] -> out[256] */
void orc_byte_adjust_freqs(const uint64_t* freqs, uint16_t* out);
long orc_interp_encode(uint32_t* dst, const uint32_t* in, size_t n, size_t u);
long orc_interp_decode(const uint32_t* src, uint32_t* out, size_t n, size_t u);

/* whole-stream coders: ans_X_compress / ans_X_decompress */
long orc_compress(int method, uint32_t f, uint8_t* dst, size_t cap, const uint32_t* src, size_t n);
int orc_decompress(int method, uint32_t f, uint32_t* dst, size_t n, const uint8_t* src, size_t bytes);

/* model objects + per-block coder (one global model, 4-state loop per sub-range) */
orc_model* orc_model_create(int method, uint32_t f, const uint32_t* in, size_t n);
orc_model* orc_model_load(int method, uint32_t f, const uint8_t* prelude);
long orc_model_info(orc_model* m, uint64_t* frame_size, uint32_t* nfreqs_out);
long orc_model_serialize(orc_model* m, uint8_t* out);
long orc_model_rfold_info(orc_model* m, uint64_t* sigma, uint32_t* most_frequent_out);
long orc_encode_block(orc_model* m, const uint32_t* in, size_t n, uint8_t* dst);
int orc_decode_block(orc_model* m, const uint8_t* src, size_t bytes, uint32_t* out, size_t n);
void orc_model_free(orc_model* m);

/* src/pseudo_adaptive.cpp:83-125, one block of the driver with the dense-rank alphabet remap: returns the bytes
 * written ([u32 sigma][u32 universe][interp of the running sums of the present values] + T::encode of the ranks),
 * *alphabet_bytes = the part before T::encode; -2 when the running sums leave 32 bits */
long orc_pa_remap_block(int method, uint32_t f, const uint32_t* in, size_t n, uint8_t* dst, size_t cap,
    uint32_t* remapped_out, size_t* alphabet_bytes);

/* single-thread timing helper for bench.py's cpu_baseline ("kind": "port") */
int orc_time(int method, uint32_t f, const uint32_t* in, size_t n, int runs, double* enc_ns,
    double* dec_ns, size_t* bytes);

#ifdef __cplusplus
}
#endif
#endif
