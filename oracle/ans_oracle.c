/* TEST INFRASTRUCTURE ONLY -- never linked into, or loaded by, the product library.
 *
 * Plain-C restatement of the reference's semi-static large-alphabet ANS path
 * (mpetri/ans-large-alphabet @416c7e7).  Every function cites the reference
 * file:line (under /root/reference/include) whose behaviour it restates.
 * Parity status: PINNED against the unmodified reference (see ans_oracle.h).
 *
 * Floating point: the reference is built with GCC's default -ffp-contract=fast on an
 * FMA-capable x86-64 (CMakeLists.txt:10-13, -march=native), which fuses exactly three
 * expressions on this path (checked in the disassembly of oracle/_ref/libansref.so):
 *   scale_freqs   0.5 + aratio * F          (ans_util.hpp:86)
 *   entropy       H0 += p * log2(p)         (util.hpp:278)
 *   cross_entropy H0 += p * log2(q)         (util.hpp:294)
 * They are written as explicit fma() here so the result does not depend on flags.
 */
#include "ans_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define ANS_K 16ULL          /* ans_byte.hpp:30  constants::K           */
#define ANS_RADIX_LOG2 32    /* ans_byte.hpp:28  constants::RADIX_LOG2  */
#define MSB_MAX_SIGMA 1280u  /* ans_msb.hpp:29                           */

/* ------------------------------------------------------------------------- mappings */

/* ans_fold.hpp:38-50 (and :52-65 for the byte count); ans_reorder_fold.hpp:36-64 is
 * the same function. radix = 8. */
uint32_t orc_fold_mapping(uint32_t f, uint32_t x, uint32_t* nbytes)
{
    const uint32_t thres = 1u << (f + 7);
    const uint32_t step = (1u << (f - 1)) * 255u;
    uint32_t off = 0, k = 0;
    while (x >= thres) {
        x >>= 8;
        off += step;
        k++;
    }
    if (nbytes)
        *nbytes = k;
    return x + off;
}

/* ans_fold.hpp:150-161 (value) and :165-175 (byte count) */
void orc_fold_undo(uint32_t f, uint32_t s, uint32_t* value, uint32_t* nbytes)
{
    const uint32_t thres = 1u << (f + 7);
    const uint32_t div = (1u << (f - 1)) * 255u;
    if (s < thres) {
        *value = s;
        *nbytes = 0;
        return;
    }
    uint32_t k = (s - thres) / div + 1;
    uint32_t x_org = s - div * k;
    *value = x_org << (8 * k);
    *nbytes = k;
}

/* ans_msb.hpp:41-73 */
uint32_t orc_msb_mapping(uint32_t x, uint32_t* nbytes)
{
    uint32_t k, s;
    if (x <= 256) {
        k = 0;
        s = x;
    } else if (x <= (1u << 16)) {
        k = 1;
        s = (x >> 8) + 256;
    } else if (x <= (1u << 24)) {
        k = 2;
        s = (x >> 16) + 512;
    } else {
        k = 3;
        s = (x >> 24) + 768;
    }
    if (nbytes)
        *nbytes = k;
    return s;
}

/* ans_msb.hpp:156-176 */
void orc_msb_undo(uint32_t s, uint32_t* value, uint32_t* nbytes)
{
    if (s <= 256) {
        *value = s;
        *nbytes = 0;
    } else if (s <= 512) {
        *value = (s - 256) << 8;
        *nbytes = 1;
    } else if (s <= 768) {
        *value = (s - 512) << 16;
        *nbytes = 2;
    } else {
        *value = (s - 768) << 24;
        *nbytes = 3;
    }
}

/* -------------------------------------------------------------------- normalisation */

typedef struct {
    uint64_t freq;
    uint32_t sym;
} fs_pair;

static int cmp_fs(const void* a, const void* b)
{
    const fs_pair* x = (const fs_pair*)a;
    const fs_pair* y = (const fs_pair*)b;
    if (x->freq != y->freq)
        return x->freq < y->freq ? -1 : 1;
    return x->sym < y->sym ? -1 : (x->sym > y->sym);
}

/* util.hpp:271-282 */
static double orc_entropy(const uint64_t* freqs, size_t nbins, size_t freq_sum)
{
    double H0 = 0.0, n = (double)freq_sum;
    for (size_t i = 0; i < nbins; i++) {
        if (freqs[i] != 0) {
            double p = (double)freqs[i] / n;
            H0 = fma(p, log2(p), H0);
        }
    }
    return -H0;
}

/* util.hpp:284-298.  Both sums go through std::accumulate(..., 0): an `int`
 * accumulator, i.e. they wrap to a signed 32-bit value. */
static double orc_cross_entropy(const uint64_t* P, size_t np, const uint32_t* Q, size_t nq)
{
    int32_t accp = 0, accq = 0;
    for (size_t i = 0; i < np; i++)
        accp = (int32_t)(uint32_t)((uint64_t)(int64_t)accp + P[i]);
    for (size_t i = 0; i < nq; i++)
        accq = (int32_t)((uint32_t)accq + Q[i]);
    double n = (double)accp, m = (double)accq, H0 = 0.0;
    for (size_t i = 0; i < np; i++) {
        if (P[i] != 0 && i < nq && Q[i] != 0) {
            double p = (double)P[i] / n;
            double q = (double)Q[i] / m;
            H0 = fma(p, log2(q), H0);
        }
    }
    return -H0;
}

/* ans_util.hpp:77-95 */
static int orc_scale_freqs(uint32_t* S, const uint64_t* F, const uint32_t* order, int64_t M, size_t sigma,
    size_t freq_sum)
{
    for (size_t cur = 0; cur < sigma; cur++) {
        uint32_t sym = order[cur];
        double aratio = (double)M / (double)freq_sum;
        S[sym] = (uint32_t)fma(aratio, (double)F[sym], 0.5);
        if (S[sym] == 0)
            S[sym] = 1;
        M -= S[sym];
        freq_sum -= F[sym];
        if (M < 0)
            break;
    }
    return M != 0;
}

/* ans_util.hpp:65-74 */
static uint64_t next_pow2(uint64_t x)
{
    if (x == 0)
        return 1;
    return 1ULL << (64 - __builtin_clzll(x));
}
static int is_pow2(uint64_t x) { return x != 0 && !(x & (x - 1)); }

/* ans_util.hpp:100-157.  out: largest_sym+1 entries. */
int orc_adjust_freqs(const uint64_t* freqs, size_t nbins, uint32_t largest_sym, int require_u16,
    uint32_t H_approx, uint32_t* out)
{
    size_t sigma = 0, freq_sum = 0;
    for (size_t i = 0; i < nbins; i++) {
        freq_sum += freqs[i];
        sigma += (freqs[i] != 0);
    }
    size_t target = sigma;
    if (!is_pow2(target))
        target = next_pow2(target);
    fs_pair* sorted = (fs_pair*)malloc((sigma + 1) * sizeof(fs_pair));
    uint32_t* order = (uint32_t*)malloc((sigma + 1) * sizeof(uint32_t));
    size_t ns = (size_t)largest_sym + 1;
    uint32_t* scaled = (uint32_t*)calloc(ns, sizeof(uint32_t));
    uint32_t* prev = (uint32_t*)calloc(ns, sizeof(uint32_t));
    size_t c = 0;
    for (size_t i = 0; i < nbins; i++)
        if (freqs[i] != 0) {
            sorted[c].freq = freqs[i];
            sorted[c].sym = (uint32_t)i;
            c++;
        }
    qsort(sorted, sigma, sizeof(fs_pair), cmp_fs);
    for (size_t i = 0; i < sigma; i++)
        order[i] = sorted[i].sym;

    double H = orc_entropy(freqs, nbins, freq_sum);
    double threshold = H * (1.0 + (double)H_approx / 1000.0);
    for (;;) {
        if (orc_scale_freqs(scaled, freqs, order, (int64_t)target, sigma, freq_sum)) {
            target *= 2;
            continue;
        }
        uint32_t maxf = 0;
        for (size_t i = 0; i < ns; i++)
            if (scaled[i] > maxf)
                maxf = scaled[i];
        double XH = orc_cross_entropy(freqs, nbins, scaled, ns);
        if (require_u16 && maxf >= 65535u) {
            memcpy(scaled, prev, ns * sizeof(uint32_t));
            break;
        }
        if (XH < threshold)
            break;
        target *= 2;
        memcpy(prev, scaled, ns * sizeof(uint32_t));
    }
    memcpy(out, scaled, ns * sizeof(uint32_t));
    free(sorted);
    free(order);
    free(scaled);
    free(prev);
    return (int)ns;
}

/* ------------------------------------------------------------ prelude: vbyte + interp */

/* vbyte.hpp:57-80 */
static void vbyte_put(uint8_t** out, uint32_t x)
{
    while (x >= 128) {
        *(*out)++ = (uint8_t)((x & 127) | 128);
        x >>= 7;
    }
    *(*out)++ = (uint8_t)x;
}
/* vbyte.hpp:82-95 */
static uint32_t vbyte_get(const uint8_t** in)
{
    uint32_t x = 0, shift = 0;
    for (;;) {
        uint8_t c = *(*in)++;
        x += (uint32_t)(c & 127) << shift;
        if (!(c & 128))
            return x;
        shift += 7;
    }
}

/* bits.hpp:146-218 bit_stream: integers are packed LSB-first into consecutive 32-bit
 * words; the last word is written only if partly filled.  Its unused high bits are
 * whatever the reference's uninitialised 3-word buffer held; they are 0 here, and the
 * tests mask them when comparing with the reference. The destination may be unaligned
 * (ans_util.hpp:52), so words go through memcpy. */
typedef struct {
    uint8_t* p;
    uint64_t acc;
    unsigned nbits;
    size_t bytes;
} bitw;

static void bitw_put(bitw* w, uint32_t val, unsigned bits)
{
    if (bits == 0)
        return;
    if (bits < 32)
        val &= (1u << bits) - 1u;
    w->acc |= (uint64_t)val << w->nbits;
    w->nbits += bits;
    if (w->nbits >= 32) {
        uint32_t word = (uint32_t)w->acc;
        memcpy(w->p, &word, 4);
        w->p += 4;
        w->bytes += 4;
        w->acc >>= 32;
        w->nbits -= 32;
    }
}
static size_t bitw_flush(bitw* w)
{
    if (w->nbits != 0) {
        uint32_t word = (uint32_t)w->acc;
        memcpy(w->p, &word, 4);
        w->p += 4;
        w->bytes += 4;
        w->nbits = 0;
        w->acc = 0;
    }
    return w->bytes;
}

typedef struct {
    const uint8_t* p;
    uint64_t acc;
    unsigned nbits;
    size_t words;
} bitr;

static uint32_t bitr_get(bitr* r, unsigned bits)
{
    if (bits == 0)
        return 0;
    if (r->nbits < bits) {
        uint32_t word;
        memcpy(&word, r->p, 4);
        r->p += 4;
        r->words++;
        r->acc |= (uint64_t)word << r->nbits;
        r->nbits += 32;
    }
    uint32_t v = (uint32_t)(r->acc & ((bits == 32) ? 0xFFFFFFFFull : ((1ull << bits) - 1)));
    r->acc >>= bits;
    r->nbits -= bits;
    return v;
}

static unsigned hi_bit(uint64_t x) { return x ? 63u - (unsigned)__builtin_clzll(x) : 0u; } /* bits.hpp:33 */

/* interp.hpp:28-47: centred minimal binary code of val in [1,u] */
static void write_center_mid(bitw* w, uint64_t val, uint64_t u)
{
    if (u == 1)
        return;
    uint64_t b = hi_bit(u - 1) + 1;
    uint64_t d = 2 * u - (1ull << b);
    val = val + (u - (d >> 1));
    if (val > u)
        val -= u;
    uint32_t m = (uint32_t)((1ull << b) - u);
    if (val <= m) {
        bitw_put(w, (uint32_t)(val - 1), (unsigned)(b - 1));
    } else {
        val += m;
        bitw_put(w, (uint32_t)((val - 1) >> 1), (unsigned)(b - 1));
        bitw_put(w, (uint32_t)((val - 1) & 1), 1);
    }
}

/* interp.hpp:48-65 */
static uint64_t read_center_mid(bitr* r, uint64_t u)
{
    uint64_t b = (u == 1) ? 0 : hi_bit(u - 1) + 1;
    uint64_t d = 2 * u - (1ull << b);
    uint64_t val = 1;
    if (u != 1) {
        uint64_t m = (1ull << b) - u;
        val = (uint64_t)bitr_get(r, (unsigned)(b - 1)) + 1;
        if (val > m)
            val = (2 * val + bitr_get(r, 1)) - m - 1;
    }
    val = val + (d >> 1);
    if (val > u)
        val -= u;
    return val;
}

/* interp.hpp:67-82 */
static void enc_interp(bitw* w, const uint32_t* in, size_t n, size_t low, size_t high)
{
    if (n == 0)
        return;
    uint64_t h = (n + 1) >> 1, n1 = h - 1, n2 = n - h;
    uint64_t v = (uint64_t)in[h - 1] + 1;
    write_center_mid(w, v - low - n1 + 1, high - n2 - low - n1 + 1);
    enc_interp(w, in, n1, low, v - 1);
    enc_interp(w, in + h, n2, v + 1, high);
}

/* interp.hpp:84-98 */
static void dec_interp(bitr* r, uint32_t* out, size_t n, size_t low, size_t high)
{
    if (n == 0)
        return;
    uint64_t h = (n + 1) >> 1, n1 = h - 1, n2 = n - h;
    uint64_t v = low + n1 - 1 + read_center_mid(r, high - n2 - low - n1 + 1);
    out[h - 1] = (uint32_t)(v - 1);
    if (n1)
        dec_interp(r, out, n1, low, v - 1);
    if (n2)
        dec_interp(r, out + h, n2, v + 1, high);
}

/* interp.hpp:100-108 */
long orc_interp_encode(uint32_t* dst, const uint32_t* in, size_t n, size_t u)
{
    bitw w = { (uint8_t*)dst, 0, 0, 0 };
    enc_interp(&w, in, n, 1, u + 1);
    return (long)bitw_flush(&w);
}

/* interp.hpp:110-118 */
long orc_interp_decode(const uint32_t* src, uint32_t* out, size_t n, size_t u)
{
    bitr r = { (const uint8_t*)src, 0, 0, 0 };
    dec_interp(&r, out, n, 1, u + 1);
    return (long)r.words;
}

/* ans_util.hpp:46-63 */
long orc_serialize_interp(const uint32_t* nfreqs, size_t count, uint64_t frame_size, uint8_t* out)
{
    uint8_t* p = out;
    uint32_t max_sym = (uint32_t)(count - 1);
    vbyte_put(&p, max_sym);
    *p++ = (uint8_t)log2((double)frame_size);
    uint32_t* inc = (uint32_t*)malloc(count * sizeof(uint32_t));
    inc[0] = nfreqs[0];
    for (size_t s = 1; s <= max_sym; s++)
        inc[s] = inc[s - 1] + nfreqs[s] + 1;
    bitw w = { p, 0, 0, 0 };
    enc_interp(&w, inc, count, 1, frame_size + count + 1 + 1);
    p += bitw_flush(&w);
    free(inc);
    return p - out;
}

/* ans_util.hpp:25-42 ; out may be NULL to query the count */
long orc_load_interp(const uint8_t* in, uint32_t* out)
{
    const uint8_t* p = in;
    uint32_t max_sym = vbyte_get(&p);
    uint64_t frame_size = 1ull << (*p++);
    size_t count = (size_t)max_sym + 1;
    if (!out)
        return (long)count;
    bitr r = { p, 0, 0, 0 };
    dec_interp(&r, out, count, 1, frame_size + count + 1 + 1);
    uint32_t prev = out[0];
    for (size_t s = 1; s <= max_sym; s++) {
        uint32_t cur = out[s];
        out[s] = cur - prev - 1;
        prev = cur;
    }
    return (long)count;
}

/* ans_byte.hpp:41-97 -- the byte coder's own normaliser: frame = min(4096, 64 x distinct symbols) rounded up to a power
 * of two; truncated shares with a fudge factor that shrinks by 0.01 per try until they fit; the slack goes out from the
 * top symbol down in proportion, the rest to the largest share. */
void orc_byte_adjust_freqs(const uint64_t* freqs, uint16_t* adj)
{
    size_t uniq = 0, initial_sum = 0;
    for (size_t i = 0; i < 256; i++) {
        initial_sum += freqs[i];
        uniq += freqs[i] != 0;
        adj[i] = 0;
    }
    size_t target = uniq * 64; /* :52-58 */
    if (target > 4096)
        target = 4096;
    if (!is_pow2(target))
        target = (size_t)next_pow2(target);
    double c = (double)target / (double)initial_sum;
    size_t cur = (size_t)-1;
    double fudge = 1.0;
    while (cur > target) { /* :63-72 */
        fudge -= 0.01;
        cur = 0;
        for (size_t sym = 0; sym < 256; sym++) {
            adj[sym] = (uint16_t)(fudge * (double)freqs[sym] * c);
            if (adj[sym] == 0 && freqs[sym] != 0)
                adj[sym] = 1;
            cur += adj[sym];
        }
    }
    int64_t excess = (int64_t)target - (int64_t)cur; /* :73-86 */
    for (size_t sym = 0; sym < 256; sym++) {
        uint16_t ncnt = adj[255 - sym];
        if (ncnt == 0)
            continue;
        double ratio = (double)excess / (double)cur;
        size_t adder = (size_t)(ratio * (double)ncnt);
        if (adder > (size_t)excess)
            adder = (size_t)excess;
        excess -= (int64_t)adder;
        cur -= ncnt;
        adj[255 - sym] = (uint16_t)(adj[255 - sym] + adder);
    }
    if (excess != 0) { /* :87-97 */
        size_t max_freq = 0, max_sym = 0;
        for (size_t sym = 0; sym < 256; sym++)
            if (adj[sym] > max_freq) {
                max_freq = adj[sym];
                max_sym = sym;
            }
        adj[max_sym] = (uint16_t)(adj[max_sym] + excess);
    }
}

/* ---------------------------------------------------------------------------- models */

struct orc_model {
    int method;
    uint32_t f;
    uint32_t* nfreqs; /* max_sym + 1 entries */
    size_t nsyms;
    uint64_t M, L;
    uint32_t logM;
    /* encoder side */
    uint32_t* base;
    uint32_t* mapping; /* rfold: value -> reordered value (ans_reorder_fold.hpp:92-106) */
    size_t mapping_n;
    uint32_t* most_frequent;
    size_t n_mf;
    uint64_t sigma;
    /* decoder side: one entry per slot */
    uint32_t* d_freq;
    uint32_t* d_off;
    uint32_t* d_val; /* sym (int) or unfolded value | nbytes<<30 */
};

static size_t nbins_of(int method, uint32_t f, uint32_t max_value)
{
    switch (method) {
    case ORC_INT: return (size_t)max_value + 1; /* ans_int.hpp:46 */
    case ORC_BYTE: return 256;                   /* ans_byte.hpp:25 */
    case ORC_MSB: return MSB_MAX_SIGMA;          /* ans_msb.hpp:79 */
    default: return (size_t)1 << (f + 9);        /* ans_fold.hpp:70 */
    }
}

static inline uint32_t map_symbol(const orc_model* m, uint32_t x, uint32_t* k)
{
    switch (m->method) {
    case ORC_INT:
    case ORC_BYTE: *k = 0; return x;
    case ORC_MSB: return orc_msb_mapping(x, k);
    case ORC_FOLD: return orc_fold_mapping(m->f, x, k);
    default: return orc_fold_mapping(m->f, m->mapping[x], k); /* ans_reorder_fold.hpp:158-159 */
    }
}

typedef struct {
    uint64_t count;
    uint32_t sym;
} cs_pair;
static int cmp_cs(const void* a, const void* b) /* (-count, sym) ascending, ans_reorder_fold.hpp:79-85 */
{
    const cs_pair* x = (const cs_pair*)a;
    const cs_pair* y = (const cs_pair*)b;
    if (x->count != y->count)
        return x->count > y->count ? -1 : 1;
    return x->sym < y->sym ? -1 : (x->sym > y->sym);
}

static void finish_encoder(orc_model* m)
{
    uint64_t sum = 0;
    for (size_t s = 0; s < m->nsyms; s++)
        sum += m->nfreqs[s];
    m->M = sum; /* ans_fold.hpp:80-81 */
    m->L = ANS_K * m->M;
    m->logM = (uint32_t)hi_bit(m->M);
    m->base = (uint32_t*)malloc(m->nsyms * sizeof(uint32_t));
    uint64_t cur = 0;
    for (size_t s = 0; s < m->nsyms; s++) { /* ans_fold.hpp:84-90 */
        m->base[s] = (uint32_t)cur;
        cur += m->nfreqs[s];
    }
}

/* ans_fold.hpp:68-94 ; ans_int.hpp:40-68 ; ans_msb.hpp:76-101 ; ans_reorder_fold.hpp:69-130 */
orc_model* orc_model_create(int method, uint32_t f, const uint32_t* in, size_t n)
{
    orc_model* m = (orc_model*)calloc(1, sizeof(orc_model));
    m->method = method;
    m->f = f;
    uint32_t max_value = 0;
    if (method == ORC_INT || method == ORC_RFOLD)
        for (size_t i = 0; i < n; i++)
            if (in[i] > max_value)
                max_value = in[i];
    if (method == ORC_RFOLD) {
        size_t nv = (size_t)max_value + 1;
        uint64_t* cnt = (uint64_t*)calloc(nv, sizeof(uint64_t));
        for (size_t i = 0; i < n; i++)
            cnt[in[i]]++;
        size_t sigma = 0;
        for (size_t v = 0; v < nv; v++)
            sigma += cnt[v] != 0;
        cs_pair* pr = (cs_pair*)malloc((sigma + 1) * sizeof(cs_pair));
        size_t c = 0;
        for (size_t v = 0; v < nv; v++)
            if (cnt[v]) {
                pr[c].count = cnt[v];
                pr[c].sym = (uint32_t)v;
                c++;
            }
        qsort(pr, sigma, sizeof(cs_pair), cmp_cs);
        m->sigma = sigma;
        size_t thres = (size_t)1 << (f + 7);
        m->mapping = (uint32_t*)malloc(nv * sizeof(uint32_t));
        m->mapping_n = nv;
        if (sigma < thres) {
            for (size_t v = 0; v < nv; v++)
                m->mapping[v] = (uint32_t)v;
        } else {
            for (size_t v = 0; v < nv; v++)
                m->mapping[v] = (uint32_t)(v + thres);
            m->most_frequent = (uint32_t*)malloc(thres * sizeof(uint32_t));
            m->n_mf = thres;
            for (size_t i = 0; i < thres; i++) {
                m->mapping[pr[i].sym] = (uint32_t)i;
                m->most_frequent[i] = pr[i].sym;
            }
        }
        free(pr);
        free(cnt);
    }
    size_t nbins = nbins_of(method, f, max_value);
    uint64_t* freqs = (uint64_t*)calloc(nbins, sizeof(uint64_t));
    uint32_t max_sym = 0;
    for (size_t i = 0; i < n; i++) {
        uint32_t k, s = map_symbol(m, in[i], &k);
        freqs[s]++;
        if (s > max_sym)
            max_sym = s;
    }
    if (method == ORC_BYTE) { /* ans_byte.hpp:99-120: all 256 entries, the byte normaliser */
        uint16_t nf[256];
        orc_byte_adjust_freqs(freqs, nf);
        m->nsyms = 256;
        m->nfreqs = (uint32_t*)calloc(256, sizeof(uint32_t));
        for (size_t s = 0; s < 256; s++)
            m->nfreqs[s] = nf[s];
        free(freqs);
        finish_encoder(m);
        return m;
    }
    m->nsyms = (size_t)max_sym + 1;
    m->nfreqs = (uint32_t*)calloc(m->nsyms, sizeof(uint32_t));
    orc_adjust_freqs(freqs, nbins, max_sym, method != ORC_INT, 1, m->nfreqs);
    free(freqs);
    finish_encoder(m);
    return m;
}

long orc_model_info(orc_model* m, uint64_t* frame_size, uint32_t* nfreqs_out)
{
    *frame_size = m->M;
    if (nfreqs_out)
        memcpy(nfreqs_out, m->nfreqs, m->nsyms * sizeof(uint32_t));
    return (long)m->nsyms;
}

long orc_model_rfold_info(orc_model* m, uint64_t* sigma, uint32_t* mf)
{
    if (m->method != ORC_RFOLD)
        return -1;
    *sigma = m->sigma;
    if (mf && m->n_mf)
        memcpy(mf, m->most_frequent, m->n_mf * sizeof(uint32_t));
    return (long)m->n_mf;
}

/* ans_fold.hpp:96-99 ; ans_reorder_fold.hpp:132-154 */
long orc_model_serialize(orc_model* m, uint8_t* out)
{
    uint8_t* p = out;
    if (m->method == ORC_RFOLD) {
        uint32_t flag = m->n_mf ? 1u : 0u;
        memcpy(p, &flag, 4);
        p += 4;
        if (flag) {
            memcpy(p, m->most_frequent, m->n_mf * 4);
            p += m->n_mf * 4;
        }
    }
    if (m->method == ORC_BYTE) { /* ans_byte.hpp:122-137: running sums of (nfreq + 1), universe 4096 + 256, no header */
        uint32_t inc[256];
        inc[0] = m->nfreqs[0];
        for (size_t s = 1; s < 256; s++)
            inc[s] = inc[s - 1] + m->nfreqs[s] + 1;
        long ib = orc_interp_encode((uint32_t*)p, inc, 256, 4096 + 256);
        return (p - out) + ib;
    }
    p += orc_serialize_interp(m->nfreqs, m->nsyms, m->M, p);
    return p - out;
}

/* ans_fold.hpp:100-112 (ans_int.hpp:74-84, ans_msb.hpp:107-118, ans_reorder_fold.hpp:156-168):
 * k raw low bytes first, then the optional renormalisation word, then the state update. */
static inline void encode_symbol(const orc_model* m, uint64_t* state, uint32_t x, uint8_t** out)
{
    uint32_t k, s = map_symbol(m, x, &k);
    uint32_t v = (m->method == ORC_RFOLD) ? m->mapping[x] : x;
    for (uint32_t i = 0; i < k; i++) {
        *(*out)++ = (uint8_t)(v & 0xFF);
        v >>= 8;
    }
    uint64_t freq = m->nfreqs[s];
    if (*state >= ((ANS_K << ANS_RADIX_LOG2) * freq)) {
        uint32_t w = (uint32_t)(*state & 0xFFFFFFFFull);
        memcpy(*out, &w, 4);
        *out += 4;
        *state >>= ANS_RADIX_LOG2;
    }
    *state = (*state / freq) * m->M + (*state % freq) + m->base[s];
}

/* the 4-state loop of ans_fold.hpp:250-278 on a sub-range */
long orc_encode_block(orc_model* m, const uint32_t* in, size_t n, uint8_t* dst)
{
    uint8_t* out = dst;
    uint64_t st[4] = { m->L, m->L, m->L, m->L };
    size_t cur = 0;
    while ((n - cur) % 4 != 0) {
        encode_symbol(m, &st[0], in[n - cur - 1], &out);
        cur++;
    }
    while (cur != n) {
        encode_symbol(m, &st[0], in[n - cur - 1], &out);
        encode_symbol(m, &st[1], in[n - cur - 2], &out);
        encode_symbol(m, &st[2], in[n - cur - 3], &out);
        encode_symbol(m, &st[3], in[n - cur - 4], &out);
        cur += 4;
    }
    for (int i = 0; i < 4; i++) { /* ans_fold.hpp:115-120 */
        uint64_t v = st[i] - m->L;
        memcpy(out, &v, 8);
        out += 8;
    }
    return out - dst;
}

/* ans_fold.hpp:179-207 ; ans_int.hpp:116-164 ; ans_msb.hpp:179-204 ; ans_reorder_fold.hpp:235-279 */
orc_model* orc_model_load(int method, uint32_t f, const uint8_t* prelude)
{
    orc_model* m = (orc_model*)calloc(1, sizeof(orc_model));
    m->method = method;
    m->f = f;
    const uint8_t* p = prelude;
    size_t thres = (size_t)1 << (f + 7);
    if (method == ORC_RFOLD) {
        uint32_t flag;
        memcpy(&flag, p, 4);
        p += 4;
        m->most_frequent = (uint32_t*)malloc(thres * sizeof(uint32_t));
        m->n_mf = thres;
        if (flag == 1) {
            memcpy(m->most_frequent, p, thres * 4);
            p += thres * 4;
        } else {
            for (size_t i = 0; i < thres; i++)
                m->most_frequent[i] = (uint32_t)i;
        }
    }
    if (method == ORC_BYTE) { /* ans_byte.hpp:178-190 */
        m->nsyms = 256;
        m->nfreqs = (uint32_t*)malloc(256 * sizeof(uint32_t));
        orc_interp_decode((const uint32_t*)p, m->nfreqs, 256, 4096 + 256);
        uint32_t prev = m->nfreqs[0];
        for (size_t s = 1; s < 256; s++) {
            uint32_t cur = m->nfreqs[s];
            m->nfreqs[s] = cur - prev - 1;
            prev = cur;
        }
    } else {
        m->nsyms = (size_t)orc_load_interp(p, NULL);
        m->nfreqs = (uint32_t*)malloc(m->nsyms * sizeof(uint32_t));
        orc_load_interp(p, m->nfreqs);
    }
    finish_encoder(m);
    m->d_freq = (uint32_t*)malloc(m->M * sizeof(uint32_t));
    m->d_off = (uint32_t*)malloc(m->M * sizeof(uint32_t));
    m->d_val = (uint32_t*)malloc(m->M * sizeof(uint32_t));
    for (size_t s = 0; s < m->nsyms; s++) {
        uint32_t val = (uint32_t)s, k = 0;
        if (method == ORC_FOLD) {
            orc_fold_undo(f, (uint32_t)s, &val, &k);
        } else if (method == ORC_MSB) {
            orc_msb_undo((uint32_t)s, &val, &k);
        } else if (method == ORC_RFOLD) {
            if (s < thres) {
                val = m->most_frequent[s] + (uint32_t)thres; /* ans_reorder_fold.hpp:214-215 */
                k = 0;
            } else
                orc_fold_undo(f, (uint32_t)s, &val, &k);
        }
        if (method != ORC_INT && method != ORC_BYTE)
            val += k << 30;
        for (uint32_t j = 0; j < m->nfreqs[s]; j++) {
            size_t slot = m->base[s] + j;
            m->d_freq[slot] = m->nfreqs[s];
            m->d_off[slot] = j;
            m->d_val[slot] = val;
        }
    }
    return m;
}

/* ans_fold.hpp:216-228 + :135-147 (ans_int.hpp:173-186): renorm word first, then k raw bytes */
static inline uint32_t decode_sym(const orc_model* m, uint64_t* state, const uint8_t** in)
{
    size_t slot = (size_t)(*state & (m->M - 1));
    uint32_t val = m->d_val[slot];
    *state = (uint64_t)m->d_freq[slot] * (*state >> m->logM) + m->d_off[slot];
    if (*state < m->L) {
        uint32_t w;
        *in -= 4;
        memcpy(&w, *in, 4);
        *state = (*state << ANS_RADIX_LOG2) | w;
    }
    if (m->method == ORC_INT || m->method == ORC_BYTE)
        return val;
    static const uint32_t mask[4] = { 0x0, 0xFF, 0xFFFF, 0xFFFFFF };
    uint32_t k = val >> 30, w;
    memcpy(&w, *in - k, 4);
    uint32_t num = (val & 0x3FFFFFFFu) + (w & mask[k]);
    *in -= k;
    if (m->method == ORC_RFOLD)
        num -= 1u << (m->f + 7); /* ans_reorder_fold.hpp:301 */
    return num;
}

/* the loop of ans_fold.hpp:291-310 on one block payload */
int orc_decode_block(orc_model* m, const uint8_t* src, size_t bytes, uint32_t* out, size_t n)
{
    const uint8_t* in = src + bytes;
    uint64_t st[4];
    for (int i = 3; i >= 0; i--) { /* ans_fold.hpp:209-214, 292-295 */
        uint64_t v;
        in -= 8;
        memcpy(&v, in, 8);
        st[i] = v + m->L;
    }
    size_t i = 0, fast = n - (n % 4);
    while (i != fast) {
        out[i] = decode_sym(m, &st[3], &in);
        out[i + 1] = decode_sym(m, &st[2], &in);
        out[i + 2] = decode_sym(m, &st[1], &in);
        out[i + 3] = decode_sym(m, &st[0], &in);
        i += 4;
    }
    while (i != n)
        out[i++] = decode_sym(m, &st[0], &in);
    return 0;
}

void orc_model_free(orc_model* m)
{
    if (!m)
        return;
    free(m->nfreqs);
    free(m->base);
    free(m->mapping);
    free(m->most_frequent);
    free(m->d_freq);
    free(m->d_off);
    free(m->d_val);
    free(m);
}

/* ans_X_compress (ans_fold.hpp:238-281 etc.) */
long orc_compress(int method, uint32_t f, uint8_t* dst, size_t cap, const uint32_t* src, size_t n)
{
    (void)cap; /* the reference ignores dstCapacity too (ans_fold.hpp:239-240) */
    orc_model* m = orc_model_create(method, f, src, n);
    uint8_t* p = dst;
    p += orc_model_serialize(m, p);
    p += orc_encode_block(m, src, n, p);
    orc_model_free(m);
    return p - dst;
}

/* ans_X_decompress (ans_fold.hpp:283-311 etc.) */
int orc_decompress(int method, uint32_t f, uint32_t* dst, size_t n, const uint8_t* src, size_t bytes)
{
    orc_model* m = orc_model_load(method, f, src);
    orc_decode_block(m, src, bytes, dst, n);
    orc_model_free(m);
    return 0;
}

static double now_ns(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec * 1e9 + (double)ts.tv_nsec;
}

/* table_efficiency.cpp:77-101 style: min over `runs`, single thread */
/* ---- pseudo-adaptive driver with the dense-rank alphabet remap, one block (src/pseudo_adaptive.cpp:83-125) ----
 * The driver marks the values present in the block, replaces every value by its 1-based rank among them (:86-99),
 * turns the sorted list of present values into running sums held in uint32_t (:101-103), writes
 * [u32 sigma][u32 last sum + 1][interpolative code of the sums, universe last sum + 1] (:105-112) and, unless the
 * block holds one distinct value, appends T::encode of the ranks (:114-121).  Sums that leave 32 bits are outside
 * the driver's working domain (the list handed to the interpolative coder would no longer be increasing): -2. */
static int cmp_u32(const void* a, const void* b)
{
    const uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b;
    return x < y ? -1 : x > y;
}

long orc_pa_remap_block(int method, uint32_t f, const uint32_t* in, size_t n, uint8_t* dst, size_t cap,
    uint32_t* remapped_out, size_t* alphabet_bytes)
{
    if (n == 0 || !in || !dst)
        return -1;
    uint32_t* present = (uint32_t*)malloc(n * sizeof(uint32_t));
    uint32_t* ranks = remapped_out ? remapped_out : (uint32_t*)malloc(n * sizeof(uint32_t));
    long result = -1;
    memcpy(present, in, n * sizeof(uint32_t));
    qsort(present, n, sizeof(uint32_t), cmp_u32);
    size_t sigma = 1;
    for (size_t i = 1; i < n; i++)
        if (present[i] != present[sigma - 1])
            present[sigma++] = present[i];
    for (size_t k = 0; k < n; k++) { /* rank = number of present values <= in[k] */
        size_t lo = 0, hi = sigma;
        while (hi - lo > 1) {
            const size_t mid = (lo + hi) / 2;
            if (present[mid] <= in[k])
                lo = mid;
            else
                hi = mid;
        }
        ranks[k] = (uint32_t)(lo + 1);
    }
    uint64_t run = 0;
    for (size_t i = 0; i < sigma; i++) {
        run += present[i];
        if (run > 0xFFFFFFFEull) {
            result = -2;
            goto done;
        }
        present[i] = (uint32_t)run;
    }
    if (cap < 8 + 8 * sigma + 64) {
        result = -3;
        goto done;
    }
    {
        const uint32_t head[2] = { (uint32_t)sigma, present[sigma - 1] + 1u };
        memcpy(dst, head, 8);
        const long ib = orc_interp_encode((uint32_t*)(dst + 8), present, sigma, (size_t)present[sigma - 1] + 1u);
        size_t at = 8 + (size_t)ib;
        if (alphabet_bytes)
            *alphabet_bytes = at;
        if (sigma != 1) {
            const long eb = orc_compress(method, f, dst + at, cap - at, ranks, n);
            if (eb < 0)
                goto done;
            at += (size_t)eb;
        }
        result = (long)at;
    }
done:
    free(present);
    if (!remapped_out)
        free(ranks);
    return result;
}

int orc_time(int method, uint32_t f, const uint32_t* in, size_t n, int runs, double* enc_ns, double* dec_ns,
    size_t* bytes)
{
    uint8_t* out = (uint8_t*)malloc(n * 8 + 4096);
    uint32_t* rec = (uint32_t*)malloc((n + 16) * sizeof(uint32_t));
    double be = 1e300, bd = 1e300;
    long nb = 0;
    for (int r = 0; r < runs; r++) {
        double t0 = now_ns();
        nb = orc_compress(method, f, out, n * 8 + 4096, in, n);
        double t1 = now_ns();
        if (t1 - t0 < be)
            be = t1 - t0;
    }
    for (int r = 0; r < runs; r++) {
        double t0 = now_ns();
        orc_decompress(method, f, rec, n, out, (size_t)nb);
        double t1 = now_ns();
        if (t1 - t0 < bd)
            bd = t1 - t0;
    }
    int ok = memcmp(rec, in, n * sizeof(uint32_t)) == 0;
    free(out);
    free(rec);
    *enc_ns = be;
    *dec_ns = bd;
    *bytes = (size_t)nb;
    return ok ? 0 : -2;
}
