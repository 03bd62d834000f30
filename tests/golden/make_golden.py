#!/usr/bin/env python
"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libansref.so, built by
oracle/Makefile from /root/reference/include).  Run in the build container only:

    make -C oracle && python tests/golden/make_golden.py

Each vector holds a small seeded input and what the reference produced for it:
  nfreqs / frame        ans_X_encode::create  -> adjust_freqs          (ans_util.hpp:100-157)
  prelude               ans_X_encode::serialize                        (ans_util.hpp:46-63, ans_reorder_fold.hpp:132-154)
  stream                ans_X_compress whole stream                    (ans_fold.hpp:238-281 ...)
  block_payloads        the 4-state loop of ans_fold.hpp:250-278 driven per 1024-symbol block with the global model
plus known answers of the fold / msb mappings (ans_fold.hpp:38-65, ans_msb.hpp:41-73) and the
interpolative coder (interp.hpp:100-118), and blocks of the pseudo-adaptive driver with its dense-rank alphabet
remap (src/pseudo_adaptive.cpp:83-125; `python tests/golden/make_golden.py pa_remap` writes only those), and ans_byte
(ans_byte.hpp; `... make_golden.py ans_byte`).  The reference leaves the unused high bits of its last
interp word uninitialised (bits.hpp:148); they are masked to zero here so the vectors are stable.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _inputs  # noqa: E402
from _libs import M_FOLD, M_INT, M_MSB, M_RFOLD, ptr, reference, u32p  # noqa: E402
import ctypes as C  # noqa: E402

BLOCK = 1024
CASES = [
    # name, method, f, input kind, n, seed
    ("fold1_zipf20", M_FOLD, 1, "zipf20", 20011, 1),
    ("fold2_zipf20", M_FOLD, 2, "zipf20", 20011, 2),
    ("fold1_wide", M_FOLD, 1, "wide", 9001, 3),
    ("fold5_wide", M_FOLD, 5, "wide", 9001, 4),
    ("fold8_zipf20", M_FOLD, 8, "zipf20", 9001, 5),
    ("msb_zipf20", M_MSB, 0, "zipf20", 20011, 6),
    ("msb_wide", M_MSB, 0, "wide", 9001, 7),
    ("int_geom", M_INT, 0, "geom", 20011, 8),
    ("int_zipf16", M_INT, 0, "zipf16", 20011, 9),
    ("rfold1_zipf20", M_RFOLD, 1, "zipf20", 20011, 10),
    ("rfold2_zipf20", M_RFOLD, 2, "zipf20", 20011, 11),
    ("fold1_tiny5", M_FOLD, 1, "zipf16", 5, 12),
    ("int_tiny7", M_INT, 0, "zipf16", 7, 13),
    ("fold1_skew", M_FOLD, 1, "zipf12s", 20011, 14),
]


def mask_tail(prelude):
    """zero the undefined padding bits: keep only bits below the highest bit the encoder set
    deliberately is impossible to know, so instead the last interp word is stored masked to the
    bits that a clean (zero-initialised) writer produces -- see _container.prelude_equal."""
    return prelude


PA_REMAP_CASES = [
    # method, f, input kind, block length, seed   (one block of src/pseudo_adaptive.cpp:83-125 each)
    (M_INT, 0, "zipf20", 8192, 31), (M_MSB, 0, "zipf20", 8192, 32), (M_INT, 0, "geom", 4096, 33),
    (M_MSB, 0, "zipf16", 1024, 34), (M_INT, 0, "u8", 777, 35), (M_MSB, 0, "zipf12s", 16384, 36),
]


def pa_remap_vectors(ref):
    """the driver's dense-rank remap + block alphabet + T::encode, block by block"""
    out = {}
    for j, (method, f, kind, n, seed) in enumerate(PA_REMAP_CASES):
        data = _inputs.make(kind, n, seed=seed)
        blob, alpha_bytes, ranks = ref.pa_remap_block(method, f, data)
        coder_prelude = ref.model_create(method, f, ranks).serialize().size
        assert np.array_equal(ref.decompress(method, f, blob[alpha_bytes:], n), ranks)
        out[f"data{j}"], out[f"blob{j}"], out[f"ranks{j}"] = data, blob, ranks
        out[f"meta{j}"] = np.array([method, f, alpha_bytes, coder_prelude], dtype=np.int64)
    single = np.full(1000, 12345, dtype=np.uint32)  # one distinct value: alphabet only (pseudo_adaptive.cpp:114)
    blob, alpha_bytes, ranks = ref.pa_remap_block(M_INT, 0, single)
    assert blob.size == alpha_bytes
    j = len(PA_REMAP_CASES)
    out[f"data{j}"], out[f"blob{j}"], out[f"ranks{j}"] = single, blob, ranks
    out[f"meta{j}"] = np.array([M_INT, 0, alpha_bytes, 0], dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "pa_remap.npz"), count=np.int32(j + 1), **out)
    print(f"pa_remap: {j + 1} blocks")


def ans_byte_vectors(ref):
    """ans_byte (ans_byte.hpp): byte inputs, the byte normaliser's frequencies, prelude, whole stream, 1024-symbol blocks"""
    M_BYTE = 4
    rng = np.random.default_rng(77)
    cases = [
        ("uniform", rng.integers(0, 256, 20011)), ("geom", np.minimum(rng.geometric(0.04, 20011) - 1, 255)),
        ("vbyte", np.where(rng.random(9001) < 0.9, 128 + rng.geometric(0.1, 9001) % 128, rng.integers(0, 128, 9001))),
        ("few", rng.integers(3, 8, 5003)), ("one", np.full(777, 77)), ("tiny", rng.integers(0, 256, 5)),
    ]
    out = {}
    for j, (name, data) in enumerate(cases):
        data = data.astype(np.uint32)
        model = ref.model_create(M_BYTE, 0, data)
        frame, nfreqs = model.info()
        stream = ref.compress(M_BYTE, 0, data)
        assert np.array_equal(ref.decompress(M_BYTE, 0, stream, data.size), data)
        payloads = [model.encode_block(data[b : b + BLOCK]) for b in range(0, data.size, BLOCK)]
        out[f"data{j}"], out[f"nfreqs{j}"], out[f"frame{j}"] = data.astype(np.uint8), nfreqs.astype(np.uint16), np.uint64(frame)
        out[f"prelude{j}"], out[f"stream{j}"] = model.serialize(), stream
        out[f"block_sizes{j}"] = np.array([p.size for p in payloads], dtype=np.uint32)
        out[f"block_payloads{j}"] = np.concatenate(payloads)
        print(f"ans_byte {name}: n={data.size} frame={frame} stream={stream.size}B")
    np.savez_compressed(os.path.join(HERE, "ans_byte.npz"), count=np.int32(len(cases)), **out)


def main():
    ref = reference()
    if len(sys.argv) > 1 and sys.argv[1] == "pa_remap":
        pa_remap_vectors(ref)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "ans_byte":
        ans_byte_vectors(ref)
        return
    for name, method, f, kind, n, seed in CASES:
        data = _inputs.make(kind, n, seed=seed)
        model = ref.model_create(method, f, data)
        frame, nfreqs = model.info()
        prelude = model.serialize()
        stream = ref.compress(method, f, data)
        payloads = [model.encode_block(data[b : b + BLOCK]) for b in range(0, n, BLOCK)]
        sizes = np.array([p.size for p in payloads], dtype=np.uint32)
        extra = {}
        if method == M_RFOLD:
            sigma, mf = model.rfold_info()
            extra = dict(rfold_sigma=np.uint64(sigma), most_frequent=mf)
        # sanity: the reference decodes its own output
        assert np.array_equal(ref.decompress(method, f, stream, n), data)
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"), data=data, method=np.int32(method), fidelity=np.uint32(f),
            frame=np.uint64(frame), nfreqs=nfreqs, prelude=prelude, stream=stream, block_syms=np.uint32(BLOCK),
            block_sizes=sizes, block_payloads=np.concatenate(payloads), **extra)
        print(f"{name}: n={n} frame={frame} prelude={prelude.size}B stream={stream.size}B blocks={len(payloads)}")

    # mapping known answers: every boundary +-1 and a seeded sample, for f = 1..8 and msb
    rng = np.random.default_rng(99)
    xs = [0, 1, 2]
    for b in range(1, 31):
        xs += [(1 << b) - 1, 1 << b, (1 << b) + 1]
    for f in range(1, 9):
        for j in range(4):
            t = (1 << (f + 7)) << (8 * j)
            xs += [t - 1, t, t + 1]
    xs = np.array(sorted(set(x for x in xs if x < (1 << 30))) + list(rng.integers(0, 1 << 30, 4000)), dtype=np.uint32)
    L = ref.lib
    fold = np.zeros((8, xs.size), dtype=np.uint32)
    undo_v = np.zeros_like(fold)
    undo_k = np.zeros_like(fold)
    for f in range(1, 9):
        for i, x in enumerate(xs):
            m = C.c_uint32()
            L.ref_fold_mapping(f, int(x), C.byref(m))
            fold[f - 1, i] = m.value
            v, k = C.c_uint32(), C.c_uint32()
            L.ref_fold_undo(f, m.value, C.byref(v), C.byref(k))
            undo_v[f - 1, i], undo_k[f - 1, i] = v.value, k.value
    msb = np.array([L.ref_msb_mapping(int(x)) for x in xs], dtype=np.uint32)
    msb_v = np.zeros(xs.size, dtype=np.uint32)
    msb_k = np.zeros(xs.size, dtype=np.uint32)
    for i, m in enumerate(msb):
        v, k = C.c_uint32(), C.c_uint32()
        L.ref_msb_undo(int(m), C.byref(v), C.byref(k))
        msb_v[i], msb_k[i] = v.value, k.value
    np.savez_compressed(os.path.join(HERE, "mappings.npz"), x=xs, fold=fold, fold_undo_value=undo_v,
                        fold_undo_bytes=undo_k, msb=msb, msb_undo_value=msb_v, msb_undo_bytes=msb_k)
    print(f"mappings: {xs.size} values x (8 fidelities + msb)")

    # interpolative coder known answers (interp.hpp:100-118): clustered increasing lists
    out = {}
    for j, (n, u) in enumerate([(1, 10), (2, 3), (7, 1000), (100, 101), (257, 1 << 20), (1000, 5000)]):
        vals = np.sort(rng.choice(u, size=n, replace=False)).astype(np.uint32)
        words = ref.interp_encode(vals, u)
        assert np.array_equal(ref.interp_decode(words, n, u), vals)
        out[f"v{j}"], out[f"w{j}"], out[f"u{j}"] = vals, words, np.uint64(u)
    np.savez_compressed(os.path.join(HERE, "interp.npz"), **out)

    # adjust_freqs known answers on raw histograms, incl. the u16 abort and a > 2^31 total (int overflow quirk)
    hs = {}
    for j, (nb, scale, u16) in enumerate([(300, 50, 1), (1024, 1000, 1), (2048, 3, 1), (5000, 200, 0), (700, 4_000_000, 1),
                                           (64, 40_000_000, 0)]):
        h = (rng.zipf(1.3, nb).astype(np.uint64) * scale) * (rng.random(nb) < 0.8)
        h[0] += 1
        h[nb - 1] += 1
        largest = int(np.nonzero(h)[0][-1])
        hs[f"h{j}"], hs[f"u16_{j}"] = h, np.int32(u16)
        hs[f"n{j}"] = ref.adjust_freqs(h, largest, bool(u16))
    np.savez_compressed(os.path.join(HERE, "adjust_freqs.npz"), **hs)
    print("interp + adjust_freqs vectors written")
    pa_remap_vectors(ref)
    ans_byte_vectors(ref)


if __name__ == "__main__":
    main()
