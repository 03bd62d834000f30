"""not gpu: host-side entry points added in round 2.

* b200ans_scale_freqs == the reference's scale_freqs (ans_util.hpp:77-95), live against the shim where it is built and
  against a plain numpy restatement everywhere;
* oracle/_ref's histogram constructor (ref_model_from_hist, used by bench.py's full-size parity check) gives the
  model create() gives;
* untrusted preludes: the bounded reader of host/prelude.hpp never walks off the caller's buffer (b200ans_load_interp
  keeps the reference's trusting signature; the bounded reader sits behind b200ans_model_load, which needs a GPU, so
  here it is reached through the C++ test program tests/cpp/prelude_fuzz.cpp)."""
import os
import subprocess

import numpy as np
import pytest

import _inputs
from _libs import M_FOLD, M_INT, M_MSB, have_reference, reference

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _np_scale_freqs(F, mapping, M, freq_sum):
    """ans_util.hpp:77-95 in numpy scalars (float64, fused multiply-add as the reference's build contracts it)"""
    import math

    S = np.zeros(F.size, dtype=np.uint32)
    M = int(M)
    for sym in mapping:
        ratio = float(M) / float(freq_sum)
        share = int(math.fma(ratio, float(F[sym]), 0.5)) if hasattr(math, "fma") else int(0.5 + ratio * float(F[sym]))
        share &= 0xFFFFFFFF
        if share == 0:
            share = 1
        S[sym] = share
        M -= share
        freq_sum -= int(F[sym])
        if M < 0:
            break
    return S, M == 0


@pytest.mark.parametrize("kind,method,f", [("zipf20", M_FOLD, 1), ("zipf20", M_FOLD, 2), ("geom", M_INT, 0), ("wide", M_MSB, 0)])
def test_scale_freqs_matches_reference(host, kind, method, f):
    data = _inputs.make(kind, 50_000, seed=3)
    if method == M_INT:
        F = np.bincount(data).astype(np.uint64)
    else:
        nb = 1280 if method == M_MSB else 1 << (f + 9)
        F = np.zeros(nb, dtype=np.uint64)
        for v, c in zip(*np.unique(data, return_counts=True)):
            F[host.msb_map(v)[0] if method == M_MSB else host.fold_map(f, v)[0]] += np.uint64(c)
    present = np.nonzero(F)[0]
    order = present[np.lexsort((present, F[present]))].astype(np.uint32)  # ascending (count, symbol): ans_util.hpp:114-122
    sigma = order.size
    m0 = 1 << int(np.ceil(np.log2(sigma)))
    seen_fit = seen_fail = False
    for M in (m0, 2 * m0, 8 * m0, 64 * m0):
        S, fits = host.scale_freqs(F, order, M)
        if have_reference():
            S_ref, fits_ref = reference().scale_freqs(F, order, M)
            assert fits == fits_ref and (not fits or np.array_equal(S, S_ref)), M
        if hasattr(__import__("math"), "fma"):
            S_np, fits_np = _np_scale_freqs(F, order, M, int(F.sum()))
            assert fits == fits_np and (not fits or np.array_equal(S, S_np)), M
        if fits:
            assert int(S.sum()) == M and np.all(S[present] >= 1)
        seen_fit |= fits
        seen_fail |= not fits
    assert seen_fit
    with pytest.raises(Exception):
        host.scale_freqs(F, np.array([F.size], dtype=np.uint32), m0)  # mapping entry out of range


@pytest.mark.parametrize("kind,method,f", [("zipf20", M_FOLD, 1), ("zipf20", M_FOLD, 2), ("zipf20", M_MSB, 0), ("geom", M_INT, 0),
                                           ("zipf16", M_INT, 0)])
def test_reference_model_from_histogram_equals_create(ref, host, kind, method, f):
    """the constructor bench.py uses at 2^30 / 2^32 symbols == ans_X_encode::create on the input itself"""
    data = _inputs.make(kind, 120_001, seed=5)
    want = ref.model_create(method, f, data)
    if method == M_INT:
        bins = np.bincount(data).astype(np.uint64)
    else:
        nb = 1280 if method == M_MSB else 1 << (f + 9)
        bins = np.zeros(nb, dtype=np.uint64)
        for v, c in zip(*np.unique(data, return_counts=True)):
            bins[host.msb_map(v)[0] if method == M_MSB else host.fold_map(f, v)[0]] += np.uint64(c)
    got = ref.model_from_hist(method, f, bins)
    fw, nw = want.info()
    fg, ng = got.info()
    assert fw == fg and np.array_equal(nw, ng)
    assert np.array_equal(want.serialize()[:-4], got.serialize()[:-4])
    for lo in (0, 8192, 65536):
        assert np.array_equal(want.encode_block(data[lo : lo + 8192]), got.encode_block(data[lo : lo + 8192]))


def test_bounded_prelude_reader(tmp_path):
    """host/prelude.hpp ans_load_interp_checked on truncated / hostile bytes (a C++ program: the function is header-only)"""
    exe = tmp_path / "prelude_fuzz"
    src = os.path.join(ROOT, "tests", "cpp", "prelude_fuzz.cpp")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
                           "-I", os.path.join(ROOT, "ans_large_alphabet_b200", "host"), "-o", str(exe), src])
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ok" in out.stdout
