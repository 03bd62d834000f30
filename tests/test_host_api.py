"""not gpu: the product's host-side model API (b200ans_adjust_freqs / _serialize_interp / _load_interp /
_interp_encode / _interp_decode / _fold_map ... of include/b200ans.h, implemented in
ans_large_alphabet_b200/host/*.hpp) against the golden vectors of the unmodified reference and
against the oracle.  These entry points need no GPU.  Bar: bit-exact."""
import glob
import os

import numpy as np
import pytest

import _container
import _inputs
from _libs import M_FOLD, M_INT, M_MSB, M_RFOLD

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLD, "*.npz"))
               if os.path.basename(p)[:-4] not in ("mappings", "interp", "adjust_freqs", "pa_remap", "ans_byte"))


def load(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def histogram(host, method, f, data, most_frequent=None):
    """what K1 computes, restated with the product's own host mapping functions"""
    if method == M_INT:
        return np.bincount(data, minlength=int(data.max()) + 1).astype(np.uint64)
    if method == M_RFOLD:
        thres = 1 << (f + 7)
        if most_frequent is not None and most_frequent.size:
            lut = {int(v): i for i, v in enumerate(most_frequent)}
            data = np.array([lut.get(int(v), int(v) + thres) for v in data], dtype=np.uint32)
    nb = 1280 if method == M_MSB else 1 << (f + 9)
    uniq, cnt = np.unique(data, return_counts=True)
    h = np.zeros(nb, dtype=np.uint64)
    for v, c in zip(uniq, cnt):
        s = host.msb_map(v)[0] if method == M_MSB else host.fold_map(f, v)[0]
        h[s] += c
    return h


@pytest.mark.parametrize("name", CASES)
def test_normalise_and_prelude_match_reference(host, name):
    g = load(name)
    method, f, data = int(g["method"]), int(g["fidelity"]), g["data"]
    mf = g["most_frequent"] if method == M_RFOLD else None
    h = histogram(host, method, f, data, mf)
    largest = int(np.nonzero(h)[0][-1])
    nfreqs = host.adjust_freqs(h, largest, method != M_INT)
    assert np.array_equal(nfreqs, g["nfreqs"]), "adjust_freqs differs from the reference"
    assert int(nfreqs.sum()) == int(g["frame"])
    blob = host.serialize_interp(nfreqs, int(g["frame"]))
    skip = 0
    if method == M_RFOLD:  # flag word [+ most_frequent list] precede the interp part (ans_reorder_fold.hpp:132-154)
        skip = 4 + (4 * mf.size if int(g["rfold_sigma"]) >= (1 << (f + 7)) else 0)
    assert _container.prelude_equal(blob, g["prelude"][skip:]), "ans_serialize_interp differs from the reference"
    back, used = host.load_interp(g["prelude"][skip:])
    assert np.array_equal(back, nfreqs) and used == g["prelude"].size - skip


def test_adjust_freqs_vectors(host):
    g = load("adjust_freqs")
    j = 0
    while f"h{j}" in g:
        h = g[f"h{j}"]
        largest = int(np.nonzero(h)[0][-1])
        assert np.array_equal(host.adjust_freqs(h, largest, bool(g[f"u16_{j}"])), g[f"n{j}"]), j
        j += 1


def test_adjust_freqs_against_oracle(host, orc):
    rng = np.random.default_rng(17)
    for trial in range(60):
        nb = int(rng.integers(40, 2500))
        a = 0.5 + 1.4 * rng.random()
        h = np.floor(float(rng.integers(5_000, 80_000_000)) / np.arange(1, nb + 1) ** a).astype(np.uint64)
        h = h[rng.permutation(nb)] * (rng.random(nb) < 0.85)
        h[0] += 3
        largest = int(np.nonzero(h)[0][-1])
        for u16 in (True, False):
            assert np.array_equal(host.adjust_freqs(h, largest, u16), orc.adjust_freqs(h, largest, u16)), (trial, u16)


def test_adjust_freqs_h_approx(host, orc):
    """ANSsint / ANSsmsb (ans_sint.hpp, ans_smsb.hpp) only forward H_approx to adjust_freqs"""
    data = _inputs.make("zipf20", 200_000, seed=4)
    h = np.bincount(data).astype(np.uint64)
    last = None
    for ha in (1, 5, 10, 50, 100):
        a = host.adjust_freqs(h, h.size - 1, False, ha)
        assert np.array_equal(a, orc.adjust_freqs(h, h.size - 1, False, ha))
        if last is not None:
            assert a.sum() <= last  # looser target -> never a larger frame
        last = a.sum()


def test_interp_vectors(host):
    g = load("interp")
    j = 0
    while f"v{j}" in g:
        vals, words, u = g[f"v{j}"], g[f"w{j}"], int(g[f"u{j}"])
        got = host.interp_encode(vals, u)
        assert got.size == words.size
        assert _container.prelude_equal(got.view(np.uint8), words.view(np.uint8)) or np.array_equal(got, words)
        assert np.array_equal(host.interp_decode(words, vals.size, u), vals)
        j += 1


def test_interp_roundtrip_clustered(host):
    """interp_test.cpp:134-217 restated: clustered increasing list, encode -> decode"""
    rng = np.random.default_rng(3)
    gaps = np.where(rng.random(50_000) < 0.5, 1, rng.geometric(0.001, 50_000)).astype(np.uint64)
    vals = np.cumsum(gaps)
    u = int(vals[-1]) + 1
    assert u < (1 << 32)
    words = host.interp_encode(vals.astype(np.uint32), u)
    assert np.array_equal(host.interp_decode(words, vals.size, u), vals.astype(np.uint32))


def test_mappings(host):
    g = load("mappings")
    xs = g["x"]
    for f in range(1, 9):
        for i in range(0, xs.size, 5 if f > 2 else 1):
            s, k = host.fold_map(f, xs[i])
            assert s == g["fold"][f - 1, i]
            v, kb = host.fold_unmap(f, s)
            assert (v, kb) == (g["fold_undo_value"][f - 1, i], g["fold_undo_bytes"][f - 1, i]) and k == kb
    for i in range(xs.size):
        s, k = host.msb_map(xs[i])
        assert s == g["msb"][i]
        v, kb = host.msb_unmap(s)
        assert (v, kb) == (g["msb_undo_value"][i], g["msb_undo_bytes"][i]) and k == kb


def test_method_names():
    import ans_large_alphabet_b200 as P

    # methods.hpp:485,500,530-533,550-553
    assert P.method_name(M_INT) == "ANS"
    assert P.method_name(M_MSB) == "ANSmsb"
    assert P.method_name(M_FOLD, 1) == "ANSfold-1" and P.method_name(M_FOLD, 2) == "ANSfold-2"
    assert P.method_name(M_RFOLD, 2) == "ANSrfold-2"
