"""not gpu: pins the CPU oracle (oracle/ans_oracle.c, the plain-C restatement) against the golden
vectors generated from the UNMODIFIED reference (tests/golden/make_golden.py) and -- where the
reference shim was built (oracle/_ref/libansref.so, needs /root/reference at build time) --
against the reference itself on fresh seeded inputs.  Bar: bit-exact."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

import _container
import _inputs
from _libs import M_FOLD, M_INT, M_MSB, M_RFOLD, have_reference

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLD, "*.npz"))
               if os.path.basename(p)[:-4] not in ("mappings", "interp", "adjust_freqs", "pa_remap", "ans_byte"))


def load(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def split_blocks(g):
    offs = np.concatenate([[0], np.cumsum(g["block_sizes"].astype(np.int64))])
    return [g["block_payloads"][offs[i] : offs[i + 1]] for i in range(len(offs) - 1)]


def test_golden_set_is_complete():
    assert len(CASES) >= 14 and {"mappings", "interp", "adjust_freqs", "pa_remap"} <= {os.path.basename(p)[:-4] for p in glob.glob(GOLD + "/*.npz")}


@pytest.mark.parametrize("name", CASES)
def test_oracle_model_and_prelude_match_reference(orc, name):
    g = load(name)
    method, f, data = int(g["method"]), int(g["fidelity"]), g["data"]
    m = orc.model_create(method, f, data)
    frame, nfreqs = m.info()
    assert frame == int(g["frame"])
    assert np.array_equal(nfreqs, g["nfreqs"]), "normalised frequencies differ from the reference"
    assert _container.prelude_equal(m.serialize(), g["prelude"]), "prelude differs from the reference"
    if method == M_RFOLD:
        sigma, mf = m.rfold_info()
        assert sigma == int(g["rfold_sigma"]) and np.array_equal(mf, g["most_frequent"])


@pytest.mark.parametrize("name", CASES)
def test_oracle_stream_and_blocks_match_reference(orc, name):
    g = load(name)
    method, f, data = int(g["method"]), int(g["fidelity"]), g["data"]
    stream = orc.compress(method, f, data)
    pre = g["prelude"].size
    assert stream.size == g["stream"].size
    assert _container.prelude_equal(stream[:pre], g["stream"][:pre])
    assert np.array_equal(stream[pre:], g["stream"][pre:]), "whole-stream payload differs from ans_X_compress"
    assert np.array_equal(orc.decompress(method, f, g["stream"], data.size), data), "oracle cannot decode the reference stream"
    bs = int(g["block_syms"])
    m = orc.model_create(method, f, data)
    dec = orc.model_load(method, f, g["prelude"])
    for b, want in enumerate(split_blocks(g)):
        blk = data[b * bs : (b + 1) * bs]
        got = m.encode_block(blk)
        assert np.array_equal(got, want), f"block {b} payload differs from the reference coder"
        assert np.array_equal(dec.decode_block(want, blk.size), blk)


def test_oracle_mappings_match_reference(orc):
    g = load("mappings")
    L = orc.lib
    xs = g["x"]
    for f in range(1, 9):
        for i in range(0, xs.size, 3 if f > 2 else 1):
            k = C.c_uint32()
            s = L.orc_fold_mapping(f, int(xs[i]), C.byref(k))
            assert s == g["fold"][f - 1, i]
            v, kb = C.c_uint32(), C.c_uint32()
            L.orc_fold_undo(f, s, C.byref(v), C.byref(kb))
            assert (v.value, kb.value) == (g["fold_undo_value"][f - 1, i], g["fold_undo_bytes"][f - 1, i])
            assert kb.value == k.value
    for i in range(xs.size):
        k = C.c_uint32()
        s = L.orc_msb_mapping(int(xs[i]), C.byref(k))
        assert s == g["msb"][i]
        v, kb = C.c_uint32(), C.c_uint32()
        L.orc_msb_undo(s, C.byref(v), C.byref(kb))
        assert (v.value, kb.value) == (g["msb_undo_value"][i], g["msb_undo_bytes"][i])


def test_oracle_fold_identity_sweep(orc):
    """src/test.cpp:49-71 restated: undo(fold(x)) == x with the low 8*bytes bits masked, f = 1 and 2,
    every 4099th value below 2^30 plus all values below 2^18."""
    L = orc.lib
    for f in (1, 2):
        xs = np.concatenate([np.arange(0, 1 << 18, 37, dtype=np.uint64), np.arange(0, 1 << 30, 4_000_037, dtype=np.uint64)])
        for x in xs:
            k = C.c_uint32()
            s = L.orc_fold_mapping(f, int(x), C.byref(k))
            v, kb = C.c_uint32(), C.c_uint32()
            L.orc_fold_undo(f, s, C.byref(v), C.byref(kb))
            assert kb.value == k.value and v.value == (int(x) >> (8 * k.value)) << (8 * k.value)
            assert s < (1 << (f + 9))


def test_oracle_interp_matches_reference(orc):
    g = load("interp")
    j = 0
    while f"v{j}" in g:
        vals, words, u = g[f"v{j}"], g[f"w{j}"], int(g[f"u{j}"])
        got = orc.interp_encode(vals, u)
        assert got.size == words.size
        assert _container.prelude_equal(got.view(np.uint8), words.view(np.uint8)) or np.array_equal(got, words)
        assert np.array_equal(orc.interp_decode(words, vals.size, u), vals)
        j += 1
    assert j >= 6


def test_oracle_adjust_freqs_matches_reference(orc):
    g = load("adjust_freqs")
    j = 0
    while f"h{j}" in g:
        h = g[f"h{j}"]
        largest = int(np.nonzero(h)[0][-1])
        got = orc.adjust_freqs(h, largest, bool(g[f"u16_{j}"]))
        assert np.array_equal(got, g[f"n{j}"]), f"histogram {j}"
        j += 1
    assert j >= 6


# ---- live comparison against the compiled reference (fresh inputs, more shapes) -------------------
live = pytest.mark.skipif(not have_reference(), reason="oracle/_ref/libansref.so not built")


@live
@pytest.mark.parametrize("method,f", [(M_FOLD, 1), (M_FOLD, 2), (M_FOLD, 3), (M_MSB, 0), (M_INT, 0), (M_RFOLD, 1), (M_RFOLD, 2)])
@pytest.mark.parametrize("kind,n", [("zipf20", 50_003), ("geom", 40_000), ("zipf12s", 30_001), ("u16", 10_000)])
def test_oracle_equals_reference_live(orc, ref, method, f, kind, n):
    if method == M_RFOLD and kind in ("geom", "zipf12s"):
        pytest.skip("rfold with sigma < 2^(f+7): the reference itself does not round-trip (SURVEY.md section 4)")
    data = _inputs.make(kind, n, seed=21)
    a, b = orc.compress(method, f, data), ref.compress(method, f, data)
    pre = ref.model_create(method, f, data).serialize().size
    assert a.size == b.size and _container.prelude_equal(a[:pre], b[:pre]) and np.array_equal(a[pre:], b[pre:])
    assert np.array_equal(orc.decompress(method, f, b, n), data)
    assert np.array_equal(ref.decompress(method, f, a, n), data)
    mo, mr = orc.model_create(method, f, data), ref.model_create(method, f, data)
    assert mo.info()[0] == mr.info()[0] and np.array_equal(mo.info()[1], mr.info()[1])
    for lo in (0, 8192, n - 4099):
        blk = data[lo : lo + 4096 + 3]
        assert np.array_equal(mo.encode_block(blk), mr.encode_block(blk))


@live
def test_oracle_adjust_freqs_live(orc, ref):
    rng = np.random.default_rng(5)
    for trial in range(40):
        nb = int(rng.integers(40, 3000))
        a = 0.6 + 1.2 * rng.random()
        h = np.floor(float(rng.integers(10_000, 50_000_000)) / np.arange(1, nb + 1) ** a).astype(np.uint64)
        h = h[rng.permutation(nb)] * (rng.random(nb) < 0.8)
        h[int(rng.integers(0, nb))] += 1
        h[0] += 2
        largest = int(np.nonzero(h)[0][-1])
        for u16 in (True, False):
            assert np.array_equal(orc.adjust_freqs(h, largest, u16), ref.adjust_freqs(h, largest, u16)), (trial, u16)


def remap_blocks_equal(a, a_alpha, b, b_alpha, coder_prelude):
    """[sigma][universe][interp] and the coder's prelude up to the reference's undefined padding bits, the rest exactly"""
    if a.size != b.size or a_alpha != b_alpha:
        return False
    if not _container.prelude_equal(a[:a_alpha], b[:b_alpha]):
        return False
    if a.size == a_alpha:
        return True
    at = a_alpha + coder_prelude
    return _container.prelude_equal(a[a_alpha:at], b[b_alpha:at]) and np.array_equal(a[at:], b[at:])


def test_oracle_pa_remap_blocks_match_reference(orc):
    """src/pseudo_adaptive.cpp:83-125 block by block: ranks, block alphabet and coder output of the oracle against
    vectors written by the reference's own interpolative coder and ans_X_compress; the ranks decode and map back"""
    g = load("pa_remap")
    for j in range(int(g["count"])):
        method, f, alpha_bytes, coder_prelude = (int(x) for x in g[f"meta{j}"])
        data, want = g[f"data{j}"], g[f"blob{j}"]
        blob, ab, ranks = orc.pa_remap_block(method, f, data)
        assert np.array_equal(ranks, g[f"ranks{j}"]), j
        assert remap_blocks_equal(blob, ab, want, alpha_bytes, coder_prelude), j
        # and back: alphabet from the running sums, values from the ranks
        sigma, universe = (int(x) for x in want[:8].view(np.uint32))
        sums = orc.interp_decode(want[8:alpha_bytes].view(np.uint32), sigma, universe).astype(np.int64)
        alphabet = np.diff(np.concatenate([[0], sums]))
        assert universe == sums[-1] + 1 and np.array_equal(alphabet, np.unique(data))
        got = orc.decompress(method, f, want[alpha_bytes:], data.size) if sigma != 1 else np.ones(data.size, np.uint32)
        assert np.array_equal(alphabet[got.astype(np.int64) - 1], data), j


@live
@pytest.mark.parametrize("method,f", [(M_INT, 0), (M_MSB, 0), (M_FOLD, 1)])
def test_oracle_pa_remap_live(orc, ref, method, f):
    rng = np.random.default_rng(17)
    for kind, n in (("zipf20", 8192), ("zipf16", 1280), ("geom", 65536), ("u8", 129), ("zipf12s", 4096)):
        data = _inputs.make(kind, n, seed=int(rng.integers(1, 1000)))
        if rng.random() < 0.5:
            data[int(rng.integers(0, n))] = 0  # a block that holds the value 0 (pseudo_adaptive.cpp:92-93)
        a, b = orc.pa_remap_block(method, f, data), ref.pa_remap_block(method, f, data)
        pre = ref.model_create(method, f, b[2]).serialize().size
        assert np.array_equal(a[2], b[2]) and remap_blocks_equal(a[0], a[1], b[0], b[1], pre), (kind, n)
    wide = (rng.integers(1 << 20, 1 << 24, size=4096)).astype(np.uint32)  # running sums beyond 32 bits
    assert orc.pa_remap_block(method, f, wide) is None and ref.pa_remap_block(method, f, wide) is None


def test_oracle_and_host_ans_byte_match_reference(orc, host):
    """ans_byte (ans_byte.hpp): the C restatement (ORC_BYTE) and the product's host normaliser against the reference's
    frequencies, prelude, whole stream and 1024-symbol blocks (tests/golden/ans_byte.npz)"""
    g = load("ans_byte")
    for j in range(int(g["count"])):
        data = g[f"data{j}"].astype(np.uint32)
        m = orc.model_create(4, 0, data)
        frame, nfreqs = m.info()
        assert frame == int(g[f"frame{j}"]) and np.array_equal(nfreqs, g[f"nfreqs{j}"].astype(np.uint32))
        got = host.byte_adjust_freqs(np.bincount(data, minlength=256))
        assert np.array_equal(got, g[f"nfreqs{j}"]), "b200ans_byte_adjust_freqs differs from the reference"
        pre = g[f"prelude{j}"]
        assert _container.prelude_equal(m.serialize(), pre)
        stream = orc.compress(4, 0, data)
        assert stream.size == g[f"stream{j}"].size and np.array_equal(stream[pre.size :], g[f"stream{j}"][pre.size :])
        assert np.array_equal(orc.decompress(4, 0, g[f"stream{j}"], data.size), data)
        offs = np.concatenate([[0], np.cumsum(g[f"block_sizes{j}"].astype(np.int64))])
        dec = orc.model_load(4, 0, pre)
        for b in range(len(offs) - 1):
            blk = data[b * 1024 : (b + 1) * 1024]
            want = g[f"block_payloads{j}"][offs[b] : offs[b + 1]]
            assert np.array_equal(m.encode_block(blk), want), f"case {j} block {b}"
            assert np.array_equal(dec.decode_block(want, blk.size), blk)
