"""-m gpu: pseudo-adaptive mode (one model per block, src/pseudo_adaptive.cpp:64-177) against the reference.

Bar: for every block, prelude(b) + payload(b) of the container equals what the reference's ans_X_compress() writes
for that block on its own (prelude modulo the reference's uninitialised padding bits, payload byte for byte), the
reference decodes every block, and the container round-trips."""
import numpy as np
import pytest

import _container
import _inputs
from _libs import M_FOLD, M_INT, M_MSB, have_reference, oracle, reference

pytestmark = pytest.mark.gpu


def cpu():
    return reference() if have_reference() else oracle()


def split(blob, n, model_syms):
    """(prelude bytes, payload bytes) per block from a pseudo-adaptive container"""
    assert np.frombuffer(blob[:4].tobytes(), dtype=np.uint32)[0] == 0x4E413242
    nm = (n + model_syms - 1) // model_syms
    sizes = np.frombuffer(blob[64 : 64 + 8 * nm].tobytes(), dtype=np.uint32)
    pre, pay = sizes[:nm], sizes[nm:]
    at = (64 + 8 * nm + 15) // 16 * 16
    out = []
    pres = []
    for m in range(nm):
        pres.append(blob[at : at + pre[m]])
        at += (int(pre[m]) + 15) // 16 * 16
    for m in range(nm):
        out.append((pres[m], blob[at : at + pay[m]]))
        at += (int(pay[m]) + 15) // 16 * 16
    assert at == blob.size
    return out


@pytest.mark.parametrize("method,f", [(M_FOLD, 1), (M_FOLD, 2), (M_MSB, 0)])
@pytest.mark.parametrize("model_syms,n", [(1024, 20_000), (8192, 100_003), (65536, 300_001), (8192, 8192 * 3)])
def test_blocks_equal_reference(codec, method, f, model_syms, n):
    data = _inputs.make("zipf20", n, seed=7)
    blob = codec.pa_compress(method, f, model_syms, data).copy()
    blocks = split(blob, n, model_syms)
    ref = cpu()
    total = 0
    for m, (pre, pay) in enumerate(blocks):
        chunk = data[m * model_syms : (m + 1) * model_syms]
        want = ref.compress(method, f, chunk)
        assert pre.size + pay.size == want.size, (m, pre.size, pay.size, want.size)
        assert _container.prelude_equal(pre, want[: pre.size]), m
        assert np.array_equal(pay, want[pre.size :]), f"block {m}: payload differs from the reference coder"
        if m % 7 == 0:
            got = ref.decompress(method, f, np.concatenate([pre, pay]), chunk.size)
            assert np.array_equal(got, chunk)
        total += want.size
    assert np.array_equal(codec.pa_decompress(method, f, blob, n), data)
    assert blob.size >= total and blob.size <= total + 64 + len(blocks) * (8 + 30) + 16


def test_large_round_trip_and_gain(codec):
    """per-block models pay off on data whose distribution drifts"""
    rng = np.random.default_rng(5)
    n = 1 << 23
    scale = np.repeat(2 ** rng.integers(2, 18, size=n // 65536), 65536)
    data = (rng.exponential(1.0, size=n) * scale).astype(np.uint32)
    blob = codec.pa_compress(M_FOLD, 1, 65536, data).copy()
    assert np.array_equal(codec.pa_decompress(M_FOLD, 1, blob, n), data)
    codec.set_block_syms(8192)
    semi = codec.compress(M_FOLD, 1, data).copy()
    assert blob.size < semi.size, (blob.size, semi.size)


def test_edges_and_errors(codec):
    import ans_large_alphabet_b200 as P

    blob = codec.pa_compress(M_FOLD, 1, 8192, np.zeros(0, np.uint32)).copy()
    assert blob.size == 64 and codec.pa_decompress(M_FOLD, 1, blob, 0).size == 0
    one = np.full(5000, 77, np.uint32)  # a single distinct symbol per block
    blob = codec.pa_compress(M_FOLD, 1, 1024, one).copy()
    assert np.array_equal(codec.pa_decompress(M_FOLD, 1, blob, one.size), one)
    data = _inputs.make("zipf20", 50_000, seed=1)
    blob = codec.pa_compress(M_MSB, 0, 2048, data).copy()
    for bad in (lambda: codec.pa_compress(M_INT, 0, 8192, data), lambda: codec.pa_compress(M_FOLD, 3, 8192, data),
                lambda: codec.pa_compress(M_FOLD, 1, 1000, data), lambda: codec.pa_decompress(M_FOLD, 1, blob, data.size),
                lambda: codec.pa_decompress(M_MSB, 0, blob, data.size - 1), lambda: codec.pa_decompress(M_MSB, 0, blob[:-40], data.size),
                lambda: codec.pa_decompress(M_MSB, 0, blob[:50], data.size), lambda: codec.decompress(M_MSB, 0, blob, data.size)):
        with pytest.raises(P.B200AnsError):
            bad()
    with pytest.raises(P.B200AnsError):
        codec.pa_compress(M_FOLD, 1, 8192, np.array([1, 2, 1 << 30, 4] * 600, dtype=np.uint32))
    assert np.array_equal(codec.pa_decompress(M_MSB, 0, blob, data.size), data)


def test_host_fallback_of_the_normaliser_gives_the_same_bytes(codec, monkeypatch):
    """pa_fit_kernel hands models with a decision margin below 1e-9 to the host's adjust_freqs; forcing
    every model down that road must not change a byte"""
    import ans_large_alphabet_b200 as P

    data = _inputs.make("zipf20", 150_001, seed=3)
    want = codec.pa_compress(M_FOLD, 1, 2048, data).copy()
    monkeypatch.setenv("B200ANS_PA_MARGIN", "1e30")
    other = P.Codec(device=0)
    try:
        got = other.pa_compress(M_FOLD, 1, 2048, data).copy()
    finally:
        other.close()
    assert np.array_equal(got, want)
