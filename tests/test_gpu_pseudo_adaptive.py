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
@pytest.mark.parametrize("path", [0, 1])  # 0: tables built in shared memory (pa_kernels.cuh); 1: tables in global memory
def test_blocks_equal_reference(codec, method, f, model_syms, n, path):
    codec.set_kernel_path(path)
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


@pytest.mark.parametrize("method,f", [(M_FOLD, 1), (M_FOLD, 2), (M_MSB, 0)])
def test_skewed_and_wide_blocks_both_table_layouts(codec, method, f):
    """blocks with one dominant symbol (frequencies beyond the per-CTA reciprocal table), blocks that use the whole
    value range (every alphabet bin in play) and near-uniform blocks (large frames): both kernel generations must
    write the same bytes, and sampled blocks must equal the reference's"""
    rng = np.random.default_rng(11)
    S = 4096
    parts = []
    for b in range(48):
        kind = b % 4
        if kind == 0:  # 97 % one value
            x = rng.integers(0, 1 << 12, size=S, dtype=np.uint32)
            x[rng.random(S) < 0.97] = 5
        elif kind == 1:  # whole range below 2^30
            x = (2.0 ** rng.uniform(0, 30, size=S)).astype(np.uint32)
            x[0] = (1 << 30) - 1
        elif kind == 2:  # uniform over 2^9 values
            x = rng.integers(0, 1 << 9, size=S, dtype=np.uint32)
        else:
            x = _inputs.make("zipf20", S, seed=100 + b)
        parts.append(x.astype(np.uint32))
    data = np.concatenate(parts + [parts[1][:777]])
    codec.set_kernel_path(0)
    blob = codec.pa_compress(method, f, S, data).copy()
    back = codec.pa_decompress(method, f, blob, data.size)
    assert np.array_equal(back, data)
    codec.set_kernel_path(1)
    blob1 = codec.pa_compress(method, f, S, data).copy()
    assert np.array_equal(blob1, blob), "shared-memory-table and global-table encoders differ"
    assert np.array_equal(codec.pa_decompress(method, f, blob, data.size), data)
    codec.set_kernel_path(0)
    ref = cpu()
    blocks = split(blob, data.size, S)
    for m in list(range(0, 8)) + [len(blocks) - 1]:
        chunk = data[m * S : (m + 1) * S]
        want = ref.compress(method, f, chunk)
        pre, pay = blocks[m]
        assert _container.prelude_equal(pre, want[: pre.size]) and np.array_equal(pay, want[pre.size :]), m


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


@pytest.mark.parametrize("runs", [1, 3, 8])
def test_host_buffer_decode_in_runs_of_models(codec, monkeypatch, runs):
    """b200ans_pa_decompress with host buffers moves the payload in runs of models (copy in / decode / copy out
    overlapped); any number of runs gives the same symbols, and a size table that does not add up is refused with
    nothing left in flight"""
    import ans_large_alphabet_b200 as P

    monkeypatch.setenv("B200ANS_PA_RUNS", str(runs))
    n = (1 << 24) + 777
    data = _inputs.make("zipf20", n, seed=21)
    blob = codec.pa_compress(M_FOLD, 1, 8192, data).copy()
    assert blob.size > (24 << 20)  # enough payload for several runs of 8 MiB
    assert np.array_equal(codec.pa_decompress(M_FOLD, 1, blob, n), data)
    nm = (n + 8191) // 8192
    bad = blob.copy()
    sizes = bad[64 : 64 + 8 * nm].view(np.uint32)
    sizes[nm + nm // 2] += 16  # one payload size off by a 16-byte unit
    with pytest.raises(P.B200AnsError):
        codec.pa_decompress(M_FOLD, 1, bad, n)
    assert np.array_equal(codec.pa_decompress(M_FOLD, 1, blob, n), data)


@pytest.mark.parametrize("heavy,lane_per_block", [((0.5,), True), ((0.3, 0.3), False)])
def test_lane_per_block_encoder_frequencies_beyond_its_reciprocal_table(codec, monkeypatch, heavy, lane_per_block):
    """The lane-per-block encoder (lpb3_kernels.cuh, PA) takes reciprocals from a table indexed by the frequency. One
    frequency beyond the table per model is served from a register; a model with two sends the whole input through the
    four-lanes-per-block encoder again. Same bytes either way (a short table stands in for frequencies above 4096)."""
    rng = np.random.default_rng(11)
    n = 8192 * 5 + 100
    data = rng.integers(2, 300, size=n, dtype=np.uint32)
    u = rng.random(n)
    lo = 0.0
    for v, p in enumerate(heavy):
        data[(u >= lo) & (u < lo + p)] = v
        lo += p
    monkeypatch.setenv("B200ANS_NO_PA3", "1")
    want = codec.pa_compress(M_FOLD, 1, 8192, data).copy()
    assert codec.stats().encode_kernel.decode() != "lpb3_encode_kernel<PA>"
    monkeypatch.delenv("B200ANS_NO_PA3")
    monkeypatch.setenv("B200ANS_PA3_RCPN", "257")
    got = codec.pa_compress(M_FOLD, 1, 8192, data).copy()
    assert (codec.stats().encode_kernel.decode() == "lpb3_encode_kernel<PA>") == lane_per_block
    assert np.array_equal(got, want)
    assert np.array_equal(codec.pa_decompress(M_FOLD, 1, got, n), data)
    monkeypatch.delenv("B200ANS_PA3_RCPN")
    got = codec.pa_compress(M_FOLD, 1, 8192, data).copy()
    assert codec.stats().encode_kernel.decode() == "lpb3_encode_kernel<PA>"
    assert np.array_equal(got, want)


@pytest.mark.parametrize("model_syms,n", [(1024, 30_000), (8192, 70_001)])
def test_plain_ans_blocks_over_dense_ranks_equal_reference(codec, model_syms, n):
    """Pseudo-adaptive plain ANS (run<ANSint> of the driver codes the dense ranks of its remap, pseudo_adaptive.cpp:235):
    every block equals ans_int_compress(block); values beyond model_syms are outside the domain."""
    import ans_large_alphabet_b200 as P

    rng = np.random.default_rng(5)
    data = np.minimum(rng.zipf(1.3, size=n), model_syms).astype(np.uint32)  # ranks 1 .. model_syms, skewed
    blob = codec.pa_compress(M_INT, 0, model_syms, data).copy()
    ref = cpu()
    for m, (pre, pay) in enumerate(split(blob, n, model_syms)):
        chunk = data[m * model_syms : (m + 1) * model_syms]
        want = ref.compress(M_INT, 0, chunk)
        assert pre.size + pay.size == want.size, m
        assert _container.prelude_equal(pre, want[: pre.size]) and np.array_equal(pay, want[pre.size :]), m
    assert np.array_equal(codec.pa_decompress(M_INT, 0, blob, n), data)
    data[17] = model_syms + 1
    with pytest.raises(P.B200AnsError) as e:
        codec.pa_compress(M_INT, 0, model_syms, data)
    assert e.value.code == -4
