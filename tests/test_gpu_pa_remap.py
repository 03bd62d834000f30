"""-m gpu: the pseudo-adaptive driver's dense-rank alphabet remap on the device (src/pseudo_adaptive.cpp:83-125, SURVEY.md
section 8 row f1).  Bar: bit-exact -- for every block the alphabet bytes ([sigma][universe][interp of the running sums])
and the coded ranks (prelude + payload of T::encode) equal what the reference's driver writes for that block."""
import numpy as np
import pytest

import _container
import _inputs
from _libs import M_INT, M_MSB, have_reference, oracle, reference

pytestmark = pytest.mark.gpu


def cpu():
    return reference() if have_reference() else oracle()


def split(blob, n, model_syms):
    """(alphabet bytes, prelude bytes, payload bytes) per block of a remapped pseudo-adaptive container"""
    nm = (n + model_syms - 1) // model_syms
    hdr = np.frombuffer(blob[:64].tobytes(), dtype=np.uint32)
    assert hdr[0] == 0x4E413242 and hdr[7] == 3, "magic / flags (pseudo-adaptive | remap)"
    alpha_total = int(np.frombuffer(blob[40:48].tobytes(), dtype=np.uint64)[0])
    a_sizes = np.frombuffer(blob[64 : 64 + 4 * nm].tobytes(), dtype=np.uint32)
    at = (64 + 4 * nm + 15) // 16 * 16
    alphas = []
    for s in a_sizes:
        alphas.append(blob[at : at + int(s)])
        at += (int(s) + 15) // 16 * 16
    assert at == (64 + 4 * nm + 15) // 16 * 16 + alpha_total
    inner = blob[at:]
    sizes = np.frombuffer(inner[64 : 64 + 8 * nm].tobytes(), dtype=np.uint32)
    pre_at = (64 + 8 * nm + 15) // 16 * 16
    pay_at = pre_at + int(sum((int(x) + 15) // 16 * 16 for x in sizes[:nm]))
    pres, pays = [], []
    for m in range(nm):
        pres.append(inner[pre_at : pre_at + sizes[m]])
        pays.append(inner[pay_at : pay_at + sizes[nm + m]])
        pre_at += (int(sizes[m]) + 15) // 16 * 16
        pay_at += (int(sizes[nm + m]) + 15) // 16 * 16
    return alphas, pres, pays


@pytest.mark.parametrize("method", [M_MSB, M_INT])  # run<ANSmsb> / run<ANSint> (pseudo_adaptive.cpp:235-236)
@pytest.mark.parametrize("kind,n,model_syms", [
    ("zipf20", 70_001, 8192), ("zipf16", 40_000, 1024), ("geom", 50_003, 4096), ("zipf12s", 100_000, 16384), ("u8", 9_000, 1024),
])
def test_blocks_equal_the_reference_driver(codec, kind, n, model_syms, method):
    if method == M_INT and model_syms > 16256:
        pytest.skip("per-block plain ANS: at most 16256 symbols per block")
    data = _inputs.make(kind, n, seed=41)
    blob = codec.pa_remap_compress(method, 0, model_syms, data).copy()
    alphas, pres, pays = split(blob, n, model_syms)
    for b in range(len(alphas)):
        block = data[b * model_syms : (b + 1) * model_syms]
        want = cpu().pa_remap_block(method, 0, block)
        assert want is not None
        ref_blob, alpha_bytes, ranks = want
        # (equal but for the uninitialised padding bits of the reference's last interp word, bits.hpp:148)
        assert _container.prelude_equal(alphas[b], ref_blob[:alpha_bytes]), f"block {b}: alphabet bytes differ from the driver's"
        if np.unique(block).size == 1:
            continue  # the driver writes the alphabet only (pseudo_adaptive.cpp:114)
        body = ref_blob[alpha_bytes:]
        assert _container.prelude_equal(pres[b], body[: pres[b].size]) and np.array_equal(pays[b], body[pres[b].size :]), f"block {b}: coded ranks differ"
    assert np.array_equal(codec.pa_remap_decompress(method, 0, blob, n), data)


def test_large_round_trip_single_value_blocks_and_domain(codec):
    import ans_large_alphabet_b200 as P

    n = 3_000_017
    data = _inputs.make("zipf20", n, seed=42)
    data[8192 * 5 : 8192 * 6] = 4242  # a block with one distinct value
    blob = codec.pa_remap_compress(M_MSB, 0, 8192, data).copy()
    assert np.array_equal(codec.pa_remap_decompress(M_MSB, 0, blob, n), data)
    plain = codec.pa_compress(M_MSB, 0, 8192, data)
    print(f"remapped: {8.0 * blob.size / n:.3f} bits/symbol, raw values: {8.0 * plain.size / n:.3f}")
    # sampled blocks against the driver
    alphas, pres, pays = split(blob, n, 8192)
    for b in (0, 5, 100, len(alphas) - 1):
        block = data[b * 8192 : (b + 1) * 8192]
        ref_blob, alpha_bytes, _ = cpu().pa_remap_block(M_MSB, 0, block)
        assert _container.prelude_equal(alphas[b], ref_blob[:alpha_bytes])
    with pytest.raises(P.B200AnsError) as e:  # value range beyond the shared-memory bitmap
        codec.pa_remap_compress(M_MSB, 0, 8192, np.array([1, 2, 3, 1 << 22] * 1000, dtype=np.uint32))
    assert e.value.code == -4
    with pytest.raises(P.B200AnsError):  # the driver has no run<ANSfold>: ANSmsb and plain ANS only
        codec.pa_remap_compress(1, 1, 8192, data[:10000])
    # per-block plain ANS (run<ANSint>) at size: round trip, sampled blocks against the driver
    blob_i = codec.pa_remap_compress(M_INT, 0, 8192, data).copy()
    assert np.array_equal(codec.pa_remap_decompress(M_INT, 0, blob_i, n), data)
    alphas_i, pres_i, pays_i = split(blob_i, n, 8192)
    for b in (0, 7, 200, len(alphas_i) - 1):
        block = data[b * 8192 : (b + 1) * 8192]
        ref_blob, alpha_bytes, _ = cpu().pa_remap_block(M_INT, 0, block)
        body = ref_blob[alpha_bytes:]
        assert _container.prelude_equal(pres_i[b], body[: pres_i[b].size]) and np.array_equal(pays_i[b], body[pres_i[b].size :]), b
    print(f"plain ANS of the ranks: {8.0 * blob_i.size / n:.3f} bits/symbol")
    bad = blob.copy()
    bad[64 + 4 * len(alphas) + 40] ^= 0x5A  # inside an alphabet
    try:
        out = codec.pa_remap_decompress(M_MSB, 0, bad, n)
        assert not np.array_equal(out, data)
    except P.B200AnsError as err:
        assert err.code in (-5, -4)
