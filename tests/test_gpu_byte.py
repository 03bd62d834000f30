"""-m gpu: ans_byte (ans_byte.hpp, the byte coder of the vbyteANS / streamvbyteANS baselines, methods.hpp:432-482) behind
the same C ABI: B200ANS_METHOD_BYTE through the uint32 entry points and the byte entry points.  Bar: bit-exact -- the
normalised frequencies, the prelude and every block payload equal the reference's; with block_syms = 0 the whole stream
equals ans_byte_compress and the reference's streams decode."""
import numpy as np
import pytest

import _container
from _libs import have_reference, reference

pytestmark = pytest.mark.gpu
M_BYTE = 4


def inputs(kind, n, seed):
    rng = np.random.default_rng(seed)
    if kind == "uniform":
        return rng.integers(0, 256, n).astype(np.uint32)
    if kind == "geom":
        return np.minimum(rng.geometric(0.04, n) - 1, 255).astype(np.uint32)
    if kind == "vbyte":  # what vbyte::encode of small gaps looks like: mostly one-byte codes with the stop bit
        return np.where(rng.random(n) < 0.9, 128 + rng.geometric(0.1, n) % 128, rng.integers(0, 128, n)).astype(np.uint32)
    if kind == "few":
        return rng.integers(3, 8, n).astype(np.uint32)
    if kind == "one":
        return np.full(n, 77, np.uint32)
    raise KeyError(kind)


@pytest.fixture(autouse=True)
def _restore(codec):
    yield
    codec.set_kernel_path(0)
    codec.set_block_syms(8192)


@pytest.mark.parametrize("path", [1, 2])
@pytest.mark.parametrize("kind", ["uniform", "geom", "vbyte", "few", "one"])
def test_blocks_byte_identical_and_roundtrip(codec, kind, path):
    if not have_reference():
        pytest.skip("needs the reference shim (ans_byte.hpp)")
    ref = reference()
    n = 300_007
    data = inputs(kind, n, seed=3)
    codec.set_kernel_path(path)
    for bs in (8192, 1920):
        codec.set_block_syms(bs)
        blob = codec.compress(M_BYTE, 0, data).copy()
        c = _container.parse(blob)
        assert (c["method"], c["n"], c["block_syms"]) == (M_BYTE, n, bs) and c["pad_ok"]
        model = ref.model_create(M_BYTE, 0, data)
        frame, nfreqs = model.info()
        got_frame, got_nf = codec.model_nfreqs()
        assert got_frame == frame and np.array_equal(got_nf, nfreqs)
        assert _container.prelude_equal(c["prelude"], model.serialize()), "prelude differs from ans_byte_encode::serialize"
        for b, payload in enumerate(c["blocks"]):
            want = model.encode_block(data[b * bs : (b + 1) * bs])
            assert np.array_equal(payload, want), f"block {b}"
        assert np.array_equal(codec.decompress(M_BYTE, 0, blob, n), data)
        dec = ref.model_load(M_BYTE, 0, c["prelude"])
        for b in (0, len(c["blocks"]) - 1):
            lo = b * bs
            assert np.array_equal(dec.decode_block(c["blocks"][b], min(bs, n - lo)), data[lo : lo + bs])


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 9, 4097, 70_001])
def test_reference_wire_format_and_byte_entry_points(codec, n):
    if not have_reference():
        pytest.skip("needs the reference shim (ans_byte.hpp)")
    ref = reference()
    data = inputs("vbyte", n, seed=n)
    raw = data.astype(np.uint8)
    want = ref.compress(M_BYTE, 0, data)
    codec.set_block_syms(0)
    got = codec.byte_compress(raw).copy()
    pre = codec.model_prelude().size
    assert got.size == want.size and _container.prelude_equal(got[:pre], want[:pre])
    assert np.array_equal(got[pre:], want[pre:]), "payload differs from ans_byte_compress"
    assert np.array_equal(codec.byte_decompress(want, n), raw), "cannot decode the reference's stream"
    assert np.array_equal(ref.decompress(M_BYTE, 0, got, n), data), "the reference cannot decode ours"
    codec.set_block_syms(8192)
    blob = codec.byte_compress(raw).copy()
    assert np.array_equal(blob, codec.compress(M_BYTE, 0, data)), "byte and uint32 entry points write different containers"
    assert np.array_equal(codec.byte_decompress(blob, n), raw)


def test_byte_domain_and_normaliser(codec, host):
    import ans_large_alphabet_b200 as P

    with pytest.raises(P.B200AnsError) as e:
        codec.compress(M_BYTE, 0, np.array([1, 2, 256, 3], dtype=np.uint32))
    assert e.value.code == -4
    assert codec.byte_compress(np.zeros(0, np.uint8)).size == 64 and codec.byte_decompress(codec.byte_compress(np.zeros(0, np.uint8)), 0).size == 0
    if have_reference():
        ref = reference()
        rng = np.random.default_rng(5)
        for t in range(40):
            d = np.minimum(rng.geometric(0.02 + 0.02 * t, 5000 + 997 * t) - 1, 255).astype(np.uint32)
            frame, nf = ref.model_create(M_BYTE, 0, d).info()
            got = host.byte_adjust_freqs(np.bincount(d, minlength=256))
            assert int(got.sum()) == frame and np.array_equal(got.astype(np.uint32), nf.astype(np.uint32))
