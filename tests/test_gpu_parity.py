"""-m gpu: the CUDA path (through the C-ABI) against the reference / the oracle.

Bar: bit-exact.  Every block payload must equal the reference coder run on that block with the
same (global) model; normalised frequencies and prelude must equal the reference's; decode must
restore the input; the reference-wire-format mode must reproduce ans_X_compress byte for byte.
"""
import numpy as np
import pytest

import _container
import _inputs
from _libs import M_FOLD, M_INT, M_MSB, M_RFOLD, have_reference, oracle, reference

pytestmark = pytest.mark.gpu

PATHS = [pytest.param(1, id="group"), pytest.param(2, id="lpb")]  # B200ANS_PATH_GROUP / _LPB: identical bytes


@pytest.fixture(autouse=True)
def _auto_path(codec):
    yield
    codec.set_kernel_path(0)


METHODS = [(M_FOLD, 1), (M_FOLD, 2), (M_MSB, 0), (M_INT, 0), (M_RFOLD, 1), (M_RFOLD, 2), (M_FOLD, 5), (M_FOLD, 8)]


def cpu():
    """the unmodified reference when its shim was built, else the pinned C restatement"""
    return reference() if have_reference() else oracle()


def ok_for(method, f, kind):
    # reference limits: rfold breaks when sigma < 2^(f+7) and values >= 2^(f+7);
    # 'wide' holds values up to 2^30-1 which rfold (value + 2^(f+7)) cannot take.
    if method == M_RFOLD and kind in ("wide", "geom", "geom.2", "u8"):
        return False
    if method == M_INT and kind == "wide":
        return False  # 2^30-entry alphabet: out of scope for a unit test
    return True


def check_container(codec, method, f, data, blob, block_syms):
    c = _container.parse(blob)
    assert (c["method"], c["n"], c["block_syms"]) == (method, data.size, block_syms)
    assert c["pad_ok"], "padding between blocks must be zero"
    model = cpu().model_create(method, f, data)
    frame, nfreqs = model.info()
    assert _container.prelude_equal(c["prelude"], model.serialize()), "prelude differs from the reference"
    for b, payload in enumerate(c["blocks"]):
        want = model.encode_block(data[b * block_syms : (b + 1) * block_syms])
        assert payload.size == want.size and np.array_equal(payload, want), f"block {b} payload differs"
    # and the CPU decoder reads our blocks
    dec = cpu().model_load(method, f, c["prelude"])
    for b in (0, len(c["blocks"]) // 2, len(c["blocks"]) - 1):
        lo = b * block_syms
        got = dec.decode_block(c["blocks"][b], min(block_syms, data.size - lo))
        assert np.array_equal(got, data[lo : lo + block_syms])
    return c


@pytest.mark.parametrize("path", PATHS)
@pytest.mark.parametrize("kind", ["zipf20", "geom", "wide", "u8", "zipf12s"])
@pytest.mark.parametrize("method,f", METHODS)
def test_blocks_byte_identical_and_roundtrip(codec, method, f, kind, path):
    if not ok_for(method, f, kind):
        pytest.skip("outside the reference's valid domain")
    codec.set_kernel_path(path)
    n = 200_003  # not a multiple of 4 nor of the block size
    data = _inputs.make(kind, n, seed=7)
    codec.set_block_syms(8192)
    blob = codec.compress(method, f, data).copy()
    check_container(codec, method, f, data, blob, 8192)
    back = codec.decompress(method, f, blob, n)
    assert np.array_equal(back, data)


@pytest.mark.parametrize("method,f", [(M_FOLD, 1), (M_INT, 0), (M_MSB, 0), (M_RFOLD, 2)])
@pytest.mark.parametrize("block_syms", [128, 1024, 4096, 65536, 14208, 1920])  # the last two: not powers of two
@pytest.mark.parametrize("path", PATHS)
def test_block_sizes(codec, method, f, block_syms, path):
    codec.set_kernel_path(path)
    data = _inputs.make("zipf16", 70_001, seed=3)
    codec.set_block_syms(block_syms)
    blob = codec.compress(method, f, data).copy()
    check_container(codec, method, f, data, blob, block_syms)
    assert np.array_equal(codec.decompress(method, f, blob, data.size), data)
    codec.set_block_syms(8192)


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 7, 8, 127, 128, 129, 8191, 8192, 8193, 8195, 16384 + 2])
@pytest.mark.parametrize("method,f", [(M_FOLD, 1), (M_INT, 0), (M_MSB, 0)])
@pytest.mark.parametrize("path", PATHS)
def test_ragged_lengths(codec, method, f, n, path):
    codec.set_kernel_path(path)
    base = _inputs.make("zipf16", 20_000, seed=11)
    data = base[:n].copy()
    if method == M_INT and np.unique(data).size < 2:
        pytest.skip("the reference hangs on single-symbol input for plain ANS")
    codec.set_block_syms(8192)
    blob = codec.compress(method, f, data).copy()
    check_container(codec, method, f, data, blob, 8192)
    assert np.array_equal(codec.decompress(method, f, blob, n), data)


@pytest.mark.parametrize("kind", ["zipf20", "geom", "wide"])
@pytest.mark.parametrize("method,f", METHODS)
@pytest.mark.parametrize("path", PATHS)
def test_reference_wire_format(codec, method, f, kind, path):
    """block_syms = 0: output == ans_X_compress(), and reference streams decode."""
    if not ok_for(method, f, kind):
        pytest.skip("outside the reference's valid domain")
    codec.set_kernel_path(path)
    n = 30_001
    data = _inputs.make(kind, n, seed=5)
    want = cpu().compress(method, f, data)
    codec.set_block_syms(0)
    try:
        got = codec.compress(method, f, data).copy()
        assert got.size == want.size
        model = cpu().model_create(method, f, data)
        pre = model.serialize().size
        assert _container.prelude_equal(got[:pre], want[:pre])
        assert np.array_equal(got[pre:], want[pre:]), "payload differs from ans_X_compress"
        assert np.array_equal(codec.decompress(method, f, want, n), data), "cannot decode the reference's stream"
        assert np.array_equal(cpu().decompress(method, f, got, n), data), "reference cannot decode our stream"
    finally:
        codec.set_block_syms(8192)


@pytest.mark.parametrize("h_approx", [5, 40, 320])
@pytest.mark.parametrize("method", [M_INT, M_MSB])
def test_h_approx_variants(codec, method, h_approx):
    """ANSsint<H> / ANSsmsb<H> (methods.hpp:584-620): same coders, adjust_freqs stops at H/1000 above the entropy."""
    if not have_reference():
        pytest.skip("needs the reference shim (ans_sint.hpp / ans_smsb.hpp)")
    ref = reference()
    if not hasattr(ref.lib, "ref_compress_h"):
        pytest.skip("reference shim was built without the H_approx entry points")
    n = 60_001
    data = _inputs.make("zipf20" if method == M_MSB else "geom", n, seed=11)
    want = ref.compress_h(method, h_approx, data)
    plain = ref.compress(method, 0, data)
    codec.set_h_approx(h_approx)
    try:
        codec.set_block_syms(0)
        got = codec.compress(method, 0, data).copy()
        assert got.size == want.size
        pre = codec.model_prelude().size
        assert _container.prelude_equal(got[:pre], want[:pre])
        assert np.array_equal(got[pre:], want[pre:]), "payload differs from ans_sint/smsb_compress<H>"
        assert np.array_equal(ref.decompress_h(method, got, n), data), "reference cannot decode our stream"
        assert np.array_equal(codec.decompress(method, 0, want, n), data), "cannot decode the reference's stream"
        if h_approx >= 40:
            assert got.size != plain.size or not np.array_equal(got, plain), "H_approx had no effect"
        codec.set_block_syms(8192)
        blob = codec.compress(method, 0, data).copy()
        assert np.array_equal(codec.decompress(method, 0, blob, n), data)
    finally:
        codec.set_h_approx(1)
        codec.set_block_syms(8192)
    with pytest.raises(Exception):
        codec.set_h_approx(0)


def test_empty_input(codec):
    codec.set_block_syms(8192)
    blob = codec.compress(M_FOLD, 1, np.zeros(0, np.uint32)).copy()
    assert blob.size == 64
    assert codec.decompress(M_FOLD, 1, blob, 0).size == 0


def test_domain_errors(codec):
    import ans_large_alphabet_b200 as P

    codec.set_block_syms(8192)
    bad = np.array([1, 2, 3, 1 << 30], dtype=np.uint32)
    for method, f in ((M_FOLD, 1), (M_MSB, 0), (M_RFOLD, 1)):
        with pytest.raises(P.B200AnsError) as e:
            codec.compress(method, f, bad)
        assert e.value.code == -4
    with pytest.raises(P.B200AnsError) as e:  # the reference never terminates on this one
        codec.compress(M_INT, 0, np.full(1000, 5, np.uint32))
    assert e.value.code == -4
    small = np.empty(8, dtype=np.uint8)
    with pytest.raises(P.B200AnsError) as e:
        codec.compress(M_FOLD, 1, _inputs.make("zipf16", 5000), out=small)
    assert e.value.code == -3
    good = codec.compress(M_FOLD, 1, _inputs.make("zipf16", 5000)).copy()
    with pytest.raises(P.B200AnsError):
        codec.decompress(M_MSB, 0, good, 5000)  # written by another method
    good[0] ^= 0xFF
    with pytest.raises(P.B200AnsError) as e:
        codec.decompress(M_FOLD, 1, good, 5000)
    assert e.value.code == -5


@pytest.mark.parametrize("method,f,kind", [(M_FOLD, 1, "zipf20"), (M_FOLD, 2, "zipf20"), (M_INT, 0, "geom"), (M_MSB, 0, "zipf20")])
@pytest.mark.parametrize("path", PATHS)
def test_large_roundtrip_properties(codec, method, f, kind, path):
    codec.set_kernel_path(path)
    """2^25 symbols: round trip, size = sum of parts, bits/symbol within 1 % of the whole-stream CPU coder,
    and sampled blocks byte-identical."""
    n = 1 << 25
    data = np.tile(_inputs.make(kind, 1 << 21, seed=2), 16)
    data[:: 4097] ^= 1  # break the exact periodicity a little
    codec.set_block_syms(8192)
    blob = codec.compress(method, f, data).copy()
    c = _container.parse(blob)
    back = codec.decompress(method, f, blob, n)
    assert np.array_equal(back, data)
    model = cpu().model_create(method, f, data)
    rng = np.random.default_rng(0)
    for b in rng.integers(0, c["n_blocks"], 24):
        want = model.encode_block(data[b * 8192 : (b + 1) * 8192])
        assert np.array_equal(c["blocks"][b], want)
    whole = cpu().compress(method, f, data).size
    assert blob.size <= whole * 1.01, (blob.size, whole)


@pytest.mark.parametrize("method,f", [(M_FOLD, 1), (M_FOLD, 2), (M_MSB, 0)])
def test_folded_decoder_with_a_large_frame(codec, monkeypatch, method, f):
    """A rank of a multi-GPU job codes its shard with the GLOBAL model; once the job reaches 2^32 symbols adjust_freqs
    keeps doubling the frame until a frequency nears 16 bits (M = 2^17 .. 2^19), and the 8-byte slot table no longer
    fits in shared memory.  The decoder then runs on the rank/select index (no slot table is built at all), from shared
    or from global memory, on both kernel families; the model must be the reference's for that histogram."""
    import torch

    import ans_large_alphabet_b200 as P

    host = P.host_api()
    n = 300_007
    data = _inputs.make("zipf20", n, seed=9)
    nbins = 1280 if method == M_MSB else 1 << (f + 9)
    bins = np.zeros(nbins, dtype=np.uint64)
    vals, cnts = np.unique(data, return_counts=True)
    for v, c in zip(vals, cnts):
        s = host.msb_map(int(v))[0] if method == M_MSB else host.fold_map(f, int(v))[0]
        bins[s] += np.uint64(c)
    bins *= np.uint64(1 << 15)  # the same shape, 9.8e9 symbols in total
    largest = int(np.nonzero(bins)[0][-1])
    want_nf = cpu().adjust_freqs(bins[: largest + 1], largest, True)
    codec.model_build(method, f, bins[: largest + 1])
    frame, nf = codec.model_nfreqs()
    assert np.array_equal(nf, want_nf) and frame == int(want_nf.sum()) and frame >= 1 << 15, frame
    d_in = torch.from_numpy(data.view(np.int32)).cuda()
    d_blob = torch.empty((codec.bound(method, n) + 15) // 16 * 16, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    nb = codec.encode_device(d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
    blob = d_blob[:nb].cpu().numpy()
    for path in (1, 2):
        codec.set_kernel_path(path)
        for no_rank in ("0", "1"):  # the index in shared memory / in global memory (frames whose index does not fit)
            monkeypatch.setenv("B200ANS_RANK_GLOBAL", no_rank)
            for pure in ("0", "1"):  # rank records with whole-group answers (ANSfold only; others ignore the switch)
                monkeypatch.setenv("B200ANS_PURE", pure)
                assert np.array_equal(codec.decompress(method, f, blob, n), data), (path, no_rank, pure)


@pytest.mark.parametrize("kind", ["geom", "gaps", "skewed"])
def test_plain_ans_rank_records_with_whole_group_answers(codec, monkeypatch, kind):
    """DEC_INT_SMEM_PURE (large inputs take it by themselves): groups of 32 frame slots inside one symbol answer from
    their own record.  Dense and sparse alphabets, a dominant symbol, ragged last block; same symbols as without."""
    rng = np.random.default_rng(3)
    n = 1_000_003
    if kind == "geom":
        data = _inputs.make("geom", n, seed=4)
    elif kind == "gaps":
        data = (_inputs.make("geom", n, seed=5) * 3 + 1).astype(np.uint32)  # two of three ids never occur
    else:
        data = rng.integers(0, 3000, size=n, dtype=np.uint32)
        data[rng.random(n) < 0.9] = 17
    codec.set_kernel_path(2)
    blob = codec.compress(M_INT, 0, data).copy()
    for pure in ("0", "1"):
        monkeypatch.setenv("B200ANS_PURE", pure)
        assert np.array_equal(codec.decompress(M_INT, 0, blob, n), data), pure
