"""-m gpu: ANSrfold with the model built on the device ("K4"; ans_reorder_fold.hpp:75-115):
radix select of the 2^(f+7) most frequent values with the reference's tie-break, the value reordering, the folded
histogram derived from the raw one, and the L3 encoder that looks the hot values up through a perfect hash in shared
memory.  Bar: bit-exact -- prelude (which carries most_frequent[]) and every block payload equal the reference's."""
import numpy as np
import pytest

import _container
import _inputs
from _libs import M_RFOLD, have_reference, oracle, reference

pytestmark = pytest.mark.gpu


def cpu():
    return reference() if have_reference() else oracle()


def scattered_zipf(n, log2_alphabet, seed):
    """Zipf(1.0) ranks pushed through a bijection of [0, 2^log2_alphabet): the hot values lie anywhere in the range"""
    r = _inputs.zipf(n, 1 << log2_alphabet, 1.0, seed, offset=0).astype(np.uint64)
    return ((r * np.uint64(0x9E3779B1)) & np.uint64((1 << log2_alphabet) - 1)).astype(np.uint32)


def ties_at_the_boundary(f, seed, tie_mult=3):
    """far more values share the K-th largest count than fit: the K selected ones must be the smallest values"""
    K = 1 << (f + 7)
    rng = np.random.default_rng(seed)
    hot = rng.permutation(5 * K)[: K // 2].astype(np.uint32) + 7      # K/2 values, 9 times each
    tie = (rng.permutation(40 * K)[: tie_mult * K].astype(np.uint32) + 5 * K + 7)  # 3K values, 4 times each: K/2 of them get in
    cold = np.arange(50 * K, 50 * K + 4 * K, dtype=np.uint32)           # once each
    data = np.concatenate([np.repeat(hot, 9), np.repeat(tie, 4), cold])
    rng.shuffle(data)
    return data


def host_prelude(codec, f, data):
    """the staged HOST route: raw bins on the host -> b200ans_rfold_prepare -> folded histogram over the input -> model"""
    import torch

    raw = np.bincount(data, minlength=int(data.max()) + 1).astype(np.uint64)
    codec.rfold_prepare(f, raw)
    nb = 1 << (f + 9)
    d_in = torch.from_numpy(data.view(np.int32)).cuda()
    d_bins = torch.zeros(nb, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    codec.histogram_device(M_RFOLD, f, d_in.data_ptr(), data.size, d_bins.data_ptr(), nb)
    torch.cuda.synchronize()
    bins = d_bins.cpu().numpy().view(np.uint64)
    codec.model_build(M_RFOLD, f, bins)
    return codec.model_prelude().copy(), bins


@pytest.mark.parametrize("f", [1, 2, 3, 5, 8])
@pytest.mark.parametrize("kind", ["scattered", "ties", "zipf20"])
def test_device_selection_equals_host_and_reference(codec, f, kind):
    import torch

    if kind == "scattered":
        data = scattered_zipf(400_003, 22, seed=f)
    elif kind == "ties":
        data = ties_at_the_boundary(f, seed=f)
    else:
        data = _inputs.make("zipf20", 300_001, seed=f)
    if np.unique(data).size < (1 << (f + 7)):
        pytest.skip("fewer distinct values than 2^(f+7): covered by the identity-mapping test")
    want_pre, want_bins = host_prelude(codec, f, data)
    # device route
    nvalues = int(data.max()) + 1 + 1000  # padded range: trailing zero counts must not matter
    d_raw = torch.zeros(nvalues, dtype=torch.int64, device="cuda")
    d_raw[: int(data.max()) + 1] = torch.from_numpy(np.bincount(data).astype(np.int64)).cuda()
    d_fold = torch.full((1 << (f + 9),), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    codec.rfold_prepare_device(f, d_raw.data_ptr(), nvalues, d_fold.data_ptr())
    torch.cuda.synchronize()
    got_bins = d_fold.cpu().numpy().view(np.uint64)
    assert np.array_equal(got_bins, want_bins), "folded histogram derived from the raw bins differs"
    codec.model_build(M_RFOLD, f, got_bins)
    got_pre = codec.model_prelude().copy()
    assert np.array_equal(got_pre, want_pre), "most_frequent[] / prelude differ from the host selection"
    ref_pre = cpu().model_create(M_RFOLD, f, data).serialize()
    assert _container.prelude_equal(got_pre, ref_pre), "prelude differs from the reference"


@pytest.mark.parametrize("f,kind,n,path", [
    (1, "scattered", 600_011, 2), (2, "scattered", 600_011, 2), (3, "scattered", 600_011, 2),  # L3 encoder + hash
    (2, "ties", 0, 2), (2, "scattered", 600_011, 1), (5, "scattered", 300_007, 2), (2, "zipf20", 70_001, 2),
])
def test_blocks_byte_identical(codec, f, kind, n, path):
    data = scattered_zipf(n, 22, seed=3) if kind == "scattered" else ties_at_the_boundary(f, 5) if kind == "ties" else _inputs.make(kind, n, seed=3)
    if np.unique(data).size < (1 << (f + 7)):
        pytest.skip("fewer distinct values than 2^(f+7)")
    codec.set_kernel_path(path)
    try:
        for bs in (8192, 1920):
            codec.set_block_syms(bs)
            blob = codec.compress(M_RFOLD, f, data).copy()
            c = _container.parse(blob)
            model = cpu().model_create(M_RFOLD, f, data)
            assert _container.prelude_equal(c["prelude"], model.serialize())
            for b, payload in enumerate(c["blocks"]):
                want = model.encode_block(data[b * bs : (b + 1) * bs])
                assert np.array_equal(payload, want), f"block {b}"
            assert np.array_equal(codec.decompress(M_RFOLD, f, blob, data.size), data)
    finally:
        codec.set_kernel_path(0)
        codec.set_block_syms(8192)


def test_large_alphabet_round_trip(codec):
    """2^25 symbols over a 2^24-value range (config 3's alphabet): round trip, sampled blocks against the reference,
    size within 1 % of the reference's whole stream"""
    n = 1 << 25
    data = scattered_zipf(n, 24, seed=1)
    codec.set_block_syms(8192)
    blob = codec.compress(M_RFOLD, 2, data).copy()
    assert codec.stats().encode_kernel == b"lpb3_encode_kernel"
    c = _container.parse(blob)
    assert np.array_equal(codec.decompress(M_RFOLD, 2, blob, n), data)
    model = cpu().model_create(M_RFOLD, 2, data)
    assert _container.prelude_equal(c["prelude"], model.serialize())
    rng = np.random.default_rng(0)
    for b in rng.integers(0, c["n_blocks"], 24):
        assert np.array_equal(c["blocks"][b], model.encode_block(data[b * 8192 : (b + 1) * 8192]))
    whole = cpu().compress(M_RFOLD, 2, data).size
    assert blob.size <= whole * 1.01, (blob.size, whole)


def test_identity_mapping_and_its_domain(codec):
    """fewer than 2^(f+7) distinct values: identity mapping (ans_reorder_fold.hpp:94-97); valid only below 2^(f+7)"""
    import ans_large_alphabet_b200 as P

    rng = np.random.default_rng(2)
    data = rng.integers(0, 200, 50_000, dtype=np.uint32)  # 200 distinct values < 256
    blob = codec.compress(M_RFOLD, 1, data).copy()
    want = cpu().model_create(M_RFOLD, 1, data).serialize()
    assert _container.prelude_equal(_container.parse(blob)["prelude"], want)
    assert np.array_equal(codec.decompress(M_RFOLD, 1, blob, data.size), data)
    data[7] = 100_000
    with pytest.raises(P.B200AnsError) as e:
        codec.compress(M_RFOLD, 1, data)
    assert e.value.code == -4


@pytest.mark.parametrize("f,kind,n,expect_fast", [
    (2, "scattered", 3_000_017, True), (1, "scattered", 1_500_001, True), (2, "zipf20", 2_000_003, True), (3, "scattered", 6_000_011, True),
    (2, "ties", 0, True),        # every value is a candidate (3840 distinct): exact counts, ties by value
    (2, "ties30", 0, False),     # 15 360 values share the K-th count: the bound cannot prove the selection
    (2, "flat", 400_000, False),  # uniform values: no candidate stands out
])
def test_fast_route_is_exact_or_declines(codec, monkeypatch, f, kind, n, expect_fast):
    """B200ANS_MODEL_RFOLD_FAST: candidates from a sample, one pass with exact candidate counts and a hashed upper bound
    for everything else; the route answers only when the bound proves the selection -- the container must equal the one
    the raw histogram gives, byte for byte, either way"""
    if kind == "scattered":
        data = scattered_zipf(n, 22, seed=9)
    elif kind == "ties":
        data = ties_at_the_boundary(f, 7)
    elif kind == "ties30":
        data = ties_at_the_boundary(f, 7, tie_mult=30)
    elif kind == "flat":
        data = np.random.default_rng(4).integers(0, 1 << 16, n, dtype=np.uint32)
    else:
        data = _inputs.make(kind, n, seed=9)
    codec.set_block_syms(8192)
    monkeypatch.setenv("B200ANS_RF_FAST", "0")
    want = codec.compress(M_RFOLD, f, data).copy()
    assert codec.stats().model_flags & 2 == 0
    monkeypatch.setenv("B200ANS_RF_FAST", "1")
    got = codec.compress(M_RFOLD, f, data).copy()
    assert bool(codec.stats().model_flags & 2) == expect_fast
    assert np.array_equal(got, want)
    assert np.array_equal(codec.decompress(M_RFOLD, f, got, data.size), data)
    ref_pre = cpu().model_create(M_RFOLD, f, data).serialize()
    assert _container.prelude_equal(_container.parse(got)["prelude"], ref_pre)
