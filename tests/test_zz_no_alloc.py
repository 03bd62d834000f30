"""-m gpu, collected last on purpose: b200ans_decode_blocks_noalloc -- the decode entry
point that allocates nothing (per-block offsets in a file-scope device array, include/b200ans.h) -- against the
existing entry point b200ans_decompress_device, through the C ABI.  Bar: bit-exact (integer work), on two calls in a row
that share the static scratch (the second with fewer blocks, another method and another block size, so stale contents of
the first would show), then the size limit."""
import numpy as np
import pytest

import _container
import _inputs
from _libs import M_FOLD, M_INT

pytestmark = pytest.mark.gpu

MAX_BLOCKS = 65536  # B200ANS_NOALLOC_MAX_BLOCKS


def _both_ways(codec, method, f, data, block_syms, free_must_hold):
    """container by the existing path; decode it by b200ans_decompress_device and by the no-alloc entry point"""
    import torch

    n = data.size
    codec.set_block_syms(block_syms)
    blob = codec.compress(method, f, data).copy()
    c = _container.parse(blob)
    table_at = 64 + (c["prelude"].size + 3) // 4 * 4
    payload_at = c["payload_at"]
    d_blob = torch.zeros((blob.size + 15) // 16 * 16 + 64, dtype=torch.uint8, device="cuda")
    d_blob[: blob.size] = torch.from_numpy(blob).cuda()
    d_ref = torch.zeros(n, dtype=torch.int32, device="cuda")
    d_new = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    codec.decompress_device(method, f, d_ref.data_ptr(), n, d_blob.data_ptr(), blob.size)  # existing entry point
    # set-up (may allocate): the model of this container; then the call under test, which may not
    codec.model_load(method, f, c["prelude"])
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    codec.decode_blocks_noalloc(d_blob.data_ptr() + payload_at, c["payload_bytes"], d_blob.data_ptr() + table_at, n, d_new.data_ptr())
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    if free_must_hold:
        assert free1 >= free0, f"device memory shrank by {free0 - free1} bytes across a call that must not allocate"
    assert torch.equal(d_new, d_ref), "no-alloc decode differs from b200ans_decompress_device"
    assert np.array_equal(d_new.cpu().numpy().view(np.uint32), data)
    return c["n_blocks"]


def test_noalloc_decode_matches_existing_entry_point_twice():
    import ans_large_alphabet_b200 as P

    codec = P.Codec(device=0, block_syms=8192)
    try:
        # call 1: ANSfold-1, 2 000 003 symbols in blocks of 1024 -> 1954 offsets in the static array (ragged last block)
        big = _inputs.make("zipf20", 2_000_003, seed=31)
        nb1 = _both_ways(codec, M_FOLD, 1, big, 1024, free_must_hold=False)  # (first use resolves the symbol address)
        # call 2, right behind it: plain ANS, fewer and larger blocks -- entries of call 1 beyond them are still there
        small = _inputs.make("geom", 300_001, seed=32)
        nb2 = _both_ways(codec, M_INT, 0, small, 8192, free_must_hold=True)
        assert nb2 < nb1
        # and the first again (more blocks than the call before): the array is rewritten, not appended to
        _both_ways(codec, M_FOLD, 1, big, 1024, free_must_hold=True)
    finally:
        codec.close()


def test_noalloc_decode_rejects_what_it_cannot_hold():
    import torch

    import ans_large_alphabet_b200 as P

    codec = P.Codec(device=0, block_syms=128)
    try:
        data = _inputs.make("zipf16", 50_000, seed=33)
        blob = codec.compress(M_FOLD, 1, data).copy()
        c = _container.parse(blob)
        d = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")
        out = torch.zeros(16, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        with pytest.raises(P.B200AnsError) as e:  # no model yet
            codec.decode_blocks_noalloc(d.data_ptr(), d.numel(), d.data_ptr(), 1000, out.data_ptr())
        assert e.value.code == -6
        codec.model_load(M_FOLD, 1, c["prelude"])
        too_many = 128 * (MAX_BLOCKS + 1)  # one block more than the static scratch holds; refused before anything is read
        with pytest.raises(P.B200AnsError) as e:
            codec.decode_blocks_noalloc(d.data_ptr(), d.numel(), d.data_ptr(), too_many, out.data_ptr())
        assert e.value.code == -1 and "65536" in str(e.value)
        with pytest.raises(P.B200AnsError) as e:  # sizes that add up to more than the payload the caller vouches for
            d_blob = torch.zeros((blob.size + 15) // 16 * 16 + 64, dtype=torch.uint8, device="cuda")
            d_blob[: blob.size] = torch.from_numpy(blob).cuda()
            d_out = torch.zeros(data.size, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            table_at = 64 + (c["prelude"].size + 3) // 4 * 4
            codec.decode_blocks_noalloc(d_blob.data_ptr() + c["payload_at"], c["payload_bytes"] - 16, d_blob.data_ptr() + table_at,
                                        data.size, d_out.data_ptr())
        assert e.value.code == -5
        # the context still works
        codec.decode_blocks_noalloc(d_blob.data_ptr() + c["payload_at"], c["payload_bytes"], d_blob.data_ptr() + table_at, data.size,
                                    d_out.data_ptr())
        torch.cuda.synchronize()
        assert np.array_equal(d_out.cpu().numpy().view(np.uint32), data)
    finally:
        codec.close()
