"""-m gpu: host-buffer calls on PAGEABLE memory (what the reference's callers hold: std::vector, methods.hpp:535-546).
Large copies go through the library's page-locked staging buffers on a few host threads (b200ans.cu staged_h2d /
staged_d2h); the bytes must not depend on it."""
import numpy as np
import pytest

import _inputs
from _libs import M_FOLD, M_INT, M_MSB

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("method,f", [(M_FOLD, 1), (M_INT, 0), (M_MSB, 0)])
@pytest.mark.parametrize("threads", [1, 3, 8])
def test_staged_copies_give_the_same_bytes(codec, monkeypatch, method, f, threads):
    n = (1 << 22) + 4321  # 16 MiB of symbols: several 8 MiB chunks, the last one ragged
    data = _inputs.make("geom" if method == M_INT else "zipf20", n, seed=9)
    monkeypatch.setenv("B200ANS_NO_STAGING", "1")
    want = codec.compress(method, f, data).copy()
    monkeypatch.delenv("B200ANS_NO_STAGING")
    monkeypatch.setenv("B200ANS_STAGE_MIN_MB", "1")
    monkeypatch.setenv("B200ANS_STAGE_THREADS", str(threads))
    got = codec.compress(method, f, data).copy()
    assert np.array_equal(got, want)
    assert np.array_equal(codec.decompress(method, f, got, n), data)
    if method != M_INT:
        blob = codec.pa_compress(method, f, 8192, data).copy()
        monkeypatch.setenv("B200ANS_NO_STAGING", "1")
        assert np.array_equal(codec.pa_compress(method, f, 8192, data), blob)
        assert np.array_equal(codec.pa_decompress(method, f, blob, n), data)
