"""not gpu: the C-ABI library loads and exports every symbol include/b200ans.h declares; without a GPU
coding calls fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200ans.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200ans_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    import ans_large_alphabet_b200 as P
    from ans_large_alphabet_b200 import _capi

    lib = P.load_library()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200ans.h but not exported by libb200ans.so"
        assert n in _capi.SIGNATURES, f"{n} has no ctypes signature"
    assert set(_capi.SIGNATURES) <= set(names), "ctypes binds symbols the header does not declare"


def test_no_oracle_in_product():
    """the product library must not link or load anything under oracle/"""
    import subprocess

    import ans_large_alphabet_b200 as P

    out = subprocess.run(["ldd", P.library_path()], capture_output=True, text=True).stdout
    assert "ansoracle" not in out and "ansref" not in out
    for root, _, files in os.walk(os.path.join(ROOT, "ans_large_alphabet_b200")):
        for f in files:
            if f.endswith((".py", ".hpp", ".cu", ".cuh", ".h")):
                src = open(os.path.join(root, f)).read()
                assert "oracle/" not in src.replace("oracle/ ", "") or f == "__init__.py", f"{f} mentions oracle/"
                assert "libansoracle" not in src and "libansref" not in src


def test_fails_loudly_without_gpu():
    import torch

    import ans_large_alphabet_b200 as P

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(P.B200AnsError) as e:
        P.Codec(device=0)
    assert e.value.code == -2  # B200ANS_E_CUDA: there is no CPU coding path
    assert "no CPU fallback" in str(e.value)


def test_strerror_and_bound():
    import ans_large_alphabet_b200 as P

    lib = P.load_library()
    assert lib.b200ans_strerror(0) == b"ok"
    assert b"domain" in lib.b200ans_strerror(-4)
    assert lib.b200ans_compress_bound(None, 1, 1 << 20) > 7 * (1 << 20)


def test_suggested_block_size_fills_whole_rounds():
    """b200ans_suggest_block_syms is host arithmetic (no GPU): blocks of the chosen size fill whole rounds of
    148 SMs x 16 warps x 32 blocks, ties go to the larger block, small inputs get the smallest size allowed"""
    import ans_large_alphabet_b200 as P

    slots = 148 * 16

    def cost(n, bs):
        tasks = ((n + bs - 1) // bs + 31) // 32
        return max(1, (tasks + slots - 1) // slots) * bs

    assert P.suggest_block_syms(1 << 30, 8192, 16384) == 14208  # one round of 16 warps per SM
    for n in (1 << 30, 1 << 32, 10**9, 3 * 10**8 + 17):
        for lo, hi in ((8192, 16384), (4096, 8192), (8192, 8192)):
            bs = P.suggest_block_syms(n, lo, hi)
            assert lo <= bs <= hi and bs % 128 == 0
            assert cost(n, bs) == min(cost(n, b) for b in range(lo, hi + 1, 128))
            assert all(cost(n, b) > cost(n, bs) for b in range(bs + 128, hi + 1, 128)), "a larger block ties"
    assert P.suggest_block_syms(10**7, 8192, 16384) == 8192  # one round whatever the size: most parallelism
    assert P.suggest_block_syms(1 << 30, 100, 1000) == 128 * 7 or P.suggest_block_syms(1 << 30, 100, 1000) % 128 == 0
    assert P.suggest_block_syms(1 << 30, 16384, 8192) == 0
