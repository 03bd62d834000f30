"""The C++ drop-in boundary: include/b200ans_methods.hpp (adaptor structs with the reference's methods.hpp
signatures + the ans_util.hpp model API) compiled into a table_efficiency-style driver.

not gpu: the driver compiles and links against the in-tree C-ABI library.
gpu:     it runs -- every adaptor round-trips through encode()/decode() on the GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ans_large_alphabet_b200")
EXE = os.path.join(ROOT, "tests", "cpp", "table_efficiency_b200")


def build():
    subprocess.check_call(["make", "-C", PKG, "libb200ans.so"])
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "table_efficiency_b200.cpp"), "-o", EXE,
                           "-L", PKG, "-lb200ans", "-Wl,-rpath," + PKG])
    return EXE


def test_adaptors_compile_and_link():
    exe = build()
    assert os.path.exists(exe)
    syms = subprocess.run(["nm", "-D", "--undefined-only", exe], capture_output=True, text=True).stdout
    for s in ("b200ans_compress", "b200ans_decompress", "b200ans_adjust_freqs", "b200ans_serialize_interp", "b200ans_load_interp"):
        assert s in syms, s


@pytest.mark.gpu
def test_adaptors_run_like_table_efficiency():
    exe = EXE if os.path.exists(EXE) else build()
    out = subprocess.run([exe, "3000017"], capture_output=True, text=True, timeout=600)
    print(out.stdout, out.stderr)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ALL OK" in out.stdout and "MISMATCH" not in out.stdout
    for name in ("ANS ", "ANSmsb", "ANSfold-1", "ANSfold-2", "ANSrfold-1"):
        assert name in out.stdout
