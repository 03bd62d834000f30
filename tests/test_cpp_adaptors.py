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


TOOL = os.path.join(ROOT, "apps", "b200ans_tool")


def build_tool():
    subprocess.check_call(["make", "-C", PKG, "libb200ans.so"])
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "apps", "b200ans_tool.cpp"), "-o", TOOL,
                           "-L", PKG, "-lb200ans", "-Wl,-rpath," + PKG])
    return TOOL


def test_tool_compiles_and_links():
    exe = build_tool()
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode != 0 and "usage:" in out.stderr


@pytest.mark.gpu
def test_tool_files_round_trip(tmp_path):
    """the file-level driver: .u32 / .txt readers (util.hpp:160-192), the table_efficiency / table_effectiveness
    row layout for a directory of inputs, and compress -> container on disk -> decompress"""
    import numpy as np

    import ans_large_alphabet_b200 as P

    exe = TOOL if os.path.exists(TOOL) else build_tool()
    rng = np.random.default_rng(3)
    a = (rng.geometric(0.01, size=200_003) - 1).astype(np.uint32)
    b = (rng.zipf(1.3, size=150_001) % (1 << 20)).astype(np.uint32)
    a.tofile(tmp_path / "a.u32")
    b.tofile(tmp_path / "b.u32")
    np.savetxt(tmp_path / "a.txt", a[:5000], fmt="%u")

    out = subprocess.run([exe, "effectiveness", "-i", str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    lines = out.stdout.splitlines()
    assert lines[0] == "ANS  &" and lines[1].endswith("&") and lines[2].endswith("\\\\ ")
    codec = P.Codec(device=0)
    want = 8.0 * codec.compress(P.METHOD_INT, 0, a).size / a.size
    assert abs(float(lines[1].split()[0]) - want) < 1e-4
    codec.close()

    out = subprocess.run([exe, "efficiency", "-i", str(tmp_path / "b.u32")], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "\\method{ANSfold-1}  &" in out.stdout and "ERROR" not in out.stderr

    for method, src, text in (("ANSfold-1", "b.u32", False), ("ANS", "a.txt", True), ("ANSmsb", "a.u32", False)):
        blob, back = tmp_path / "x.b2an", tmp_path / "x.u32"
        cmd = [exe, "compress", "-m", method, "-i", str(tmp_path / src), "-o", str(blob)] + (["-t"] if text else [])
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, out.stdout + out.stderr
        info = P.host_api().container_info(np.fromfile(blob, dtype=np.uint8))
        expect = a[:5000] if text else (a if src == "a.u32" else b)
        assert info["n_symbols"] == expect.size and info["total_bytes"] == os.path.getsize(blob)
        out = subprocess.run([exe, "decompress", "-i", str(blob), "-o", str(back)], capture_output=True, text=True, timeout=600)
        assert out.returncode == 0, out.stdout + out.stderr
        assert method in out.stdout
        assert np.array_equal(np.fromfile(back, dtype=np.uint32), expect)
