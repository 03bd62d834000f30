"""Small round trips through every kernel family, meant to run under compute-sanitizer (tools/gpu_sanitize.sh):
memcheck for out-of-bounds / misaligned accesses, racecheck for shared-memory hazards, initcheck for reads of
uninitialised device memory.  Sizes are small because the tools slow kernels down 10-100x; the env overrides pick
the kernel generations that AUTO would only choose for large inputs."""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _inputs  # noqa: E402

import ans_large_alphabet_b200 as P  # noqa: E402

M_INT, M_FOLD, M_MSB, M_RFOLD = 0, 1, 2, 3


def round_trip(codec, method, f, data, label):
    blob = codec.compress(method, f, data).copy()
    back = codec.decompress(method, f, blob, data.size)
    assert np.array_equal(back, data), label
    print(f"ok {label}: {data.size} symbols -> {blob.size} bytes", flush=True)


def main():
    only = sys.argv[1] if len(sys.argv) > 1 else ""
    codec = P.Codec(device=0)
    n = int(os.environ.get("SAN_SYMBOLS", "300001"))
    zipf = _inputs.make("zipf20", n, seed=1)
    geom = _inputs.make("geom", n, seed=2)
    cases = []
    for path in (P.PATH_GROUP, P.PATH_LPB, P.PATH_AUTO):
        for method, f, data, name in ((M_INT, 0, geom, "ANS"), (M_FOLD, 1, zipf, "fold1"), (M_FOLD, 2, zipf, "fold2"),
                                      (M_FOLD, 4, zipf, "fold4"), (M_MSB, 0, zipf, "msb"), (M_RFOLD, 1, zipf, "rfold1"),
                                      (M_INT, 0, zipf, "ANS-zipf")):
            cases.append((f"path{path}-{name}", path, method, f, data, {}))
    # generations AUTO keeps for large inputs: pure rank records, rank index in global memory
    cases.append(("pure-ANS", P.PATH_AUTO, M_INT, 0, geom, {"B200ANS_PURE": "1"}))
    cases.append(("nopure-ANS", P.PATH_AUTO, M_INT, 0, geom, {"B200ANS_PURE": "0"}))
    cases.append(("nointdec-ANS", P.PATH_GROUP, M_INT, 0, geom, {"B200ANS_NO_INTDEC": "1"}))
    cases.append(("nopipe-fold1", P.PATH_AUTO, M_FOLD, 1, zipf, {"B200ANS_NO_PIPELINE": "1"}))
    for label, path, method, f, data, env in cases:
        if only and only not in label:
            continue
        for k, v in env.items():
            os.environ[k] = v
        for bs in (1024, 8192):
            codec.set_kernel_path(path)
            codec.set_block_syms(bs)
            round_trip(codec, method, f, data, f"{label}-bs{bs}")
        for k in env:
            del os.environ[k]
    # pseudo-adaptive mode, both table layouts
    for path in (P.PATH_AUTO, P.PATH_GROUP):
        for method, f in ((M_FOLD, 1), (M_FOLD, 2), (M_MSB, 0)):
            label = f"pa-path{path}-m{method}f{f}"
            if only and only not in label:
                continue
            codec.set_kernel_path(path)
            blob = codec.pa_compress(method, f, 1024, zipf).copy()
            assert np.array_equal(codec.pa_decompress(method, f, blob, n), zipf), label
            print(f"ok {label}", flush=True)
    codec.close()
    print("sanitize: all round trips ok")


if __name__ == "__main__":
    main()
