#!/usr/bin/env python
"""Pseudo-adaptive mode timing (config C4's per-GPU shard): per stage times and bits/symbol next to the semi-static coder.
usage: python tools/pabench.py [--n 134217728] [--model-syms 8192,16384,65536] [--method fold1]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402


def main():
    import torch

    import ans_large_alphabet_b200 as P

    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 27)
    ap.add_argument("--model-syms", default="8192,16384,65536")
    ap.add_argument("--workload", default="fold1")
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    codec = P.Codec(device=0, block_syms=8192, stream=stream.cuda_stream)
    method, f, gen, _, desc = bench.WORKLOADS[args.workload]
    n = args.n
    d_in = bench.generate_device(gen, n, dev, seed=0)
    d_out = torch.empty(n, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        cap = codec.bound(method, n)
        d_blob = torch.empty((cap + 15) // 16 * 16, dtype=torch.uint8, device=dev)
        semi = codec.compress_device(method, f, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
    stream.synchronize()
    for S in [int(x) for x in args.model_syms.split(",")]:
        cap = codec.pa_bound(method, S, n)
        d_blob = torch.empty((cap + 15) // 16 * 16, dtype=torch.uint8, device=dev)
        best = None
        for it in range(1 + args.reps):
            with torch.cuda.stream(stream):
                stream.synchronize()
                t0 = time.perf_counter()
                nb = codec.pa_compress_device(method, f, S, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
                stream.synchronize()
                t1 = time.perf_counter()
                s = codec.stats()
                codec.pa_decompress_device(method, f, d_out.data_ptr(), n, d_blob.data_ptr(), nb)
                stream.synchronize()
                t2 = time.perf_counter()
                s2 = codec.stats()
            row = dict(model_syms=S, n=n, models=(n + S - 1) // S, bits_per_symbol=8.0 * nb / n, semi_static_bits_per_symbol=8.0 * semi / n,
                       enc_wall_ms=(t1 - t0) * 1e3, dec_wall_ms=(t2 - t1) * 1e3, hist_ms=s.ms_histogram, enc_host_model_ms=s.ms_model,
                       enc_kernel_ms=s.ms_encode, compact_ms=s.ms_compact, dec_host_model_ms=s2.ms_model, dec_kernel_ms=s2.ms_decode,
                       enc_gbs=4.0 * n / (t1 - t0) / 1e9, dec_gbs=4.0 * n / (t2 - t1) / 1e9)
            if it and (best is None or row["enc_wall_ms"] + row["dec_wall_ms"] < best["enc_wall_ms"] + best["dec_wall_ms"]):
                best = row
        best["ok"] = bool(torch.equal(d_in, d_out))
        best["workload"] = desc
        print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in best.items()}), flush=True)
    codec.close()


if __name__ == "__main__":
    main()
