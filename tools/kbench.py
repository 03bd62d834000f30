#!/usr/bin/env python
"""Kernel-level timing matrix (development aid; bench.py is the contract benchmark).
usage: python tools/kbench.py [--n 1073741824] [--workloads fold1,c2] [--paths 1,2] [--block-syms 8192]
Prints one line per (workload, path): per-kernel CUDA-event times from b200ans_get_stats and the
roofline fraction of each (algorithmic bytes / time / measured HBM peak)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402


def main():
    import torch

    import ans_large_alphabet_b200 as P

    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=0)
    ap.add_argument("--workloads", default="fold1,ans30,c2")
    ap.add_argument("--paths", default="1,2")
    ap.add_argument("--block-syms", type=int, default=0, help="0 = b200ans_suggest_block_syms(n, 8192, 16384)")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--bins-scale", type=int, default=1, help="folded coders: multiply the histogram by this before the "
                    "model is built (4 on 2^30 symbols gives the frame adjust_freqs picks for a 2^32-symbol job)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    peak, _ = bench.measured_peak()
    stream = torch.cuda.Stream(device=dev)
    codec = P.Codec(device=0, block_syms=args.block_syms or 8192, stream=stream.cuda_stream)
    for wl in args.workloads.split(","):
        method, f, gen, n, desc = bench.WORKLOADS[wl]
        if args.n:
            n = args.n
        codec.set_block_syms(args.block_syms or codec.suggest_block_syms(n, 8192, 16384))
        d_in = bench.generate_device(gen, n, dev, seed=0)
        d_blob = torch.empty((codec.bound(method, n) + 15) // 16 * 16, dtype=torch.uint8, device=dev)
        d_out = torch.empty(n, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        for path in [int(p) for p in args.paths.split(",")]:
            codec.set_kernel_path(path)
            acc = dict(hist=0.0, enc=0.0, compact=0.0, dec=0.0)
            nb = 0
            for it in range(3 + args.reps):
                with torch.cuda.stream(stream):
                    if args.bins_scale > 1:
                        nbins = 1 << (f + 9) if method == bench.M_FOLD else 1024
                        d_bins = torch.zeros(nbins, dtype=torch.int64, device=dev)
                        codec.histogram_device(method, f, d_in.data_ptr(), n, d_bins.data_ptr(), nbins)
                        stream.synchronize()
                        codec.model_build(method, f, (d_bins.cpu().numpy().astype("uint64") * args.bins_scale))
                        nb = codec.encode_device(d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
                    else:
                        nb = codec.compress_device(method, f, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
                    s = codec.stats()
                    codec.decompress_device(method, f, d_out.data_ptr(), n, d_blob.data_ptr(), nb)
                    s2 = codec.stats()
                if it >= 3:
                    acc["hist"] += s.ms_histogram / args.reps
                    acc["enc"] += s.ms_encode / args.reps
                    acc["compact"] += s.ms_compact / args.reps
                    acc["dec"] += s2.ms_decode / args.reps
            stream.synchronize()
            ok = bool(torch.equal(d_in, d_out)) if args.check or True else None
            c = s.payload_bytes / n
            alg = dict(hist=4.0 * n, enc=(4.0 + c) * n, compact=2.0 * c * n, dec=(4.0 + c) * n)
            row = {"workload": wl, "n": n, "path": {0: "auto", 1: "group", 2: "lpb"}[path], "ok": ok, "bps": 8.0 * nb / n}
            for k in acc:
                row[k + "_ms"] = round(acc[k], 4)
                row[k + "_frac"] = round(alg[k] / (acc[k] * 1e6) / peak, 4) if acc[k] > 0 else None
            print(json.dumps(row), flush=True)
        del d_in, d_blob, d_out
    codec.close()


if __name__ == "__main__":
    main()
