#!/usr/bin/env python
"""How the ANSrfold pre-pass (raw histogram + selection) scales with the value range: 1e9 Zipf(1.0) symbols over
2^20 .. 2^24 values, and with the hot values scattered over the range instead of sitting at its bottom."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench
import ans_large_alphabet_b200 as P

dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
codec = P.Codec(device=0, block_syms=13312, stream=stream.cuda_stream)
n = int(os.environ.get("N", 1_000_000_000))
for lg, scatter in ((20, 0), (22, 0), (23, 0), (24, 0), (24, 1)):
    d_in = bench.generate_device(("zipf", 1 << lg, 1.0, 0), n, dev, seed=0)
    if scatter:
        d_in = ((d_in.to(torch.int64) * 0x9E3779B1) & ((1 << lg) - 1)).to(torch.int32)
    d_blob = torch.empty((codec.bound(3, n) + 15) // 16 * 16, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    for it in range(3):
        with torch.cuda.stream(stream):
            nb = codec.compress_device(3, 2, d_blob.data_ptr(), d_blob.numel(), d_in.data_ptr(), n)
            s = codec.stats()
    print(f"values 2^{lg} scatter={scatter}: model pass {s.ms_histogram:.3f} ms, encode {s.ms_encode:.3f} ms ({s.encode_kernel.decode()}), bps {8.0 * nb / n:.4f}", flush=True)
    del d_in, d_blob
