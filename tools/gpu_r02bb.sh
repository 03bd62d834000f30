#!/bin/bash
out=gpurun_out/${1:-r02bb}; mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_pseudo_adaptive.py -m gpu -x -q 2>&1 | tail -1
timeout 300 python - <<P
import numpy as np, time, sys, os
sys.path.insert(0, ".")
import ans_large_alphabet_b200 as P
rng = np.random.default_rng(3)
n = (1 << 26) + 12345
data = np.minimum(rng.zipf(1.2, size=n), 1 << 20).astype(np.uint32)
c = P.Codec(device=0)
t0 = time.perf_counter(); blob = c.pa_compress(1, 1, 8192, data).copy(); t1 = time.perf_counter()
back = c.pa_decompress(1, 1, blob, n)
print("pa pageable staged: enc", round(4*n/(t1-t0)/1e9,2), "GB/s ok", bool(np.array_equal(back, data)))
os.environ["B200ANS_NO_STAGING"] = "1"
t0 = time.perf_counter(); blob2 = c.pa_compress(1, 1, 8192, data).copy(); t1 = time.perf_counter()
print("plain: enc", round(4*n/(t1-t0)/1e9,2), "GB/s same bytes", bool(np.array_equal(blob, blob2)))
c.close()
P
