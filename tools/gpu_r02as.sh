#!/bin/bash
# staging threads for pageable host buffers: 2 / 4 / 8
tag=${1:-r02as}
out=gpurun_out/$tag
mkdir -p $out
for t in 2 4 8; do
B200ANS_STAGE_THREADS=$t timeout 300 python - <<P
import numpy as np, time, sys
sys.path.insert(0, ".")
import ans_large_alphabet_b200 as P
rng = np.random.default_rng(1)
n = 1 << 28
data = np.minimum(rng.zipf(1.2, size=n), 1 << 20).astype(np.uint32)
c = P.Codec(device=0, block_syms=8192)
blob = np.zeros(3 * n + (64 << 20), dtype=np.uint8)
back = np.zeros(n, dtype=np.uint32)
best = (0, 0)
for it in range(3):
    t0 = time.perf_counter(); nb = c.compress_ptr(1, 1, blob.ctypes.data, blob.size, data.ctypes.data, n); t1 = time.perf_counter()
    c.decompress_ptr(1, 1, back.ctypes.data, n, blob.ctypes.data, nb); t2 = time.perf_counter()
    best = max(best, (4*n/(t1-t0)/1e9, 4*n/(t2-t1)/1e9))
print("threads $t: enc", round(best[0],2), "GB/s dec", round(best[1],2), "GB/s both", round(8*n/(4*n/best[0]+4*n/best[1])/1e9*1e9/1e9,2), "ok", bool(np.array_equal(back, data)), flush=True)
c.close()
P
done
