#!/bin/bash
# staged copies for pageable host buffers: parity suites that go through host buffers, then the bench lines
tag=${1:-r02ar}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_staged.py -m gpu -x -q > $out/pytest.log 2>&1
echo "pytest exit $?"; tail -3 $out/pytest.log
timeout 600 python - > $out/big_pageable.log 2>&1 <<P
import numpy as np, time, sys
sys.path.insert(0, ".")
import ans_large_alphabet_b200 as P
rng = np.random.default_rng(1)
n = 1 << 27
data = np.minimum(rng.zipf(1.2, size=n), 1 << 20).astype(np.uint32)
c = P.Codec(device=0, block_syms=8192)
for it in range(3):
    t0 = time.perf_counter(); blob = c.compress(1, 1, data); t1 = time.perf_counter(); back = c.decompress(1, 1, blob, n); t2 = time.perf_counter()
    print("pageable 2^27 fold1: enc", round(4*n/(t1-t0)/1e9,2), "GB/s dec", round(4*n/(t2-t1)/1e9,2), "GB/s ok", bool(np.array_equal(back, data)), flush=True)
import os
os.environ["B200ANS_NO_STAGING"] = "1"
for it in range(2):
    t0 = time.perf_counter(); blob2 = c.compress(1, 1, data); t1 = time.perf_counter(); back = c.decompress(1, 1, blob2, n); t2 = time.perf_counter()
    print("no staging: enc", round(4*n/(t1-t0)/1e9,2), "GB/s dec", round(4*n/(t2-t1)/1e9,2), "GB/s ok", bool(np.array_equal(back, data)), "same bytes", bool(np.array_equal(blob, blob2)), flush=True)
c.close()
P
cat $out/big_pageable.log | tail -6
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err
echo "bench exit $?"
python - <<P
import json
d=json.loads(open("$out/bench.json").read().strip().splitlines()[-1])
e=d["e2e"]
print("bench", round(d["value"],1), d["roundtrip_ok"], "e2e", round(e["value"],2), "pageable", e.get("pageable_gbs"))
P
