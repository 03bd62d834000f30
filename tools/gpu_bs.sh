#!/bin/bash
# block-size sweep of one bench workload: value, stage times and bits per symbol per block size
tag=${1:-bs}; out=gpurun_out/$tag; mkdir -p $out
for bs in ${BS:-1024 2048 4096 8192}; do
  timeout 600 python bench.py --workload ${WL:-c2} --block-syms $bs --steps 10 --warmup 3 --no-cpu --no-targets > $out/bench_${WL:-c2}_bs$bs.json 2> $out/bench_bs$bs.err
  python - <<P
import json
d=json.load(open("$out/bench_${WL:-c2}_bs$bs.json")); k=d["kernels"]
print($bs, round(d["value"],1), "enc", round(d["encode_ms"],3), "dec", round(d["decode_ms"],3), "bps", round(d["bits_per_symbol"],4), {n:round(v["ms"],3) for n,v in k.items()}, "e2e", round(d["e2e"]["value"],2))
P
done
