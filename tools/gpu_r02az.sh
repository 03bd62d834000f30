#!/bin/bash
# configs 3 (ANSfold-2 / ANSrfold-2, 10^9 symbols) with the final library
tag=${1:-r02az}
out=gpurun_out/$tag
mkdir -p $out
for wl in c3 c3r; do
timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-targets > $out/bench_$wl.json 2> $out/bench_$wl.err
python - <<P
import json
d=json.loads(open("$out/bench_$wl.json").read().strip().splitlines()[-1])
print("$wl", round(d["value"],1), round(d["ms_per_step"],3), {k:(v["kernel"],round(v["ms"],3),round(v["frac"],3)) for k,v in d["kernels"].items()}, d["roundtrip_ok"], d["parity"]["ok"], "e2e", round(d["e2e"]["value"],1), "host", round(d["host_model_ms_per_step"],3))
P
done
