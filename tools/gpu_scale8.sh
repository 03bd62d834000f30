#!/bin/bash
# the default bench (config 5: ONE 2^32-symbol job) at N = 8 and N = 4 on one box
out=gpurun_out/${1:-r02m}; mkdir -p $out
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 --no-targets > $out/bench_c5_n$n.json 2> $out/bench_c5_n$n.err; echo "N=$n exit $?"; tail -2 $out/bench_c5_n$n.err
  python - <<PY
import json
try:
    d=json.loads(open("$out/bench_c5_n$n.json").read().strip().splitlines()[-1])
    print("N=$n", round(d["value"],1), "GB/s", round(d["ms_per_step"],3), "ms", d["scaling"], "e2e", round(d["e2e"]["value"],1), "allreduce", d.get("allreduce_ms_per_step"), "host", d.get("host_model_ms_per_step"), {k:(round(v["ms"],3),round(v["frac"],3)) for k,v in d["kernels"].items()}, d.get("roundtrip_ok"), (d.get("parity") or {}).get("ok"))
except Exception as e: print("ERR", e)
PY
done
