#!/bin/bash
tag=${1:-r02ao}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -2 $out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.log 2>&1
echo "smoke exit $?"
timeout 200 python tools/pabench.py --remap --remap-method int --model-syms 8192 --reps 2 2>/dev/null | tail -1 | tee $out/pabench_int.json | cut -c1-330
timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 > $out/bench_c4.json 2> $out/bench_c4.err
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err
echo "bench exit $?"
python - <<P
import json
for f in ("bench_c4","bench"):
    d=json.loads(open("$out/"+f+".json").read().strip().splitlines()[-1])
    print(f, round(d["value"],1), round(d["ms_per_step"],3), {k:(v["kernel"],round(v["ms"],3),round(v["frac"],3)) for k,v in d["kernels"].items()}, d["roundtrip_ok"], "e2e", round(d["e2e"]["value"],2), "launches", d["gpu_launches"])
P
