#!/bin/bash
# the driver's round-end sequence: GPU suite, smoke, reference arm, default bench
tag=${1:-r02final}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -3 $out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke.log 2>&1
echo "smoke exit $?"; tail -2 $out/smoke.log
timeout 900 python bench.py --impl reference > $out/bench_ref.json 2> $out/bench_ref.err
echo "reference arm exit $?"; tail -c 600 $out/bench_ref.json
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err
echo "bench exit $?"
python - <<P
import json
d=json.loads(open("$out/bench.json").read().strip().splitlines()[-1])
print(d["metric"], d["value"], d["unit"], d["ms_per_step"], d["config"]["workload"][:60])
print("roofline", d["roofline"]["kernel"], d["roofline"]["frac"], d["roofline"]["traffic"], d["roofline"].get("traffic_over_algorithmic"))
print("e2e", d["e2e"]["value"], d["e2e"].get("copy_only_gbs"), d["e2e"].get("copy_floor_gbs"))
print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["kind"], "launches", d["gpu_launches"], "clocks", d["clocks"])
print("targets", {k:(round(v["kernels"]["encode"]["frac"],3), round(v["kernels"]["decode"]["frac"],3)) for k,v in d.get("north_star_2p30",{}).items() if isinstance(v,dict) and "kernels" in v})
P
