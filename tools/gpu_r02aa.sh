#!/bin/bash
# full GPU suite + C4 / default bench lines after the pseudo-adaptive work of r02v..r02z
tag=${1:-r02aa}
out=gpurun_out/$tag
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -4 $out/pytest_gpu.log
timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 > $out/bench_c4.json 2> $out/bench_c4.err
echo "bench c4 exit $?"
python - <<P
import json
d=json.loads(open("$out/bench_c4.json").read().strip().splitlines()[-1])
print("c4", d["value"], d["ms_per_step"], d["encode_ms"], d["decode_ms"], {k:(v["kernel"],round(v["ms"],3),round(v["frac"],3)) for k,v in d["kernels"].items()}, d["roundtrip_ok"], d["parity"], d["e2e"]["value"], d.get("driver_remap",{}).get("encode_ms"), d.get("driver_remap",{}).get("decode_ms"))
P
timeout 600 python bench.py > $out/bench_c5.json 2> $out/bench_c5.err
echo "bench default exit $?"
python - <<P
import json
d=json.loads(open("$out/bench_c5.json").read().strip().splitlines()[-1])
print("c5", d["value"], d["ms_per_step"], {k:(v["kernel"],round(v["ms"],3),round(v["frac"],3)) for k,v in d["kernels"].items()}, d["roundtrip_ok"], d["parity"]["ok"], d["e2e"]["value"])
P
timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 > $out/bench_c2.json 2> $out/bench_c2.err
python - <<P
import json
d=json.loads(open("$out/bench_c2.json").read().strip().splitlines()[-1])
print("c2", d["value"], d["ms_per_step"], {k:(v["kernel"],round(v["ms"],3),round(v["frac"],3)) for k,v in d["kernels"].items()}, d["roundtrip_ok"], d["parity"]["ok"], d["e2e"]["value"])
P
