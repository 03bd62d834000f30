#!/bin/bash
tag=${1:-r02al}
out=gpurun_out/$tag
mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_pa_remap.py tests/test_gpu_pseudo_adaptive.py -m gpu -x -q > $out/pytest.log 2>&1
echo "pytest exit $?"; tail -3 $out/pytest.log
timeout 200 python tools/pabench.py --remap --remap-method int --model-syms 8192 --reps 2 2>/dev/null | tee $out/pabench_int.json | cut -c1-400
timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu > $out/bench_c4.json 2> $out/bench_c4.err
python - <<P
import json
d=json.loads(open("$out/bench_c4.json").read().strip().splitlines()[-1])
print("c4", d["value"], d["ms_per_step"], d["encode_ms"], d["decode_ms"], {k:(v["kernel"],round(v["ms"],3)) for k,v in d["kernels"].items()}, d["roundtrip_ok"])
P
