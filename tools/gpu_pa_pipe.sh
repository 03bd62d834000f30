#!/bin/bash
# host-buffer decode pipeline of the pseudo-adaptive mode: parity tests, then the C4 bench line per number of runs
tag=${1:-pa_pipe}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_pseudo_adaptive.py -m gpu -x -q > $out/pytest_pa.log 2>&1; echo "pytest exit $?"; tail -5 $out/pytest_pa.log
for r in ${RUNS:-1 2 4 8}; do
  B200ANS_PA_RUNS=$r timeout 600 python bench.py --workload c4 --steps 5 --warmup 3 --no-cpu --no-targets > $out/bench_c4_runs$r.json 2> $out/bench_c4_runs$r.err
  echo "runs $r exit $?"; python - <<P
import json
d=json.load(open("$out/bench_c4_runs$r.json")); print(d["value"], d["e2e"]["value"], d["e2e"]["ms_per_step"])
P
done
