#!/bin/bash
# pseudo-adaptive iteration: parity tests of the mode + stage timings (optionally under env overrides)
tag=${1:-pa}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_pseudo_adaptive.py -m gpu -x -q > $out/pytest_pa.log 2>&1; echo "pytest exit $?"; tail -5 $out/pytest_pa.log
timeout 600 python tools/pabench.py --model-syms 1024,8192,65536 --reps 3 > $out/pabench.log 2>&1; echo "pabench exit $?"; cat $out/pabench.log
for blog in 8 9; do
  echo "== BLOG $blog"; B200ANS_PA_BLOG=$blog timeout 600 python tools/pabench.py --model-syms 8192 --reps 3 2>&1 | tail -1
done
for w in 10 12; do
  echo "== WARPS $w"; B200ANS_PA_WARPS=$w timeout 600 python tools/pabench.py --model-syms 8192 --reps 3 2>&1 | tail -1
done
