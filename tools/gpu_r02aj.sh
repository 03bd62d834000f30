#!/bin/bash
# slot table twice in shared memory (DEC_FOLD_SMEM2): 2^30-symbol targets with and without, parity tests
tag=${1:-r02aj}
out=gpurun_out/$tag
mkdir -p $out
for tw in 1 0; do
  B200ANS_DEC_TWICE=$tw timeout 300 python tools/kbench.py --workloads fold1,fold2,msb --paths 2 --reps 5 --check > $out/kbench_twice$tw.txt 2>&1
  echo "twice=$tw"; grep -o '"workload": "[a-z0-9]*"\|"ok": [a-z]*\|"dec_ms": [0-9.]*\|"dec_frac": [0-9.]*' $out/kbench_twice$tw.txt | paste - - - -
done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $out/pytest_parity.log 2>&1
echo "pytest parity exit $?"; tail -3 $out/pytest_parity.log
