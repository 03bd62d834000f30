#!/bin/bash
# r02h: L3 decoder with TMA stores (A/B) + parity of the lane-per-block paths
tag=${1:-r02h}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -k "lpb or large or alloc or frame or rank" > $out/pytest_parity.log 2>&1; echo "parity exit $?"; tail -3 $out/pytest_parity.log
for ts in 0 1; do
  B200ANS_L3_TSTORE=$ts timeout 600 python tools/kbench.py --workloads fold1,fold2,ans30 --paths 2 --reps 3 > $out/kbench_ts$ts.txt 2>&1; echo "tstore=$ts"; cut -c1-400 $out/kbench_ts$ts.txt | tail -3
  B200ANS_L3_TSTORE=$ts timeout 300 python tools/kbench.py --workloads c5w --paths 2 --reps 3 --bins-scale 4 > $out/kbench_c5m_ts$ts.txt 2>&1; tail -1 $out/kbench_c5m_ts$ts.txt
done
