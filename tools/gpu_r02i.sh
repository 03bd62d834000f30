#!/bin/bash
# r02i: group records for large-frame folded decoders (A/B on the C5 model) + the large-frame parity test
tag=${1:-r02i}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -k "frame or alloc" > $out/pytest.log 2>&1; echo "tests exit $?"; tail -3 $out/pytest.log
for g in 0 1; do
  B200ANS_GROUP=$g timeout 300 python tools/kbench.py --workloads c5w --paths 2 --reps 3 --bins-scale 4 > $out/kbench_c5m_g$g.txt 2>&1; echo "group=$g"; tail -1 $out/kbench_c5m_g$g.txt
done
B200ANS_GROUP=1 timeout 300 python tools/kbench.py --workloads fold1 --paths 2 --reps 3 --bins-scale 16 > $out/kbench_fold1_big.txt 2>&1; tail -1 $out/kbench_fold1_big.txt
