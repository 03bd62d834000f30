#!/bin/bash
# Next-round experiment: the encoder with up to 22 warps per SM (B200ANS_L3_ENC_WIDE=1, lpb3_kernels.cuh L3CfgWide).
# Parity first (the byte-identical block tests with the wide shape forced), then stage times next to the default.
tag=${1:-wide}; out=gpurun_out/$tag; mkdir -p $out
B200ANS_L3_ENC_WIDE=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "byte_identical or block_sizes or large_roundtrip" > $out/pytest_wide.log 2>&1
echo "pytest (wide) exit $?"; tail -3 $out/pytest_wide.log
for wide in 0 1; do
  for wl in ${WLS:-fold1 fold2 ans30}; do
    echo "== wide=$wide $wl"
    B200ANS_L3_ENC_WIDE=$wide timeout 600 python tools/kbench.py --workloads $wl --paths 0 --reps 3 2>&1 | tail -1 | cut -c1-300
  done
done
