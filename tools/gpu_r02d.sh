#!/bin/bash
# r02d: ANSrfold on the device (tests + C3r line) and the decoder's cooperative 64-byte ring refill (A/B)
tag=${1:-r02d}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_rfold.py -x -q > $out/pytest_rfold.log 2>&1; echo "rfold tests exit $?"; tail -15 $out/pytest_rfold.log
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "rfold or RFOLD or 3-" > $out/pytest_parity_rfold.log 2>&1; echo "parity(rfold) exit $?"; tail -5 $out/pytest_parity_rfold.log
for coop in 0 1; do
  B200ANS_L3_COOP=$coop timeout 600 python tools/kbench.py --workloads fold1,fold2,ans30 --paths 2 --reps 3 > $out/kbench_coop$coop.txt 2>&1; echo "coop=$coop"; cut -c1-400 $out/kbench_coop$coop.txt | tail -3
  B200ANS_L3_COOP=$coop timeout 300 python tools/kbench.py --workloads c5w --paths 2 --reps 3 --bins-scale 4 > $out/kbench_c5m_coop$coop.txt 2>&1; tail -1 $out/kbench_c5m_coop$coop.txt
done
B200ANS_L3_COOP=1 timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "lpb" > $out/pytest_parity_coop.log 2>&1; echo "parity(coop=1) exit $?"; tail -3 $out/pytest_parity_coop.log
timeout 900 python bench.py --workload c3r --no-targets > $out/bench_c3r.json 2> $out/bench_c3r.err; echo "bench c3r exit $?"; tail -3 $out/bench_c3r.err
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/r02d/bench_c3r.json").read().strip().splitlines()[-1]); k=d["kernels"]
    print('c3r value', round(d['value'],1), 'enc', round(d["encode_ms"],3), 'dec', round(d["decode_ms"],3), {n:(v["kernel"], round(v["ms"],3), round(v["frac"],3)) for n,v in k.items()}, d.get('roundtrip_ok'), 'parity', (d.get('parity') or {}).get('ok'))
except Exception as e: print('ERR', e)
PY
