#!/bin/bash
# r02j: ANSrfold fast route: tests, probe over value ranges / scattered hot values, C3r line
tag=${1:-r02j}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_rfold.py -x -q > $out/pytest_rfold.log 2>&1; echo "rfold tests exit $?"; tail -12 $out/pytest_rfold.log
timeout 600 python tools/rfold_hist_probe.py > $out/probe.txt 2>&1; tail -6 $out/probe.txt
timeout 900 python bench.py --workload c3r --no-targets > $out/bench_c3r.json 2> $out/bench_c3r.err; echo "bench c3r exit $?"; tail -3 $out/bench_c3r.err
python - <<PY
import json
try:
    d=json.loads(open("$out/bench_c3r.json").read().strip().splitlines()[-1]); k=d["kernels"]
    print('c3r value', round(d['value'],1), 'enc', round(d["encode_ms"],3), 'dec', round(d["decode_ms"],3), {n:(v["kernel"], round(v["ms"],3), round(v["frac"],3)) for n,v in k.items()}, d.get('roundtrip_ok'), 'parity', (d.get('parity') or {}).get('ok'), 'e2e', d['e2e']['value'])
except Exception as e: print('ERR', e)
PY
