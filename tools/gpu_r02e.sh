#!/bin/bash
# r02e: ANSrfold with 32-bit raw bins: tests, C3r line, launch list
tag=${1:-r02e}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_rfold.py -x -q > $out/pytest_rfold.log 2>&1; echo "rfold tests exit $?"; tail -3 $out/pytest_rfold.log
timeout 900 python bench.py --workload c3r --no-targets > $out/bench_c3r.json 2> $out/bench_c3r.err; echo "bench c3r exit $?"; tail -3 $out/bench_c3r.err
python - <<'PY'
import json
try:
    d=json.loads(open("gpurun_out/r02e/bench_c3r.json").read().strip().splitlines()[-1]); k=d["kernels"]
    print('c3r value', round(d['value'],1), 'enc', round(d["encode_ms"],3), 'dec', round(d["decode_ms"],3), {n:(v["kernel"], round(v["ms"],3), round(v["frac"],3)) for n,v in k.items()}, d.get('roundtrip_ok'), 'parity', (d.get('parity') or {}).get('ok'), 'e2e', d['e2e']['value'])
except Exception as e: print('ERR', e)
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/c3r_launches.csv python bench.py --workload c3r --no-targets --steps 1 --warmup 1 --no-cpu --no-parity > $out/ncu_c3r.log 2>&1; echo "ncu exit $?"
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open("gpurun_out/r02e/c3r_launches.csv")) if len(r)>5]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
for r in rows[1:][-45:]:
    print(r[ki][:70], r[vi], r[ui])
PY
