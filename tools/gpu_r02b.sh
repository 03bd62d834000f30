#!/bin/bash
# r02b: full GPU test suite + the bench default (C5, fixed 2^32-symbol job) + C3 / C3r / C2 lines.
tag=${1:-r02b}; out=gpurun_out/$tag; mkdir -p $out
t0=$(date +%s)
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $? ($(( $(date +%s) - t0 )) s)"; tail -3 $out/pytest_gpu.log
[ "${SKIP_BENCH:-0}" = 1 ] && exit 0
t0=$(date +%s)
timeout 900 python bench.py > $out/bench_c5.json 2> $out/bench_c5.err; echo "bench default exit $? ($(( $(date +%s) - t0 )) s)"; tail -3 $out/bench_c5.err
for wl in ${WLS:-c3 c3r c2}; do
  t0=$(date +%s)
  timeout 900 python bench.py --workload $wl --no-targets > $out/bench_$wl.json 2> $out/bench_$wl.err; echo "bench $wl exit $? ($(( $(date +%s) - t0 )) s)"; tail -2 $out/bench_$wl.err
done
for f in $out/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); k=d["kernels"]
    print(sys.argv[1].split('/')[-1], d["config"]["block_syms"], 'value', round(d['value'],1), 'enc', round(d["encode_ms"],3), 'dec', round(d["decode_ms"],3), 'bps', round(d["bits_per_symbol"],4), {n:(v["kernel"], round(v["ms"],3), round(v["frac"],3)) for n,v in k.items()}, 'e2e', round(d['e2e']['value'],2), 'copy_only', round(d['e2e']['copy_only_gbs'],2), d.get('roundtrip_ok'), 'parity', (d.get('parity') or {}).get('ok'))
    for n,v in (d.get("north_star_2p30") or {}).items():
        print('   ', n, {kk:(round(vv["ms"],3), round(vv["frac"],3)) for kk,vv in v.get("kernels",{}).items()}, v.get("parity",{}).get("ok"), v.get("error"))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
done
