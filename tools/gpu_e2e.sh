#!/bin/bash
tag=${1:-e2e}; out=gpurun_out/$tag; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $out/pytest_gpu.log
timeout 600 python bench.py --no-targets > $out/bench_c2.json 2> $out/bench_c2.err; echo "bench c2 exit $?"
B200ANS_NO_PIPELINE=1 timeout 600 python bench.py --no-targets --no-cpu > $out/bench_c2_nopipe.json 2> $out/bench_c2_nopipe.err; echo "bench c2 nopipe exit $?"
timeout 600 python bench.py --workload c4 --no-targets --no-cpu > $out/bench_c4.json 2> $out/bench_c4.err; echo "bench c4 exit $?"
OUT=$out python - <<'PY'
import json,sys
for f in ("bench_c2","bench_c2_nopipe","bench_c4"):
    try:
        d=json.load(open(f"%s/%s.json" % (__import__("os").environ.get("OUT","gpurun_out/e2e"), f)))
    except Exception as e:
        print(f, "ERR", e); continue
    print(f, "value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "enc", round(d["encode_ms"],3), "dec", round(d["decode_ms"],3), "e2e", d["e2e"])
PY
