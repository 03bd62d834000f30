#!/bin/bash
tag=${1:-r02aq}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err
echo "bench exit $?"; tail -3 $out/bench.err
timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 > $out/bench_c2.json 2> $out/bench_c2.err
echo "bench c2 exit $?"
python - <<P
import json
for f in ("bench","bench_c2"):
    d=json.loads(open("$out/"+f+".json").read().strip().splitlines()[-1])
    e=d["e2e"]
    print(f, round(d["value"],1), round(d["ms_per_step"],3), d["roundtrip_ok"], "e2e", round(e["value"],2), "pageable", e.get("pageable_gbs"), "copy_only", round(e["copy_only_gbs"],1), "floor", round(e["copy_floor_gbs"],1), "launches", d["gpu_launches"])
P
