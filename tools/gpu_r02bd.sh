#!/bin/bash
out=gpurun_out/${1:-r02bd}; mkdir -p $out
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err
echo "bench exit $?"
python - <<P
import json
d=json.loads(open("$out/bench.json").read().strip().splitlines()[-1])
e=d["e2e"]
print("bench", round(d["value"],1), round(d["ms_per_step"],3), d["roundtrip_ok"], d["parity"]["ok"], "e2e", round(e["value"],2), "pageable", round(e["pageable_gbs"],2), "frac", round(d["roofline"]["frac"],3), "launches", d["gpu_launches"])
P
