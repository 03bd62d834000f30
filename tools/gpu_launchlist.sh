#!/bin/bash
# launch list (ncu gpu__time_duration, cold and serialised) of a command: gpu_launchlist.sh tag name skip count -- cmd...
tag=$1; name=$2; skip=$3; count=$4; shift 5
out=gpurun_out/$tag; mkdir -p $out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s $skip -c $count --csv --log-file $out/${name}_launches.csv "$@" > $out/ncu_$name.log 2>&1; echo "ncu exit $?"
python - <<PY
import csv
rows=[r for r in csv.reader(open("$out/${name}_launches.csv")) if len(r)>5]
h=rows[0]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
for r in rows[1:]:
    print(r[ki][:64], r[vi])
PY
