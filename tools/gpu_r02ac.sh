#!/bin/bash
# DRAM traffic per launch (ncu, dram__bytes_read/write) of the kernels of the default job (C5, 2^32 symbols) and of C4
tag=${1:-r02ac}
out=gpurun_out/$tag
mkdir -p $out
for wl in c5 c4; do
  timeout 600 python bench.py --workload $wl --steps 1 --warmup 1 --no-targets --no-cpu --no-parity > $out/plain_$wl.json 2> $out/plain_$wl.err || { echo "plain $wl failed"; tail -3 $out/plain_$wl.err; continue; }
  timeout 1200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:'histogram_kernel|encode_kernel|compact_kernel|decode_kernel' -c 24 -o $out/traffic_$wl -f python bench.py --workload $wl --steps 1 --warmup 1 --no-targets --no-cpu --no-parity > $out/ncu_$wl.log 2>&1
  echo "ncu $wl exit $?"
done
ls -la $out
