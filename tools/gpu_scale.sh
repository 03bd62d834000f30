#!/bin/bash
# N-GPU bench lines (weak scaling): usage gpu_scale.sh N tag
N=${1:-8}; tag=${2:-scale}; out=gpurun_out/$tag; mkdir -p $out
port=29600
for wl in ${WLS:-c2 fold1 c4 c5}; do
  port=$((port+1))
  extra="--no-targets --no-cpu"; [ $wl = c2 ] && extra="--no-targets"
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 10 --warmup 3 --workload $wl $extra > $out/bench_${wl}_n$N.json 2> $out/bench_${wl}_n$N.err; echo "$wl N=$N exit $?"
done
[ -z "$SKIP_REF" ] && timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29650 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $out/bench_ref_n$N.json 2> $out/bench_ref_n$N.err; echo "ref N=$N exit $?"
for f in $out/bench_*_n$N.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], d.get('n_gpus'), round(d['value'],1), d['unit'], 'ms', round(d['ms_per_step'],3), 'e2e', d.get('e2e',{}).get('value'), d.get('roundtrip_ok'))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
done
