#!/bin/bash
# eight GPUs, final library: the default job (peer-memory exchange) and config 4
out=gpurun_out/${1:-r02au}; mkdir -p $out
show() {
python - <<PY
import json
try:
    d=json.loads(open("$1").read().strip().splitlines()[-1])
    e=d["e2e"]
    print("$2", round(d["value"],1), "GB/s", round(d["ms_per_step"],3), "ms", d["scaling"], "e2e", round(e["value"],1), "copy_only", round(e["copy_only_gbs"],1), "floor", round(e["copy_floor_gbs"],1), "exchange", d.get("allreduce_ms_per_step"), "host", d.get("host_model_ms_per_step"), d.get("roundtrip_ok"), (d.get("parity") or {}).get("ok", (d.get("parity") or {}).get("all_ranks_ok")))
except Exception as ex: print("ERR", ex)
PY
}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 10 --warmup 3 --no-targets > $out/bench_c5_n8.json 2> $out/bench_c5_n8.err; echo "N=8 exit $?"; tail -2 $out/bench_c5_n8.err
show $out/bench_c5_n8.json "N=8 c5"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --workload c4 --steps 10 --warmup 3 --no-cpu > $out/bench_c4_n8.json 2> $out/bench_c4_n8.err; echo "N=8 c4 exit $?"
show $out/bench_c4_n8.json "N=8 c4"
