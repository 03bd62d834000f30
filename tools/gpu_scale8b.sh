#!/bin/bash
# eight GPUs: the default job (config 5: ONE 2^32-symbol job) with the peer-memory exchange and with NCCL, then config 4
# (pseudo-adaptive, 2^27 symbols per GPU = 2^30 over 8 GPUs, no collective)
out=gpurun_out/${1:-r02ae}; mkdir -p $out
show() {
python - <<PY
import json
try:
    d=json.loads(open("$1").read().strip().splitlines()[-1])
    print("$2", round(d["value"],1), "GB/s", round(d["ms_per_step"],3), "ms", d["scaling"], "e2e", round(d["e2e"]["value"],1), "allreduce", d.get("allreduce_ms_per_step"), "host", d.get("host_model_ms_per_step"), {k:(round(v["ms"],3),round(v["frac"],3)) for k,v in d["kernels"].items()}, d.get("roundtrip_ok"), (d.get("parity") or {}).get("ok", (d.get("parity") or {}).get("all_ranks_ok")))
except Exception as e: print("ERR", e)
PY
}
B200ANS_NO_IPC=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 10 --warmup 3 --no-targets > $out/bench_c5_n8.json 2> $out/bench_c5_n8.err; echo "N=8 peer memory exit $?"; tail -2 $out/bench_c5_n8.err
show $out/bench_c5_n8.json "N=8 peer-memory"
B200ANS_NO_IPC=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 10 --warmup 3 --no-targets --no-cpu > $out/bench_c5_n8_nccl.json 2> $out/bench_c5_n8_nccl.err; echo "N=8 nccl exit $?"
show $out/bench_c5_n8_nccl.json "N=8 nccl"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --workload c4 --steps 10 --warmup 3 > $out/bench_c4_n8.json 2> $out/bench_c4_n8.err; echo "N=8 c4 exit $?"
show $out/bench_c4_n8.json "N=8 c4"
