#!/bin/bash
# two GPUs: the histogram exchange over NVLink peer memory (CUDA IPC) against ncclAllReduce -- parity test, then the default job
tag=${1:-r02ad}
out=gpurun_out/$tag
mkdir -p $out
timeout 600 python -m pytest tests/test_sharded_gloo.py -m gpu -x -q > $out/pytest.log 2>&1
echo "pytest exit $?"; tail -5 $out/pytest.log
for ipc in 0 1; do
  B200ANS_NO_IPC=$ipc timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2954$ipc bench.py --gpus 2 --steps 10 --warmup 3 --no-targets --no-cpu > $out/bench_c5_n2_noipc$ipc.json 2> $out/bench_c5_n2_noipc$ipc.err
  echo "bench NO_IPC=$ipc exit $?"
  python - <<P
import json
d=json.loads(open("$out/bench_c5_n2_noipc$ipc.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], "allreduce", d.get("allreduce_ms_per_step"), "host", d.get("host_model_ms_per_step"), d["roundtrip_ok"], d["parity"])
P
done
