#!/bin/bash
# round-2 GPU validation, call S (8 GPUs): exchange parity at W = 4 and 8 (all sharders, 3 steps, graph replay) and the 2/4/8-GPU bench lines
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/s_gpus.txt 2>&1
for W in 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 2951$W tools/check_exchange.py \
      --sharders naive_chunk,greedy,traffic --steps 3 --out gpurun_out/s_exchange_w$W.json > gpurun_out/s_exchange_w$W.log 2>&1; echo "exchange W=$W rc=$?"
  grep -E "check_exchange|parity OK|Error" gpurun_out/s_exchange_w$W.log | tail -8 | cut -c1-220
done
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 5 \
      > gpurun_out/s_bench_n$N.json 2> gpurun_out/s_bench_n$N.err; echo "bench N=$N rc=$?"
  python - <<P
import json
try:
    d = json.load(open("gpurun_out/s_bench_n$N.json"))
    print("N=$N", d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("parity", {}).get("peer_vs_nccl"))
except Exception as ex:
    print("N=$N no line:", ex)
P
  tail -3 gpurun_out/s_bench_n$N.err | cut -c1-300
done
