#!/bin/bash
# round-2 GPU validation, call S (8 GPUs): exchange parity at W = 8 (all sharders, 3 steps, graph replay), the 2/4/8-GPU bench lines
# (weak scaling, as the driver runs them) with a kernel timeline at N = 8, and one strong-scaling line (global batch 32768) at N = 8
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/s_gpus.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 tools/check_exchange.py \
    --sharders naive_chunk,greedy,traffic --steps 3 --out gpurun_out/s_exchange_w8.json > gpurun_out/s_exchange_w8.log 2>&1; echo "exchange W=8 rc=$?"
grep -E "check_exchange|parity OK|Error" gpurun_out/s_exchange_w8.log | tail -4 | cut -c1-220
for N in 2 4 8; do
  T=""; if [ $N = 8 ]; then T="--trace gpurun_out/s_trace_n8.txt"; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 5 $T \
      > gpurun_out/s_bench_n$N.json 2> gpurun_out/s_bench_n$N.err; echo "bench N=$N rc=$?"
  python - <<P
import json
try:
    d = json.load(open("gpurun_out/s_bench_n$N.json"))
    print("N=$N", d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("parity", {}).get("peer_vs_nccl"))
except Exception as ex:
    print("N=$N no line:", ex)
P
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29538 bench.py --gpus 8 --steps 20 --warmup 5 --global-batch 32768 \
    > gpurun_out/s_bench_n8_strong.json 2> gpurun_out/s_bench_n8_strong.err; echo "bench N=8 strong rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/s_bench_n8_strong.json'));print('N=8 strong',d['value'],d['ms_per_step'],d['scaling'])"
