#!/bin/bash
# round-2 GPU validation (8 GPUs): exchange parity at W = 8, two-device torch op, N = 8 bench with a kernel timeline
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_model_gpu.py -m gpu -x -q -k two_devices 2>&1 | tail -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 tools/check_exchange.py \
    --sharders naive_chunk,greedy,traffic --steps 3 --out gpurun_out/s_exchange_w8.json > gpurun_out/s_exchange_w8.log 2>&1; echo "exchange W=8 rc=$?"
grep -E "check_exchange|parity OK|Error" gpurun_out/s_exchange_w8.log | tail -6 | cut -c1-220
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus 8 --steps 20 --warmup 5 \
    --trace gpurun_out/s_trace_n8.txt > gpurun_out/s_bench_n8.json 2> gpurun_out/s_bench_n8.err; echo "bench N=8 rc=$?"
python - <<P
import json
d = json.load(open("gpurun_out/s_bench_n8.json"))
print("N=8", d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("parity", {}).get("peer_vs_nccl"))
P
