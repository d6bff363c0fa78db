#!/bin/bash
# round-2 GPU validation, call A (2 GPUs): the 2-rank exchange checks and the N=2 bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/a_gpus.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu_gpu.py tests/test_model_gpu.py -m gpu -x -q -k "exchange or two_devices" > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/a_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/a_bench_n2.json 2> gpurun_out/a_bench_n2.err; echo "bench2 rc=$?"
tail -c 1500 gpurun_out/a_bench_n2.json; tail -8 gpurun_out/a_bench_n2.err
