#!/bin/bash
# round-2 GPU validation, call A (2 GPUs): full GPU test suite (incl. the 2-rank exchange checks), smoke, N=1 and N=2 bench
set -x
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > gpurun_out/a_gpus.txt 2>&1
free -g > gpurun_out/a_host.txt; nproc >> gpurun_out/a_host.txt; grep -m1 "model name" /proc/cpuinfo >> gpurun_out/a_host.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/a_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/a_smoke.log
tail -2 gpurun_out/a_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/a_bench_n1.json 2> gpurun_out/a_bench_n1.err; echo "bench1 rc=$?"
tail -c 600 gpurun_out/a_bench_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/a_bench_n2.json 2> gpurun_out/a_bench_n2.err; echo "bench2 rc=$?"
tail -c 600 gpurun_out/a_bench_n2.json; tail -5 gpurun_out/a_bench_n2.err
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/a_ref.json 2> gpurun_out/a_ref.err; echo "ref rc=$?"
tail -c 1200 gpurun_out/a_ref.json
