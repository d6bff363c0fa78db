#!/bin/bash
# round-2 GPU validation, call B (1 GPU): tcgen05 GEMM bring-up, the single-GPU test suite, smoke, N=1 bench, CPU arm
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
free -g > gpurun_out/b_host.txt; nproc >> gpurun_out/b_host.txt; grep -m1 "model name" /proc/cpuinfo >> gpurun_out/b_host.txt
for k in fwd dgrad wgrad; do
  timeout 120 python tools/gemm_check.py --quick --only $k > gpurun_out/b_gemm_quick_$k.log 2>&1; echo "gemm quick $k rc=$?"
  tail -3 gpurun_out/b_gemm_quick_$k.log
done
timeout 300 python tools/gemm_check.py > gpurun_out/b_gemm_full.log 2>&1; echo "gemm full rc=$?"
tail -30 gpurun_out/b_gemm_full.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/b_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/b_smoke.log 2>&1; echo "smoke rc=$?"
tail -2 gpurun_out/b_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/b_bench_n1.json 2> gpurun_out/b_bench_n1.err; echo "bench1 rc=$?"
tail -c 1500 gpurun_out/b_bench_n1.json; tail -3 gpurun_out/b_bench_n1.err
timeout 500 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/b_ref.json 2> gpurun_out/b_ref.err; echo "ref rc=$?"
tail -c 1500 gpurun_out/b_ref.json; tail -3 gpurun_out/b_ref.err
