#!/bin/bash
# round-2 GPU validation, call C (1 GPU): tcgen05 towers -- kernel tests, GEMM check (graph-timed), MLP/model tests, bench both tower paths
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q > gpurun_out/c_pytest_gemm.log 2>&1; echo "pytest gemm rc=$?"
tail -15 gpurun_out/c_pytest_gemm.log
timeout 600 python tools/gemm_check.py > gpurun_out/c_gemm_full.log 2>&1; echo "gemm full rc=$?"
tail -2 gpurun_out/c_gemm_full.log | cut -c1-300
timeout 900 python -m pytest tests/test_mlp_gpu.py tests/test_model_gpu.py -m gpu -x -q > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/c_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/c_bench_tc.json 2> gpurun_out/c_bench_tc.err; echo "bench tc rc=$?"
tail -c 300 gpurun_out/c_bench_tc.json; tail -3 gpurun_out/c_bench_tc.err
