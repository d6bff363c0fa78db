#!/bin/bash
# round-2 evidence run (1 GPU): full GPU test suite, smoke, the bench line, the reference arm, ncu launch list and one ncu --set full
# capture of a training step's kernels (each ncu pass only after the same command has exited 0 without ncu)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
free -g > gpurun_out/f_host.txt; nproc >> gpurun_out/f_host.txt; grep -m1 "model name" /proc/cpuinfo >> gpurun_out/f_host.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/f_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 --trace gpurun_out/f_trace_n1.txt > gpurun_out/f_bench_n1.json 2> gpurun_out/f_bench_n1.err; echo "bench rc=$?"
timeout 500 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/f_ref.json 2> gpurun_out/f_ref.err; echo "ref rc=$?"
S="--steps 2 --warmup 3 --no-cpu-baseline --no-gpu-reference --no-c3-probe"
timeout 300 python bench.py $S > gpurun_out/f_bench_short.json 2> gpurun_out/f_bench_short.err; echo "short bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/f_launches.csv python bench.py $S > gpurun_out/f_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"embbag_fwd_kernel|bwd_update_kernel|interact_fwd_mma|interact_bwd_mma|gemm_pair_kernel|wgrad_splitk|wgrad_reduce" --launch-skip 40 -c 24 -o gpurun_out/f_ncu_step -f python bench.py $S > gpurun_out/f_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -2 gpurun_out/f_ncu_full.log
