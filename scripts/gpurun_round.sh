#!/bin/bash
# One GPU-box visit: parity tests, a short bench, the ncu launch list and one full capture of the derivative kernel.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/gpu.txt; nproc >> gpurun_out/gpu.txt; lscpu | grep "Model name" >> gpurun_out/gpu.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
CMD="python bench.py --steps 2 --warmup 3 --submaps 8"
$CMD > gpurun_out/bench_s8.json 2> gpurun_out/bench_s8.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_s8.json
PCMD="python bench.py --steps 1 --warmup 1 --submaps 4 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $PCMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_derivatives -s 3 -c 2 -o gpurun_out/prof_deriv -f $PCMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
