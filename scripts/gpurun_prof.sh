#!/bin/bash
# ncu launch list + one full capture of the derivative kernel (after a plain run exits 0)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
PCMD="python bench.py --steps 1 --warmup 1 --submaps 4 --no-cpu-baseline"
$PCMD > gpurun_out/plain.json 2> gpurun_out/plain.err || { echo "plain run failed"; tail -5 gpurun_out/plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $PCMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_derivatives -s 3 -c 2 -o gpurun_out/prof_deriv -f $PCMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
