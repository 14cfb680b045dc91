#!/bin/bash
# ncu launch lists of the two widened rows (loop-closure refinement, ScanContext), each after its plain run exited 0
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python scripts/exp_loop_refine.py 120 16 1 > gpurun_out/loop_small.json 2> gpurun_out/loop_small.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/loop_launches.csv python scripts/exp_loop_refine.py 120 16 1 > gpurun_out/ncu_loop.log 2>&1; echo "loop ncu rc=$?"
python scripts/exp_scancontext.py 16 256 > gpurun_out/sc_small.json 2> gpurun_out/sc_small.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/sc_launches.csv python scripts/exp_scancontext.py 16 256 > gpurun_out/ncu_sc.log 2>&1; echo "sc ncu rc=$?"
