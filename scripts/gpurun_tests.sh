#!/bin/bash
# GPU parity tests + production-vs-strict experiment (iteration histogram, sensitivity)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python scripts/exp_fast_vs_strict.py 8 > gpurun_out/exp_fast_vs_strict.log 2>&1; echo "exp rc=$?"
tail -30 gpurun_out/exp_fast_vs_strict.log
