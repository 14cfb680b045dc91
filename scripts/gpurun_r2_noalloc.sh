#!/bin/bash
# ndtb_fitness_noalloc: its test alone in a fresh process, then after every other GPU test (it is collected last).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== alone"
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -15 | tee gpurun_out/r02_noalloc_alone.txt
echo "== whole GPU suite, -x"
timeout 1200 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -8 | tee gpurun_out/r02_noalloc_suite.txt
