#!/bin/bash
# First GPU call of the next round (everything below was compiled but could not be run when round 1's GPU budget ran out):
#   here, before the call:   python scripts/build_variants.py
#   then:                    gpurun --timeout 2400 -- 'bash scripts/gpurun_r2_first.sh'
# 1. the whole GPU suite and a default bench line with the in-tree library (nothing regressed);
# 2. NDTB_SORT_KEY32=1: the whole suite with the flag on, then bench A/B (scripts/gpurun_key32.sh);
# 3. the kernel variants of scripts/build_variants.py through the 8-submap bench, two passes (scripts/gpurun_variants.sh).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. suite, default library"
timeout 400 bash scripts/gpurun_suite.sh 2>&1 | tee gpurun_out/r2_suite.txt
echo "== 2. 32-bit sort keys"
timeout 500 bash scripts/gpurun_key32.sh
echo "== 3. kernel variants"
timeout 1500 bash scripts/gpurun_variants.sh 2>&1 | tee gpurun_out/r2_variants.txt
