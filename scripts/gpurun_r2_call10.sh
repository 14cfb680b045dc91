#!/bin/bash
# Round-2 GPU call 10: (a) is the batch's table-build stage launch-bound or device-bound (same targets at 1M / 250k / 62k points,
# 12 and 4 lanes); (b) the hybrid replay criterion on two other worlds (4096 registrations each, against the strict kernel).
mkdir -p gpurun_out
python scripts/exp_build_scale.py > gpurun_out/r02_c10_build_scale.txt 2>&1
NDTB_BUILD_LANES=4 python scripts/exp_build_scale.py >> gpurun_out/r02_c10_build_scale.txt 2>&1
python scripts/exp_hybrid_seeds.py 1000 2000 > gpurun_out/r02_c10_hybrid_seeds.txt 2>&1
tail -5 gpurun_out/r02_c10_build_scale.txt gpurun_out/r02_c10_hybrid_seeds.txt
