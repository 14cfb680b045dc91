#!/bin/bash
# Round-2 GPU call 11: the hybrid replay criterion on six worlds (the default job + five others), per-item diagnostics kept.
mkdir -p gpurun_out
python scripts/exp_hybrid_seeds.py 0 1000 2000 3000 4000 5000 > gpurun_out/r02_c11_hybrid_seeds.txt 2>&1
tail -n 8 gpurun_out/r02_c11_hybrid_seeds.txt | cut -c1-400
