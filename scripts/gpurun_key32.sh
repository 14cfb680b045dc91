#!/bin/bash
# NDTB_SORT_KEY32=1 (cell keys of <= 32 bits generated / sorted / compared as 32-bit words): the whole GPU suite with the flag on,
# then the default bench with the flag on and off on the same box.  Untested on a GPU when committed (round 1 ended with the
# GPU budget spent): flip the default only after this script is green.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
one() {
  NDTB_SORT_KEY32=$2 timeout 60 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'value %.0f'%d['value'], 'build_ms %.3f'%d['breakdown_ms_per_step']['voxel_build'], 'step %.3f'%d['ms_per_step'], 'launches', d['gpu_launches'])"
}
{
NDTB_SORT_KEY32=1 timeout 300 python -m pytest tests -q -x -m gpu 2>&1 | tail -3
one key32_on 1
one key32_off 0
one key32_on 1
one key32_off 0
} 2>&1 | tee gpurun_out/key32.txt
