#!/bin/bash
# NDTB_FRESH_MERGE=1 (full builds skip k_run_lookup + the prefix scan + the third blocking read): build parity tests with the
# flag on, then the short bench (8 submaps) with the flag on and off on the same box.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
one() {
  NDTB_FRESH_MERGE=$2 timeout 12 python bench.py --steps 5 --warmup 3 --submaps 8 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'value %.0f'%d['value'], 'build_ms %.3f'%d['breakdown_ms_per_step']['voxel_build'], 'step %.3f'%d['ms_per_step'], 'launches', d['gpu_launches'])"
}
{
NDTB_FRESH_MERGE=1 timeout 14 python -m pytest tests/test_gpu_parity.py -q -x -k "voxel_grid or replaced or batch_matches" 2>&1 | tail -2
one fresh_on 1
one fresh_off 0
} 2>&1 | tee gpurun_out/freshmerge.txt
