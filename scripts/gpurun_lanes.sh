#!/bin/bash
# build-lane sweep on the default 64-submap batch, then loop-refinement timing with the stage breakdown
cd $GRAFT_REPO_ROOT
one() {
  NDTB_BUILD_LANES=$1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('lanes=$1', 'value %.0f'%d['value'], 'e2e %.0f'%d['e2e']['value'], 'build_ms %.3f'%d['breakdown_ms_per_step']['voxel_build'], 'deriv_ms %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'])"
}
for l in 4 6 8 12 16; do one $l; done
python scripts/exp_loop_refine.py > gpurun_out/loop_refine.json 2> gpurun_out/loop_refine.err; echo "loop rc=$?"; cat gpurun_out/loop_refine.json
