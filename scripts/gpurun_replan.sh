#!/bin/bash
# parity suite, then the default bench at several poll intervals
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -q 2>&1 | tail -6
one() {
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'value %.0f'%d['value'], 'e2e %.0f'%d['e2e']['value'], 'build_ms %.3f'%d['breakdown_ms_per_step']['voxel_build'], 'deriv_ms %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'], 'launches', d['gpu_launches'], 'conv', d['converged_frac'], 'iters', d['mean_iterations'])"
}
for p in 4 2 3 4 2 3 6; do NDTB_POLL=$p one poll$p; done
