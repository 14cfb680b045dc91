#!/bin/bash
# parity suite + default bench line (no CPU sample)
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -q 2>&1 | tail -4
for i in 1 2; do
python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value %.0f'%d['value'], 'e2e %.0f'%d['e2e']['value'], 'build_ms %.3f'%d['breakdown_ms_per_step']['voxel_build'], 'deriv_ms %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'], 'launches', d['gpu_launches'])"
done
python scripts/exp_filter_timing.py 2>/dev/null | cut -c1-400
