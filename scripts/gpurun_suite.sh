#!/bin/bash
# whole GPU parity suite, then one default bench line (no CPU sample)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 130 python -m pytest tests -m gpu -q --durations=8 2>&1 | tail -16
timeout 40 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/bench_suite.json
python -c "
import json
d=json.load(open('gpurun_out/bench_suite.json')); print('value %.0f'%d['value'], 'e2e %.0f'%d['e2e']['value'], 'step %.3f'%d['ms_per_step'], 'launches', d['gpu_launches'])"
