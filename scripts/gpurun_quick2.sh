#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -8
python scripts/exp_filter_timing.py > gpurun_out/filter_timing.json; cat gpurun_out/filter_timing.json
python scripts/c2_sequence.py 400 40 1.0 > gpurun_out/c2_short.json 2>/dev/null
python - <<'PY'
import json
d=json.load(open('gpurun_out/c2_short.json'))
for k in ('gpu_device_localmap','gpu_incremental'): print(k, d[k])
PY
python bench.py --steps 3 --warmup 3 --submaps 8 --no-cpu-baseline | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value %.0f'%d['value'], 'e2e %.0f'%d['e2e']['value'], d['breakdown_ms_per_step'], 'frac %.3f'%d['roofline']['frac'])"
