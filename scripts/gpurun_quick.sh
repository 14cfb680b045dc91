#!/bin/bash
# parity tests + a short bench line (no profiler)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python bench.py --steps 2 --warmup 3 --submaps 8 --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ['value','ms_per_step','gpu_launches','breakdown_ms_per_step','converged_frac','mean_iterations']}); print(d['e2e']); print(d['roofline'])
PY
tail -3 gpurun_out/bench_quick.err
