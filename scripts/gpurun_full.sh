#!/bin/bash
# default-size bench (64 submaps x 64 registrations per GPU per step), both arms
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time python bench.py ) > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_full.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ['value','ms_per_step','gpu_launches','breakdown_ms_per_step','converged_frac','mean_iterations','clocks']}); print(d['e2e']); print(d['roofline']); print(d.get('cpu_baseline'))
PY
tail -4 gpurun_out/bench_full.err
( time python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 1500 gpurun_out/bench_ref.json; tail -4 gpurun_out/bench_ref.err
