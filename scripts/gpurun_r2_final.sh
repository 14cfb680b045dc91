#!/bin/bash
# Round 2, final state: what the driver runs at round end -- the GPU suite, smoke(), both bench arms -- on one fresh box.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== suite"
timeout 1200 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/r02_final_suite.txt
echo "== smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r02_final_smoke.txt
echo "== reference arm"
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_final_bench_reference.json 2> gpurun_out/r02_final_bench_reference.err; echo "rc=$?"
echo "== bench"
timeout 1200 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_final_bench_default.json 2> gpurun_out/r02_final_bench_default.err; echo "rc=$?"; tail -c 300 gpurun_out/r02_final_bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_final_bench_default.json'))
r=json.loads(open('gpurun_out/r02_final_bench_reference.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'], 'ref', r['value'], 'ratio e2e/ref', d['e2e']['value']/r['value'])
print(d['breakdown_ms_per_step']); print(d['parity_vs_cpu_sample']['all_within_gate'], d['roofline']['frac'], d['roofline']['frac_issue'], d['clocks'])
print('e2e_cpp', {k: (v.get('aligns_per_s') if isinstance(v, dict) else None) for k, v in d.get('e2e_cpp', {}).items()})
PY
