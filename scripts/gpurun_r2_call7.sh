#!/bin/bash
# Round 2, GPU call 7: Morton-ordered sources in large batches: suite, smoke, hybrid study (every item against the strict
# kernel), default bench with and without the ordering.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. suite + smoke"
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6 | tee gpurun_out/r02_c7_suite.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== 2. every item of the job against the strict kernel (hybrid default)"
timeout 600 python scripts/exp_hybrid.py 64 2>&1 | grep -v "^   " | head -12 | tee gpurun_out/r02_c7_hybrid.txt
echo "== 3. bench"
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_c7_bench_default.json 2> gpurun_out/r02_c7_bench_default.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_c7_bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_c7_bench_default.json'))
print('ordered: value', d['value'], 'e2e', d['e2e']['value'], d['breakdown_ms_per_step'], d['parity_vs_cpu_sample']['all_within_gate'], 'factorised', d['value_factorised_only']['value'], 'strict', d['value_strict']['value'])
PY
NDTB_SOURCE_ORDER=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpp-pipeline --no-cpu-baseline --no-variants | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('input order: value', d['value'], d['breakdown_ms_per_step'])"
