#!/bin/bash
# Round 2, GPU call 14: the batch default is the term-by-term kernel (three 128-thread blocks per SM).  What the driver runs at
# round end -- GPU suite, smoke(), both bench arms -- then the ncu capture of one full launch of the strict kernel and the launch
# list of one default step.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== suite"
timeout 1500 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/r02_s_suite.txt
echo "== smoke"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 | tee gpurun_out/r02_s_smoke.txt
echo "== reference arm"
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_s_bench_reference.json 2> gpurun_out/r02_s_bench_reference.err; echo "rc=$?"
echo "== bench"
timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_s_bench_default.json 2> gpurun_out/r02_s_bench_default.err; echo "rc=$?"; tail -c 300 gpurun_out/r02_s_bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_s_bench_default.json'))
r=json.loads(open('gpurun_out/r02_s_bench_reference.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'], 'ref', r['value'], 'ratio e2e/ref', d['e2e']['value']/r['value'])
print(d['breakdown_ms_per_step']); print(d['parity_vs_cpu_sample']['all_within_gate'], d['roofline']['frac'], d['roofline']['frac_issue'], d['clocks'])
for k in ('value_hybrid','value_factorised_only'):
    v=d.get(k,{}); print(k, v.get('value'), (v.get('parity_vs_cpu_sample') or {}).get('also_final_transform_within_1e-4m_1e-5rad'))
print('e2e_cpp', {k: (v.get('aligns_per_s') if isinstance(v, dict) else None) for k, v in d.get('e2e_cpp', {}).items()})
PY
echo "== ncu"
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-variants --no-cpp-pipeline"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:^k_derivatives$ -s 6 -c 1 -f -o gpurun_out/r02_s_deriv_strict_full $B > gpurun_out/ncu_strict.log 2>&1; echo "ncu strict rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_s_bench_launches.csv $B > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
ls -la gpurun_out | tail -12
