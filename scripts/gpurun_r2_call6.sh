#!/bin/bash
# Round 2, GPU call 6: suite (8-lane NN query, NN grid prebuilt under the align, fresh-merge on the radix path, 32-bit sort keys,
# build-path equality test), C2 / C1 lines, default bench.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. suite"
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -8 | tee gpurun_out/r02_c6_suite.txt
echo "== 2. C2"
timeout 400 python bench.py --workload c2 > gpurun_out/r02_c6_bench_c2.json 2>gpurun_out/r02_c6_bench_c2.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_c6_bench_c2.json'))
print('c2', d['value'], d['unit'])
for k,v in d['modes'].items(): print('  ', k, {a: v[a] for a in ('scans_per_s','align_ms_mean','target_ms_mean','fitness_ms_mean','surround_ms_mean')})
PY
NDTB_NN_PREBUILD=0 timeout 400 python bench.py --workload c2 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('no-prebuild', d['value']); v=d['modes']['distance_threshold_device_localmap']; print({a: v[a] for a in ('scans_per_s','align_ms_mean','target_ms_mean','fitness_ms_mean')})"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02_c6_c2_launches.csv examples/odom_sequence /tmp/ndt_c2/seq300.bin /tmp/c2_ncu_out.txt --device-localmap --res 1.0 --scans 60 > gpurun_out/ncu_c2.log 2>&1; echo "ncu c2 rc=$?"; ls /tmp/*.bin 2>/dev/null | head
echo "== 3. C1"
timeout 400 python bench.py --workload c1 > gpurun_out/r02_c6_bench_c1.json 2>/dev/null; tail -c 200 gpurun_out/r02_c6_bench_c1.json; echo
echo "== 4. default bench"
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_c6_bench_default.json 2> gpurun_out/r02_c6_bench_default.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_c6_bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_c6_bench_default.json'))
print('value', d['value'], 'e2e', d['e2e']['value'], d['breakdown_ms_per_step'], d['parity_vs_cpu_sample']['all_within_gate'])
print('e2e_cpp', {k: (v.get('aligns_per_s') if isinstance(v, dict) else v) for k, v in d.get('e2e_cpp', {}).items()})
PY
