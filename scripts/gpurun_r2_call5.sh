#!/bin/bash
# Round 2, GPU call 5: first-pass abort of the hybrid mode (suite + default bench), strict-kernel code-size variants, the
# single-registration workloads (c1 c3 c4) and the odometry sequence (c2, with and without the align graph).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. suite"
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -6 | tee gpurun_out/r02_c5_suite.txt
echo "== 2. strict kernel variants"
one() {  # label, lib
  NDTB_LIB=${2:+$GRAFT_REPO_ROOT/$2} python bench.py --steps 3 --warmup 2 --submaps 16 --no-cpu-baseline --no-variants --no-cpp-pipeline --strict-math 1 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'])"
}
{
for lib in $(ls variants/lib_*.so); do
  echo "parity $lib: $(NDTB_LIB=$GRAFT_REPO_ROOT/$lib timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -1)"
done
for pass in 1 2; do
one intree ""
for lib in $(ls variants/lib_*.so); do one $lib $lib; done
done
} 2>&1 | tee gpurun_out/r02_c5_strict_variants.txt
echo "== 3. default bench (hybrid, first pass stops at the replay threshold)"
timeout 900 python bench.py --no-cpp-pipeline > gpurun_out/r02_c5_bench_default.json 2> gpurun_out/r02_c5_bench_default.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_c5_bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_c5_bench_default.json'))
print('value', d['value'], 'e2e', d['e2e']['value'], d['breakdown_ms_per_step'], d['parity_vs_cpu_sample']['all_within_gate'])
PY
NDTB_REPLAY_NO_ABORT=1 timeout 600 python bench.py --no-cpp-pipeline --no-cpu-baseline --no-variants | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('no-abort value', d['value'], d['breakdown_ms_per_step'])"
echo "== 4. other workloads"
for wl in c1 c3 c4; do timeout 400 python bench.py --workload $wl > gpurun_out/r02_c5_bench_$wl.json 2>gpurun_out/r02_c5_bench_$wl.err; tail -c 250 gpurun_out/r02_c5_bench_$wl.json; echo; done
timeout 400 python bench.py --workload c2 > gpurun_out/r02_c5_bench_c2.json 2>gpurun_out/r02_c5_bench_c2.err; tail -c 400 gpurun_out/r02_c5_bench_c2.json; echo
NDTB_NO_ALIGN_GRAPH=1 timeout 400 python bench.py --workload c2 > gpurun_out/r02_c5_bench_c2_nograph.json 2>/dev/null; tail -c 300 gpurun_out/r02_c5_bench_c2_nograph.json; echo
