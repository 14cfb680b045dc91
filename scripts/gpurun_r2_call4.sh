#!/bin/bash
# Round 2, GPU call 4: suite with the register-bitonic merge kernel; build A/B + launch list; strict-kernel variants (shared-memory
# accumulators) A/B; default bench line with the C++ pipeline arm; compute-sanitizer memcheck / racecheck.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. suite"
timeout 900 python -m pytest tests -m gpu -q --durations=4 2>&1 | tail -14 | tee gpurun_out/r02_c4_suite.txt
echo "== 2. build paths"
{
timeout 120 python scripts/exp_build.py 20
NDTB_BUILD_RADIX=1 NDTB_SORT_KEY32=1 timeout 120 python scripts/exp_build.py 20
} 2>&1 | tee gpurun_out/r02_c4_build.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_c4_build_counting_launches.csv python scripts/exp_build.py 3 > /dev/null 2>&1
echo "== 3. strict kernel variants (every registration with the term-by-term kernel, 16 submaps)"
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
} 2>&1 | tee gpurun_out/r02_c4_strict_variants.txt
echo "== 4. bench lines"
timeout 900 python bench.py > gpurun_out/r02_c4_bench_default.json 2> gpurun_out/r02_c4_bench_default.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_c4_bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_c4_bench_default.json'))
print('value', d['value'], 'e2e', d['e2e']['value'], 'build', d['breakdown_ms_per_step']['voxel_build'], 'frac_issue', d['roofline'].get('frac_issue'))
print('e2e_cpp', json.dumps(d.get('e2e_cpp'))[:900])
PY
echo "== 5. compute-sanitizer"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_sanitizer_memcheck_smoke.txt 2>&1; echo "memcheck smoke rc=$?"; tail -3 gpurun_out/r02_sanitizer_memcheck_smoke.txt
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "batch_c5_shape or voxel_grid_build or edge_cases_hybrid" > gpurun_out/r02_sanitizer_memcheck_c5.txt 2>&1; echo "memcheck c5 rc=$?"; tail -3 gpurun_out/r02_sanitizer_memcheck_c5.txt
timeout 1200 compute-sanitizer --tool racecheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_sanitizer_racecheck_smoke.txt 2>&1; echo "racecheck smoke rc=$?"; tail -3 gpurun_out/r02_sanitizer_racecheck_smoke.txt
timeout 1500 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "batch_c5_shape or voxel_grid_build" > gpurun_out/r02_sanitizer_racecheck_c5.txt 2>&1; echo "racecheck c5 rc=$?"; tail -3 gpurun_out/r02_sanitizer_racecheck_c5.txt
