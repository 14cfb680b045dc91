#!/bin/bash
# Round 2, GPU call 1: (1) whole GPU suite with the new defaults (counting-sort build, hybrid batch); (2) the same build tests on
# the radix path; (3) fast-vs-strict item study + hybrid threshold sweep; (4) build A/B; (5) kernel variants; (6) raw H2D rate.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. suite (defaults)"
timeout 600 python -m pytest tests -m gpu -q --durations=8 2>&1 | tail -40 | tee gpurun_out/r02_c1_suite.txt
echo "== 2. build / parity file on the radix path"
NDTB_BUILD_RADIX=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q 2>&1 | tail -5 | tee gpurun_out/r02_c1_suite_radix.txt
echo "== 3. hybrid study"
timeout 600 python scripts/exp_hybrid.py 64 2>&1 | tee gpurun_out/r02_c1_hybrid.txt
echo "== 4. build A/B (default bench, no CPU leg)"
one() {  # label, env...
  local label=$1; shift
  env "$@" timeout 120 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$label', 'value %.0f'%d['value'], 'e2e %.0f'%d['e2e']['value'], 'build_ms %.3f'%d['breakdown_ms_per_step']['voxel_build'], 'deriv_ms %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'], 'launches', d['gpu_launches'])"
}
{
for pass in 1 2; do
one counting X=1
one radix64 NDTB_BUILD_RADIX=1
one radix32 NDTB_BUILD_RADIX=1 NDTB_SORT_KEY32=1
one counting_fastonly NDTB_BENCH_STRICT=0
done
} 2>&1 | tee gpurun_out/r02_c1_build_ab.txt
echo "== 5. kernel variants"
timeout 1200 bash scripts/gpurun_variants.sh 2>&1 | tee gpurun_out/r02_c1_variants.txt
echo "== 6. raw pinned H2D / D2H"
timeout 60 python scripts/exp_h2d.py 2>&1 | tee gpurun_out/r02_c1_h2d.txt
