#!/bin/bash
# Round 2, GPU call 2: suite with the align graph / window / multi entry points; hybrid criterion check over 4096 items; per-kernel
# timing of the counting-sort build vs the radix build; kernel variants; the new bench lines.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. suite"
timeout 900 python -m pytest tests -m gpu -q --durations=6 2>&1 | tail -30 | tee gpurun_out/r02_c2_suite.txt
echo "== 2. hybrid criterion over the full job"
timeout 600 python scripts/exp_hybrid.py 64 2>&1 | grep -v "^   " | tee gpurun_out/r02_c2_hybrid.txt
echo "== 3. build paths"
{
timeout 120 python scripts/exp_build.py 20
NDTB_BUILD_RADIX=1 NDTB_SORT_KEY32=1 timeout 120 python scripts/exp_build.py 20
NDTB_BUILD_RADIX=1 timeout 120 python scripts/exp_build.py 20
} 2>&1 | tee gpurun_out/r02_c2_build.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_build_counting_launches.csv python scripts/exp_build.py 3 > /dev/null 2>&1
NDTB_BUILD_RADIX=1 NDTB_SORT_KEY32=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_build_radix32_launches.csv python scripts/exp_build.py 3 > /dev/null 2>&1
echo "== 4. bench lines"
timeout 600 python bench.py > gpurun_out/r02_c2_bench_default.json 2> gpurun_out/r02_c2_bench_default.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r02_c2_bench_default.err
timeout 300 python bench.py --workload c2 > gpurun_out/r02_c2_bench_c2.json 2>&1; tail -c 300 gpurun_out/r02_c2_bench_c2.json
NDTB_NO_ALIGN_GRAPH=1 timeout 300 python bench.py --workload c2 > gpurun_out/r02_c2_bench_c2_nograph.json 2>&1
for wl in c1 c3 c4; do timeout 300 python bench.py --workload $wl > gpurun_out/r02_c2_bench_$wl.json 2>&1; tail -c 200 gpurun_out/r02_c2_bench_$wl.json; done
echo "== 5. kernel variants"
timeout 900 bash scripts/gpurun_variants.sh 2>&1 | tee gpurun_out/r02_c2_variants.txt
