#!/bin/bash
# Round 2, GPU call 3: suite; default bench line + reference arm; build A/B with the reworked warp merge; ncu --set full of the
# strict / factorised derivative kernels and of the merge kernel; kernel variants; hybrid criterion at the default K.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. suite"
timeout 900 python -m pytest tests -m gpu -q --durations=6 2>&1 | tail -30 | tee gpurun_out/r02_c3_suite.txt
echo "== 2. bench lines"
timeout 900 python bench.py > gpurun_out/r02_c3_bench_default.json 2> gpurun_out/r02_c3_bench_default.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r02_c3_bench_default.err
timeout 600 python bench.py --impl reference > gpurun_out/r02_c3_bench_reference.json 2> gpurun_out/r02_c3_bench_reference.err; echo "ref rc=$?"; tail -c 300 gpurun_out/r02_c3_bench_reference.json
echo "== 3. build paths"
{
timeout 120 python scripts/exp_build.py 20
NDTB_BUILD_RADIX=1 NDTB_SORT_KEY32=1 timeout 120 python scripts/exp_build.py 20
} 2>&1 | tee gpurun_out/r02_c3_build.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_c3_build_counting_launches.csv python scripts/exp_build.py 3 > /dev/null 2>&1
echo "== 4. ncu --set full"
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-variants"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:^k_derivatives_fast$ -s 6 -c 1 -f -o gpurun_out/r02_deriv_fast_full $B > gpurun_out/ncu_fast.log 2>&1; echo "ncu fast rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:^k_derivatives$ -s 3 -c 1 -f -o gpurun_out/r02_deriv_strict_full $B > gpurun_out/ncu_strict.log 2>&1; echo "ncu strict rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:^k_cs_merge$ -s 2 -c 1 -f -o gpurun_out/r02_cs_merge_full python scripts/exp_build.py 3 > gpurun_out/ncu_merge.log 2>&1; echo "ncu merge rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:^k_cs_finalize$ -s 2 -c 1 -f -o gpurun_out/r02_cs_finalize_full python scripts/exp_build.py 3 > gpurun_out/ncu_fin.log 2>&1; echo "ncu finalize rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_bench_launches.csv $B > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
echo "== 5. hybrid criterion at the default K"
timeout 600 python scripts/exp_hybrid.py 64 2>&1 | grep -v "^   " | tee gpurun_out/r02_c3_hybrid.txt
echo "== 6. kernel variants"
timeout 900 bash scripts/gpurun_variants.sh 2>&1 | tee gpurun_out/r02_c3_variants.txt
