#!/bin/bash
# Round 2, GPU call 8: the documented knobs' fallback paths (radix build, input-order sources, no align graph, no NN prebuild)
# through the parity files; ncu capture of the factorised kernel on Morton-ordered sources; launch list of the final default step.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== 1. fallback paths"
{
echo "radix build:   $(NDTB_BUILD_RADIX=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_localmap.py -m gpu -q 2>&1 | tail -1)"
echo "64-bit keys:   $(NDTB_BUILD_RADIX=1 NDTB_SORT_KEY32=0 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k 'voxel or incremental or neighbor or replaced' 2>&1 | tail -1)"
echo "input order:   $(NDTB_SOURCE_ORDER=0 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k 'batch' 2>&1 | tail -1)"
echo "no graph:      $(NDTB_NO_ALIGN_GRAPH=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_cpp_shim.py -m gpu -q 2>&1 | tail -1)"
echo "no prebuild:   $(NDTB_NN_PREBUILD=0 timeout 600 python -m pytest tests/test_cpp_shim.py tests/test_gpu_parity.py -m gpu -q -k 'fitness or trajectory or odometry' 2>&1 | tail -1)"
echo "no abort:      $(NDTB_REPLAY_NO_ABORT=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k 'batch' 2>&1 | tail -1)"
} 2>&1 | tee gpurun_out/r02_c8_fallback_paths.txt
echo "== 2. ncu"
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-variants --no-cpp-pipeline"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:^k_derivatives_fast$ -s 6 -c 1 -f -o gpurun_out/r02_final_deriv_fast_full $B > gpurun_out/ncu_fast.log 2>&1; echo "ncu fast rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:^k_derivatives$ -s 3 -c 1 -f -o gpurun_out/r02_final_deriv_strict_full $B > gpurun_out/ncu_strict.log 2>&1; echo "ncu strict rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_final_bench_launches.csv $B > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
