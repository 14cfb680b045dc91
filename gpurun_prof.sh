cd $GRAFT_REPO_ROOT
CMD="python bench.py --steps 1 --warmup 1 --submaps 8 --no-cpu-baseline"
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_derivatives -s 1 -c 2 -o gpurun_out/prof_deriv_v2 -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log | cut -c1-300
