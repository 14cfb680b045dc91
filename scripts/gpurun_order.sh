#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --submaps 8 --no-cpu-baseline > /dev/null 2>&1
for o in "" random morton:1.0 morton:2.0 morton:0.5; do
  NDT_BENCH_SOURCE_ORDER=$o python bench.py --steps 3 --warmup 3 --submaps 8 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('order=${o:-asis}', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'], 'iters %.3f'%d['mean_iterations'])"
done 2>&1 | tee gpurun_out/order.log
