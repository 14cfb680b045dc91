#!/bin/bash
# ndtb_batch_fitness over the build lanes: the tests that exercise it, then the loop-refinement stage timing.
mkdir -p gpurun_out
timeout 90 python -m pytest tests/test_loop.py tests/test_loop_pipeline.py tests/test_gpu_parity.py -m gpu -q -x -k "loop or refine or batch or fitness or pipeline" 2>&1 | tail -4
timeout 60 python scripts/exp_loop_refine.py 160 32 2 > gpurun_out/loop_refine_fitlanes.json 2>gpurun_out/loop_refine_fitlanes.err
python -c "
import json
d=json.load(open('gpurun_out/loop_refine_fitlanes.json'))
for m in ('reference','current'):
    print(m, 'ms %.2f'%d[m]['gpu_ms_per_call'], {k:round(v,2) for k,v in d[m]['stage_ms'].items()}, d[m]['accepted'], d[m]['sample_agrees_with_oracle'])
"
