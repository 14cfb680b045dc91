#!/bin/bash
# GPU parity suite, loop-refinement timing, chunk sweep of the derivative kernel on the in-tree library
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -12
python scripts/exp_loop_refine.py > gpurun_out/loop_refine.json 2> gpurun_out/loop_refine.err; echo "loop rc=$?"; cat gpurun_out/loop_refine.json; tail -3 gpurun_out/loop_refine.err
one() {
  NDTB_CHUNK=$1 python bench.py --steps 5 --warmup 3 --submaps 8 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('chunk=${1:-auto}', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'])"
}
for c in "" 512 768 1024 1280 1536; do one $c; done
