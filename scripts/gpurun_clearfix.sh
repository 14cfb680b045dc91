#!/bin/bash
# Stale dense index after a target replacement: the regression test against the library before the fix (must FAIL) and the
# fixed one (must pass), then the whole GPU suite and one default bench line.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
T=tests/test_gpu_parity.py::test_target_replaced_within_the_same_bounding_cells
echo "== old library (variants/lib_oldclear.so)"
NDTB_LIB=$GRAFT_REPO_ROOT/variants/lib_oldclear.so timeout 60 python -m pytest $T -q -x 2>&1 | tail -6
echo "== whole suite, fixed library"
timeout 130 python -m pytest tests -m gpu -q 2>&1 | tail -4
timeout 40 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/bench_clearfix.json
python -c "
import json
d=json.load(open('gpurun_out/bench_clearfix.json')); print('value %.0f'%d['value'], 'e2e %.0f'%d['e2e']['value'], 'build_ms %.3f'%d['breakdown_ms_per_step']['voxel_build'], 'step %.3f'%d['ms_per_step'], 'launches', d['gpu_launches'])"
