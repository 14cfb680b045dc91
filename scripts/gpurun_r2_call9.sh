#!/bin/bash
# Round 2, GPU call 9: strict kernel with its leading accumulators in registers (7 / 13 / 19 of 43), A/B + parity file.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
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
} 2>&1 | tee gpurun_out/r02_c9_strict_variants.txt
