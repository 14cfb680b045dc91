#!/bin/bash
# Round-2 GPU call 13: the batch default is now the term-by-term kernel.  (a) parity files of the in-tree library (new default block
# shape) and of every strict-kernel variant; (b) the variants through the short strict-only bench, two passes; (c) the batch tests
# whose modes changed.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
out=gpurun_out/r02_c13_strict_variants.txt
echo "parity intree: $(timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -1)" > $out
one() {  # label, lib
  NDTB_LIB=${2:+$GRAFT_REPO_ROOT/$2} python bench.py --steps 4 --warmup 3 --submaps 8 --no-cpu-baseline --no-variants --no-cpp-pipeline --strict-math 1 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'])"
}
for lib in $(ls variants/lib_strict_*.so); do
  echo "parity $lib: $(NDTB_LIB=$GRAFT_REPO_ROOT/$lib timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "compute_derivatives or align_c1 or batch_matches or batch_c5_shape" 2>&1 | tail -1)" >> $out
done
for pass in 1 2; do
  one intree "" >> $out
  for lib in $(ls variants/lib_strict_*.so); do one $lib $lib >> $out; done
done
cat $out
