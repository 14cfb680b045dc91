#!/bin/bash
# Round 2, GPU call 16: the latency chain of a point in the strict kernel -- next source record prefetched, DIRECT7 cell codes in
# registers behind a bit mask (no local-memory list), the dense index box in registers -- through the parity files and the short
# strict-only bench.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
out=gpurun_out/r02_c16_strict_variants.txt
: > $out
one() {  # label, lib, chunk
  NDTB_LIB=${2:+$GRAFT_REPO_ROOT/$2} NDTB_CHUNK=$3 python bench.py --steps 4 --warmup 3 --submaps 8 --no-cpu-baseline --no-variants --no-cpp-pipeline --strict-math 1 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'chunk=${3:-auto}', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'])"
}
for lib in $(ls variants/lib_strict_*.so); do
  echo "parity $lib: $(NDTB_LIB=$GRAFT_REPO_ROOT/$lib timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "compute_derivatives or align_c1 or batch_matches or batch_c5_shape" 2>&1 | tail -1)" >> $out
done
for pass in 1 2; do
  one intree "" "" >> $out
  for lib in $(ls variants/lib_strict_*.so); do one $lib $lib "" >> $out; done
done
cat $out
