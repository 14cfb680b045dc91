#!/bin/bash
# kernel-variant A/B: the in-tree library and each variants/lib_*.so through the short bench (8 submaps), two passes
cd $GRAFT_REPO_ROOT
one() {  # label, lib, chunk
  NDTB_LIB=${2:+$GRAFT_REPO_ROOT/$2} NDTB_CHUNK=$3 python bench.py --steps 5 --warmup 3 --submaps 8 --no-cpu-baseline --no-variants --strict-math 0 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'chunk=${3:-auto}', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'])"
}
# every variant keeps the arithmetic of the default kernel (only loads / register allocation differ): the parity file must stay green
for lib in $(ls variants/lib_*.so); do
  echo "parity $lib: $(NDTB_LIB=$GRAFT_REPO_ROOT/$lib timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "compute_derivatives or align_c1 or batch_matches" 2>&1 | tail -1)"
done
for pass in 1 2; do
one intree "" ""
for lib in $(ls variants/lib_*.so); do one $lib $lib ""; done
done
