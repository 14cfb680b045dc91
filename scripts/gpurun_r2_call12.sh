#!/bin/bash
# Round-2 GPU call 12: (a) factorised vs term-by-term kernel at identical inputs, default and exact-core variant; (b) the
# exact-core variant over three worlds (offenders of the factorised kernel alone, first-pass time); (c) strict-kernel variants
# at three 128-thread blocks per SM (most accumulators in registers) through the short strict-only bench.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python scripts/exp_fast_terms.py 4 > gpurun_out/r02_c12_terms_default.txt 2>&1
NDTB_LIB=$GRAFT_REPO_ROOT/variants/lib_fast_exactcore.so python scripts/exp_fast_terms.py 4 > gpurun_out/r02_c12_terms_exactcore.txt 2>&1
one() {  # label, lib
  NDTB_LIB=${2:+$GRAFT_REPO_ROOT/$2} python bench.py --steps 4 --warmup 3 --submaps 8 --no-cpu-baseline --no-variants --strict-math 1 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'])"
}
for lib in $(ls variants/lib_strict_t*.so); do
  echo "parity $lib: $(NDTB_LIB=$GRAFT_REPO_ROOT/$lib timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "compute_derivatives or align_c1 or batch_matches" 2>&1 | tail -1)" >> gpurun_out/r02_c12_strict_variants.txt
done
for pass in 1 2; do
  one intree "" >> gpurun_out/r02_c12_strict_variants.txt
  for lib in $(ls variants/lib_strict_t*.so); do one $lib $lib >> gpurun_out/r02_c12_strict_variants.txt; done
done
NDTB_LIB=$GRAFT_REPO_ROOT/variants/lib_fast_exactcore.so EXP_TAG=_exactcore python scripts/exp_hybrid_seeds.py 0 1000 5000 > gpurun_out/r02_c12_hybrid_seeds_exactcore.txt 2>&1
tail -n 3 gpurun_out/r02_c12_terms_default.txt gpurun_out/r02_c12_terms_exactcore.txt
cat gpurun_out/r02_c12_strict_variants.txt
cut -c1-300 gpurun_out/r02_c12_hybrid_seeds_exactcore.txt
