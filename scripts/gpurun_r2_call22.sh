#!/bin/bash
# Round 2, GPU call 22 (packed variants at 4 and 5 blocks per SM): the strict kernel's per-pair float algebra as packed sm_100 instructions (FMUL2 + FFMA2(a, 1, b)):
# bit-identity against the scalar build (same FP64 sums, same order), the parity files, and the short strict-only bench.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
out=gpurun_out/r02_c22_strict_packed.txt
: > $out
python scripts/exp_strict_bits.py dump /tmp/bits_intree.npz >> $out 2>&1
for lib in $(ls variants/lib_strict_packed*.so); do
  NDTB_LIB=$GRAFT_REPO_ROOT/$lib python scripts/exp_strict_bits.py dump /tmp/bits_v.npz >> $out 2>&1
  python scripts/exp_strict_bits.py compare /tmp/bits_intree.npz /tmp/bits_v.npz >> $out 2>&1
  echo "parity $lib: $(NDTB_LIB=$GRAFT_REPO_ROOT/$lib timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "compute_derivatives or align_c1 or batch_matches or batch_c5_shape" 2>&1 | tail -1)" >> $out
done
one() {  # label, lib
  NDTB_LIB=${2:+$GRAFT_REPO_ROOT/$2} python bench.py --steps 4 --warmup 3 --submaps 8 --no-cpu-baseline --no-variants --no-cpp-pipeline --strict-math 1 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'])"
}
for pass in 1 2; do
  one intree "" >> $out
  for lib in $(ls variants/lib_strict_packed*.so); do one $lib $lib >> $out; done
done
cat $out
