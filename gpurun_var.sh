cd $GRAFT_REPO_ROOT
python bench.py --steps 1 --warmup 1 --submaps 8 --no-cpu-baseline > /dev/null 2>&1
for v in a_256_1_reg_pf b_128_3_smem c_128_2_reg_pf d_128_3_reg e_128_4_smem f_64_6_reg; do
 for c in 256 512 2048; do
  NDTB_CHUNK=$c NDTB_LIB=$GRAFT_REPO_ROOT/variants/lib_$v.so python bench.py --steps 2 --warmup 1 --submaps 8 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v chunk $c', 'deriv_ms/step %.2f'%d['breakdown_ms_per_step']['derivatives'], 'step %.2f'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'])"
 done
done
