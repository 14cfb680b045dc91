cd $GRAFT_REPO_ROOT
nproc
for L in 6 8 12 16 24 32; do
NDTB_BUILD_LANES=$L python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-variants --no-cpp-pipeline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('lanes $L value %.0f build %.3f step %.3f'%(d['value'], d['breakdown_ms_per_step']['voxel_build'], d['ms_per_step']))"
done
