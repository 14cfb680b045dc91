#!/bin/bash
# kernel-variant sweep: each variants/lib_*.so through the short bench
cd $GRAFT_REPO_ROOT
python bench.py --steps 1 --warmup 1 --submaps 8 --no-cpu-baseline > /dev/null 2>&1
for lib in "" $(ls variants/lib_*.so); do
  NDTB_LIB=${lib:+$GRAFT_REPO_ROOT/$lib} python bench.py --steps 3 --warmup 3 --submaps 8 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('${lib:-default}', 'value %.0f'%d['value'], 'deriv_ms/step %.3f'%d['breakdown_ms_per_step']['derivatives'], 'step %.3f'%d['ms_per_step'], 'frac %.3f'%d['roofline']['frac'])"
done
