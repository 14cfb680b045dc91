cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -5
python bench.py --steps 2 --warmup 1 --submaps 8 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print({k:d[k] for k in ['value','ms_per_step','gpu_launches','breakdown_ms_per_step']}); print(d['e2e']); print(d['roofline'])"
