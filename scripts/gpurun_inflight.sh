cd $GRAFT_REPO_ROOT
for args in "--inflight 1 --inflight-e2e 2" "--inflight 1 --inflight-e2e 3"; do
python bench.py --steps 6 --warmup 3 --no-cpu-baseline $args 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$args', 'value %.0f'%d['value'], 'e2e %.0f'%d['e2e']['value'], 'frac %.3f'%d['roofline']['frac'], d['breakdown_ms_per_step'], d['ms_per_step'])"
done
