#!/bin/bash
# Round 2, GPU call 17: the single-registration lines (C1 / C3 / C4) and the odometry sequence (C2) with the final binary (the
# strict kernel -- what a handle runs -- changed its block shape), then the whole GPU suite once more (new step-history test).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for w in c1 c3 c4 c2; do
  timeout 900 python bench.py --workload $w > gpurun_out/r02_s_bench_$w.json 2> gpurun_out/r02_s_bench_$w.err; echo "$w rc=$?"; tail -c 200 gpurun_out/r02_s_bench_$w.err
done
python - <<'PY'
import json
for w in ('c1','c3','c4','c2'):
    try:
        d=json.loads(open('gpurun_out/r02_s_bench_%s.json'%w).read().strip().splitlines()[-1])
        print(w, d.get('value'), d.get('unit'), json.dumps({k:v for k,v in d.items() if k in ('build_ms','align_ms','e2e')})[:300])
    except Exception as e: print(w,'ERR',e)
PY
echo "== suite"
timeout 1500 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -4 | tee gpurun_out/r02_s2_suite.txt
