#!/bin/bash
# the driver's multi-GPU launch at N=2 with default flags (both arms)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref2 rc=$?"; tail -c 300 gpurun_out/bench_ref_n2.json
s=$(date +%s)
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_default_n2.json 2> gpurun_out/bench_default_n2.err; echo "n2 rc=$? wall=$(( $(date +%s) - s ))s"
tail -2 gpurun_out/bench_default_n2.err | cut -c1-300
python - gpurun_out/bench_default_n2.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('n_gpus',d['n_gpus'],'value %.0f'%d['value'],'e2e %.0f'%d['e2e']['value'],'ms/step %.2f'%d['ms_per_step'],'frac %.3f'%d['roofline']['frac'], 'build', d['breakdown_ms_per_step'], 'clocks', d['clocks'])
PY
