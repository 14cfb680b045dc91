#!/bin/bash
# the driver's scaling launch at N=8 with default flags
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l; nproc
for N in 8; do
  s=$(date +%s)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_default_n$N.json 2> gpurun_out/bench_default_n$N.err; echo "n$N rc=$? wall=$(( $(date +%s) - s ))s"
  tail -2 gpurun_out/bench_default_n$N.err | cut -c1-300
  python - gpurun_out/bench_default_n$N.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], 'n_gpus',d['n_gpus'],'value %.0f'%d['value'],'e2e %.0f'%d['e2e']['value'],'ms/step %.2f'%d['ms_per_step'],'frac %.3f'%d['roofline']['frac'], 'build', d['breakdown_ms_per_step'], 'clocks', d['clocks'])
PY
done
