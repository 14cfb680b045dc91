#!/bin/bash
# Round 2, eight GPUs: the default bench the way the driver launches it (strong scaling: one 4096-registration job over 8 ranks).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for N in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_n8box_bench_n$N.json 2> gpurun_out/r02_n8box_bench_n$N.err; echo "N=$N rc=$?"; tail -c 400 gpurun_out/r02_n8box_bench_n$N.err
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpp-pipeline > gpurun_out/r02_n8box_bench_n1.json 2> gpurun_out/r02_n8box_bench_n1.err; echo "N=1 rc=$?"
python - <<'PY'
import json
for f in ('n1','n4','n8'):
    try:
        d=json.loads(open('gpurun_out/r02_n8box_bench_%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value', d['value'], 'e2e', d['e2e']['value'], 'ms', d['ms_per_step'], 'cpu', d.get('cpu_baseline',{}).get('value'), 'parity', d.get('parity_vs_cpu_sample',{}).get('all_within_gate'), 'weak', (d.get('weak') or {}).get('value'), (d.get('weak') or {}).get('e2e'), d['breakdown_ms_per_step'])
    except Exception as e: print(f, 'ERR', e)
PY
