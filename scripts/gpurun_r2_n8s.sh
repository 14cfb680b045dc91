#!/bin/bash
# Round 2, eight GPUs, final binary (term-by-term kernel as the batch default): the default bench at N=8 launched the way the driver
# launches it (strong scaling: one 4096-registration job cut over the ranks; the weak figure beside it).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 6 --warmup 3 --no-cpp-pipeline --cpu-aligns 256 > gpurun_out/r02_s_n8box_bench_n8.json 2> gpurun_out/r02_s_n8box_bench_n8.err; echo "rc=$?"; tail -c 400 gpurun_out/r02_s_n8box_bench_n8.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r02_s_n8box_bench_n8.json').read().strip().splitlines()[-1])
    print('n8 value', d['value'], 'e2e', d['e2e']['value'], 'ms', d['ms_per_step'], 'scaling', d['scaling'], 'parity', d.get('parity_vs_cpu_sample',{}).get('all_within_gate'), 'weak', json.dumps(d.get('weak'))[:200], 'hybrid', d.get('value_hybrid',{}).get('value'))
    print(d['breakdown_ms_per_step'])
except Exception as e: print('ERR', e)
PY
