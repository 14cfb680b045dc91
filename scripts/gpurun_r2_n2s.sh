#!/bin/bash
# Round 2, two GPUs, final binary (term-by-term kernel as the batch default): the multi-device tests, the new step-history test,
# and the default bench at N=2 launched the way the driver launches it (strong scaling: one 4096-registration job cut over the ranks).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
echo "== 1. multi-device tests + step history"
timeout 900 python -m pytest tests/test_shard.py tests/test_gpu_parity.py -m gpu -q -k "shard or multi or step_history or batch_matches" 2>&1 | tail -5 | tee gpurun_out/r02_s_n2_tests.txt
echo "== 2. N=2"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 8 --warmup 3 > gpurun_out/r02_s_n2box_bench_n2.json 2> gpurun_out/r02_s_n2box_bench_n2.err; echo "rc=$?"; tail -c 600 gpurun_out/r02_s_n2box_bench_n2.err
python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r02_s_n2box_bench_n2.json').read().strip().splitlines()[-1])
    print('n2 value', d['value'], 'e2e', d['e2e']['value'], 'ms', d['ms_per_step'], 'scaling', d['scaling'], 'cpu', d.get('cpu_baseline',{}).get('value'), 'parity', d.get('parity_vs_cpu_sample',{}).get('all_within_gate'), 'weak', json.dumps(d.get('weak'))[:200], 'hybrid', d.get('value_hybrid',{}).get('value'))
except Exception as e: print('ERR', e)
PY
