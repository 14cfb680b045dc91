#!/bin/bash
# Round 2, two GPUs: the multi-device tests, then the default bench at N=1 and N=2 on the same box (strong scaling: one 4096-
# registration job cut over the ranks), and the reference arm launched the way the driver launches it.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
echo "== 1. multi-device tests"
timeout 900 python -m pytest tests/test_shard.py -m gpu -q 2>&1 | tail -5 | tee gpurun_out/r02_n2_tests.txt
echo "== 2. N=1"
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_n2box_bench_n1.json 2> gpurun_out/r02_n2box_bench_n1.err; echo "rc=$?"; tail -c 300 gpurun_out/r02_n2box_bench_n1.err
echo "== 3. N=2"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_n2box_bench_n2.json 2> gpurun_out/r02_n2box_bench_n2.err; echo "rc=$?"; tail -c 600 gpurun_out/r02_n2box_bench_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 5 --warmup 3 --impl reference > gpurun_out/r02_n2box_bench_ref_n2.json 2> gpurun_out/r02_n2box_bench_ref_n2.err; echo "rc=$?"; tail -c 300 gpurun_out/r02_n2box_bench_ref_n2.err
python - <<'PY'
import json
for f in ('n1','n2'):
    try:
        d=json.loads(open('gpurun_out/r02_n2box_bench_%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value', d['value'], 'e2e', d['e2e']['value'], 'ms', d['ms_per_step'], 'scaling', d['scaling'], 'cpu', d.get('cpu_baseline',{}).get('value'), 'parity', d.get('parity_vs_cpu_sample',{}).get('all_within_gate'), 'weak', json.dumps(d.get('weak'))[:200])
    except Exception as e: print(f, 'ERR', e)
PY
tail -c 400 gpurun_out/r02_n2box_bench_ref_n2.json
