#!/bin/bash
# Round 2, eight GPUs again: whole submaps dealt round-robin over the ranks, gather staged through pinned buffers.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_shard.py -m gpu -q 2>&1 | tail -3
for N in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-variants > gpurun_out/r02_n8box2_bench_n$N.json 2> gpurun_out/r02_n8box2_bench_n$N.err; echo "N=$N rc=$?"; tail -c 300 gpurun_out/r02_n8box2_bench_n$N.err
done
python - <<'PY'
import json
for f in ('n4','n8'):
    try:
        d=json.loads(open('gpurun_out/r02_n8box2_bench_%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value', d['value'], 'e2e', d['e2e']['value'], 'ms', d['ms_per_step'], d['breakdown_ms_per_step'])
    except Exception as e: print(f, 'ERR', e)
PY
