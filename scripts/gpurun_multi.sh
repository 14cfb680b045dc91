#!/bin/bash
# N-GPU bench as the driver launches it (one rank per GPU, NCCL), then the 1-GPU line on the same box
cd $GRAFT_REPO_ROOT
N=${1:-2}
mkdir -p gpurun_out
python bench.py --gpus 1 --steps 3 --warmup 3 --submaps 16 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "n1 rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 --submaps 16 --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "n$N rc=$?"
tail -3 gpurun_out/bench_n$N.err
for f in gpurun_out/bench_n1.json gpurun_out/bench_n$N.json; do python - $f <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], 'n_gpus',d['n_gpus'],'value %.0f'%d['value'],'e2e %.0f'%d['e2e']['value'],'ms/step %.2f'%d['ms_per_step'],'frac %.3f'%d['roofline']['frac'])
PY
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus $N --steps 2 --warmup 1 --submaps 16 --cpu-aligns 64 2>/dev/null | tail -1 | cut -c1-200
