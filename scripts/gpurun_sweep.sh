#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --durations=12 2>&1 | tail -25
bash scripts/gpurun_variants.sh 2>&1 | tee gpurun_out/variants.log
python scripts/c2_sequence.py 400 40 1.0 > gpurun_out/c2_short.json 2> gpurun_out/c2_short.err; echo "c2 rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c2_short.json'))
for k in ('gpu_device_localmap','gpu_incremental','cpu_reference_all_cores'):
    print(k, d[k])
PY
