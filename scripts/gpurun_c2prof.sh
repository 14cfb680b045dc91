#!/bin/bash
# C2 (single-handle path): filter timing, the 1000-scan sequence, and the ncu launch list of a 60-scan device-localmap run
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python scripts/exp_filter_timing.py > gpurun_out/filter_timing.json 2> gpurun_out/filter_timing.err; echo "filter rc=$?"; cat gpurun_out/filter_timing.json
python scripts/c2_sequence.py 1000 100 1.0 > gpurun_out/c2_sequence.json 2> gpurun_out/c2_sequence.err; echo "c2 rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c2_sequence.json'))
for k,v in d.items(): print(k, v)
PY
python -c "
import sys; sys.path.insert(0,'.')
from xchu_slam_b200 import synth
synth.write_sequence('/tmp/seq60.bin', n_scans=60, n_points=20000)"
./examples/odom_sequence /tmp/seq60.bin /tmp/o60.txt --res 1.0 --device-localmap > /dev/null 2>&1; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/c2_launches.csv ./examples/odom_sequence /tmp/seq60.bin /tmp/o60n.txt --res 1.0 --device-localmap > gpurun_out/ncu_c2.log 2>&1; echo "ncu rc=$?"
