#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
cat > /tmp/f1.py <<'PY'
import sys; sys.path.insert(0,'.')
from xchu_slam_b200 import ndt, synth
w = synth.World(synth.BASE_SEED); poses = w.trajectory(80, seed=synth.BASE_SEED & 0xFFFF)
s = w.scan(poses[50], 500)
g = ndt.NormalDistributionsTransform()
for _ in range(3): print(g.filterScan(s))
PY
python /tmp/f1.py
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/filter_launches.csv python /tmp/f1.py > gpurun_out/ncu_filter.log 2>&1; echo "ncu rc=$?"
python scripts/exp_filter_timing.py
