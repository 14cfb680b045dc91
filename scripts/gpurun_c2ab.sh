#!/bin/bash
# C2 A/B on one box: in-tree library vs variants/oldmm (the per-warp-atomics k_minmax), device-localmap mode, 3 passes each
cd $GRAFT_REPO_ROOT
python -c "
import sys; sys.path.insert(0,'.')
from xchu_slam_b200 import synth
synth.write_sequence('/tmp/seq400.bin', n_scans=400, n_points=20000)"
for pass in 1 2 3; do
  for v in new old; do
    if [ $v = old ]; then export LD_LIBRARY_PATH=$GRAFT_REPO_ROOT/variants/oldmm; else unset LD_LIBRARY_PATH; fi
    ./examples/odom_sequence /tmp/seq400.bin /tmp/o_$v.txt --res 1.0 --device-localmap > /dev/null 2>&1; echo "$v: $(tail -1 /tmp/o_$v.txt | cut -c1-200)"
  done
done
unset LD_LIBRARY_PATH
cut -d" " -f1-10 /tmp/o_new.txt > /tmp/a; cut -d" " -f1-10 /tmp/o_old.txt > /tmp/b; cmp /tmp/a /tmp/b && echo "trajectories identical"
nproc; cat /proc/loadavg
