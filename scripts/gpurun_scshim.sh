#!/bin/bash
# SCManager drop-in header: caller-level parity of the two builds of examples/sc_sequence.cpp, then both timed on one sequence.
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_cpp_shim.py -x -q -m gpu -k sc_same > gpurun_out/scshim_test.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/scshim_test.txt
python -c "
from xchu_slam_b200 import synth
synth.write_sequence('/tmp/s.bin', n_scans=150, n_points=20000)"
./examples/sc_sequence /tmp/s.bin gpurun_out/sc_dev.txt --revisit 3,10,20,40,60
./examples/sc_sequence_ref /tmp/s.bin gpurun_out/sc_ref.txt --revisit 3,10,20,40,60 > /dev/null
grep -v '^#' gpurun_out/sc_dev.txt > /tmp/a; grep -v '^#' gpurun_out/sc_ref.txt > /tmp/b
cmp /tmp/a /tmp/b && echo IDENTICAL >> gpurun_out/scshim_test.txt
tail -q -n 1 gpurun_out/sc_dev.txt gpurun_out/sc_ref.txt >> gpurun_out/scshim_test.txt
cat gpurun_out/scshim_test.txt
