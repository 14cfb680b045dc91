#!/bin/bash
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -15
python scripts/c2_sequence.py 1000 100 1.0 > gpurun_out/c2_sequence.json 2> gpurun_out/c2_sequence.err; echo "c2 rc=$?"
tail -c 5000 gpurun_out/c2_sequence.json; tail -5 gpurun_out/c2_sequence.err
