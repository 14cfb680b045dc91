#!/bin/bash
# The driver's round-end sequence at N=1: reference arm then our arm, both with default flags; wall-clock of each.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
s=$(date +%s); python bench.py --impl reference > gpurun_out/bench_ref_default.json 2> gpurun_out/bench_ref_default.err; echo "ref rc=$? wall=$(( $(date +%s) - s ))s"
s=$(date +%s); python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$? wall=$(( $(date +%s) - s ))s"
tail -c 1500 gpurun_out/bench_ref_default.json; tail -c 3500 gpurun_out/bench_default.json; tail -5 gpurun_out/bench_default.err
