#!/bin/bash
# usage: gpurun_prof_kernel.sh <kernel regex> <skip> <count> <out name> [bench args]
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
K=$1; S=$2; C=$3; O=$4; shift 4
PCMD="python bench.py --steps 1 --warmup 1 --submaps 4 --no-cpu-baseline $@"
ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c $C -o gpurun_out/$O -f $PCMD > gpurun_out/ncu_$O.log 2>&1; echo "ncu rc=$?"
