#!/bin/bash
# Round-end sequence at N=1: parity suite, reference arm, our arm (default flags), 8-submap line, ncu launch list + full capture,
# filter timing, C2 sequence
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -4
s=$(date +%s); python bench.py --impl reference > gpurun_out/bench_ref_default.json 2> gpurun_out/bench_ref_default.err; echo "ref rc=$? wall=$(( $(date +%s) - s ))s"
s=$(date +%s); python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$? wall=$(( $(date +%s) - s ))s"
python bench.py --steps 5 --warmup 3 --submaps 8 > gpurun_out/bench_s8.json 2> gpurun_out/bench_s8.err; echo "s8 rc=$?"
tail -c 1200 gpurun_out/bench_ref_default.json; tail -c 3800 gpurun_out/bench_default.json; tail -3 gpurun_out/bench_default.err
PCMD="python bench.py --steps 1 --warmup 1 --submaps 4 --no-cpu-baseline"
$PCMD > gpurun_out/plain.json 2> gpurun_out/plain.err || { echo "plain run failed"; tail -5 gpurun_out/plain.err; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv $PCMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_derivatives -s 3 -c 2 -o gpurun_out/prof_deriv -f $PCMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
python scripts/exp_filter_timing.py > gpurun_out/filter_timing.json 2> gpurun_out/filter_timing.err; echo "filter rc=$?"; cat gpurun_out/filter_timing.json
python scripts/c2_sequence.py 1000 100 1.0 > gpurun_out/c2_sequence.json 2> gpurun_out/c2_sequence.err; echo "c2 rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/c2_sequence.json'))
for k,v in d.items(): print(k, v)
PY
