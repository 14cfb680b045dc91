"""Experiment: is the table-build stage of a batch bound by the device or by the host's launch rate?  The same 64 targets at
1M, 250k and 62k points each (strided subsamples): a launch-bound stage keeps its time, a device-bound one shrinks with the points.
    python scripts/exp_build_scale.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xchu_slam_b200 import ndt, synth  # noqa: E402

nsub = int(os.environ.get("EXP_SUBMAPS", 64))
w = synth.batch_workload(n_submaps=nsub, scans_per_submap=8, hyps_per_scan=8, cache_dir="/tmp/ndt_b200_bench_cache")
for stride in (1, 4, 16):
    tg = [t[::stride].copy() for t in w["targets"]]
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2, strict_math=0)
    ms = []
    for r in range(5):
        b.clear()
        for t in tg:
            b.add_target(t)
        b.add_source(w["sources"][0])
        b.wait_uploads()
        b.run_raw([0], [0], w["guesses"][:1])
        ms.append(b.stats()["build_ms"])
    print("lanes=%s  %d targets x %d points: build stage ms %s  -> best %.3f (%.1f us per build)" %
          (os.environ.get("NDTB_BUILD_LANES", "default"), nsub, len(tg[0]), ["%.2f" % m for m in ms], min(ms), 1e3 * min(ms) / nsub), flush=True)
