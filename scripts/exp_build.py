"""Experiment: device time of ONE full voxel-table build (setInputTarget of a 1M-point submap) on a single handle, and of 64 builds
through a batch's build lanes.  Select the path with NDTB_BUILD_RADIX / NDTB_SORT_KEY32 in the environment.
    python scripts/exp_build.py [reps]
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xchu_slam_b200 import ndt, synth  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
nsub = int(os.environ.get("EXP_SUBMAPS", 8))
w = synth.batch_workload(n_submaps=nsub, scans_per_submap=8, hyps_per_scan=8, cache_dir="/tmp/ndt_b200_bench_cache")
g = ndt.NormalDistributionsTransform()
g.setResolution(1.0)
ms = []
for r in range(reps):
    g.setInputTarget(w["targets"][r % nsub])
    ms.append(g.lastTiming()["build_ms"])
print("path radix=%s key32=%s  single-handle build (upload + table) ms: first %.3f  median %.3f  min %.3f" %
      (os.environ.get("NDTB_BUILD_RADIX", "0"), os.environ.get("NDTB_SORT_KEY32", "0"), ms[0], float(np.median(ms[2:])), min(ms)))
if reps > 4:
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2, strict_math=0)
    best = 1e9
    for r in range(4):
        b.clear()
        for t in w["targets"]:
            b.add_target(t)
        b.add_source(w["sources"][0])
        b.wait_uploads()
        b.run_raw([0], [0], w["guesses"][:1])
        best = min(best, b.stats()["build_ms"])
    print("  batch: %d builds on the lanes: %.3f ms (%.1f us per build)" % (nsub, best, 1e3 * best / nsub))
