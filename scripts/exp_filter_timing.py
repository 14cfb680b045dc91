"""Timing of the device filter_node front end (ndtb_filter_scan) on raw 64-beam scans, beside the oracle's restatement."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle
from xchu_slam_b200 import ndt, synth
w = synth.World(synth.BASE_SEED)
poses = w.trajectory(80, seed=synth.BASE_SEED & 0xFFFF)
scans = [w.scan(poses[50 + k], 500 + k) for k in range(8)]
g = ndt.NormalDistributionsTransform()
out = {}
for leaf in (0.5, 0.2):
    for s in scans[:2]:
        g.filterScan(s, leaf=leaf)
    t = []
    for rep in range(5):
        for s in scans:
            t0 = time.perf_counter(); c = g.filterScan(s, leaf=leaf); t.append((time.perf_counter() - t0) * 1e3)
    t0 = time.perf_counter(); ro = oracle.filter_scan(scans[-1], leaf=leaf); tc = (time.perf_counter() - t0) * 1e3
    assert c == ro["counts"] and np.array_equal(g.getSource(), ro["cloud"])
    out["leaf_%.1f" % leaf] = {"raw_points": int(scans[-1].shape[0]), "counts_gate_downsample_sor": list(c), "gpu_ms_median": float(np.median(t)),
                               "gpu_ms_p90": float(np.percentile(t, 90)),
                               "cpu_oracle_ms": tc, "cpu_note": "oracle restatement, exhaustive k-NN over all host cores: a checker, not a tuned kd-tree baseline"}
print(json.dumps(out))
