"""Experiment: production vs strict derivative kernel over a batch of registrations (C5 shape)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xchu_slam_b200 import ndt, synth
nsub = int(sys.argv[1]) if len(sys.argv) > 1 else 2
method = int(sys.argv[2]) if len(sys.argv) > 2 else 2
w = synth.batch_workload(n_submaps=nsub, scans_per_submap=8, hyps_per_scan=8, cache_dir="/tmp/ndt_b200_bench_cache")
res = {}
for strict in (0, 1):
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=method, strict_math=strict)
    for t in w["targets"]: b.add_target(t)
    for s in w["sources"]: b.add_source(s)
    t0 = time.time(); r = b.run(w["item_target"], w["item_source"], w["guesses"]); dt = time.time() - t0
    res[strict] = r
    print("strict", strict, "time", dt, "stats", {k: v for k, v in b.stats().items() if k in ("deriv_ms", "total_ms")})
a, s = res[0], res[1]
it_eq = np.array([x["iterations"] == y["iterations"] for x, y in zip(a, s)])
dt = np.array([np.max(np.abs(x["T"][:3, 3] - y["T"][:3, 3])) for x, y in zip(a, s)])
dr = np.array([np.max(np.abs(x["T"][:3, :3] - y["T"][:3, :3])) for x, y in zip(a, s)])
print("items", len(a), "iterations equal", it_eq.mean(), "max dT all", dt.max(), "max dR all", dr.max())
print("with equal iterations: max dT", dt[it_eq].max(), "max dR", dr[it_eq].max())
print("iteration histogram", np.bincount([x["iterations"] for x in s]))
bad = np.flatnonzero(~it_eq)
for i in bad[:10]:
    print(i, a[i]["iterations"], s[i]["iterations"], dt[i], dr[i])
print("dT percentiles", np.percentile(dt, [50, 90, 99, 100]))
its = np.array([x["iterations"] for x in s])
for i in np.flatnonzero(dt > 1e-6):
    print("item", i, "iters", its[i], "dT", dt[i], "dR", dr[i], "score fast/strict", a[i]["score"], s[i]["score"])
