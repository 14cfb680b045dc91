"""Experiment: factorised vs term-by-term derivative kernel over the C5 batch, item by item, and what the hybrid mode's replay
criterion costs at several thresholds.  Writes gpurun_out/r02_hybrid_items.npz (per-item iterations of both kernels, final-
transform differences, loop diagnostics) and prints a summary.
    python scripts/exp_hybrid.py [n_submaps]
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xchu_slam_b200 import ndt, synth  # noqa: E402

nsub = int(sys.argv[1]) if len(sys.argv) > 1 else 64
w = synth.batch_workload(n_submaps=nsub, scans_per_submap=8, hyps_per_scan=8, cache_dir="/tmp/ndt_b200_bench_cache")
n = len(w["item_target"])


def run(strict, env=None, reps=3):
    for k, v in (env or {}).items():
        os.environ[k] = str(v)
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2, strict_math=strict)
    best = None
    for _ in range(reps):
        b.clear()
        for t in w["targets"]:
            b.add_target(t)
        for s in w["sources"]:
            b.add_source(s)
        r = b.run_raw(w["item_target"], w["item_source"], w["guesses"])
        st = b.stats()
        if best is None or st["total_ms"] < best[1]["total_ms"]:
            best = (r.copy(), st, b.diag(n))
    for k in (env or {}):
        os.environ.pop(k, None)
    return best


def diff(a, s):
    Ta, Ts = a["T"].reshape(-1, 4, 4), s["T"].reshape(-1, 4, 4)      # column-major: translation in row 3
    dt = np.max(np.abs(Ta[:, 3, :3] - Ts[:, 3, :3]), axis=1)
    dr = np.max(np.abs(Ta[:, :3, :3] - Ts[:, :3, :3]), axis=(1, 2))
    return dt, dr


rs, ss, _ = run(1)
rf, sf, df = run(0)
dt, dr = diff(rf, rs)
off = (rf["iterations"] != rs["iterations"]) | (rf["converged"] != rs["converged"]) | (dt >= 1e-4) | (dr >= 1e-5)
print("items", n, "strict total_ms %.2f deriv_ms %.2f | fast total_ms %.2f deriv_ms %.2f" % (ss["total_ms"], ss["deriv_ms"], sf["total_ms"], sf["deriv_ms"]))
print("fast vs strict: outside the gate", int(off.sum()), "iteration mismatch", int((rf["iterations"] != rs["iterations"]).sum()))
print("iteration histogram (strict)", np.bincount(rs["iterations"]).tolist())
print("offenders: (item, it_fast, it_strict, dT, dR, borderline, n_rise, last_norm)")
for i in np.flatnonzero(off):
    print("  ", i, rf["iterations"][i], rs["iterations"][i], "%.2e" % dt[i], "%.2e" % dr[i], df["borderline"][i], df["n_rise"][i], "%.3e" % df["last_norm"][i])
print("min fast iterations among offenders:", int(rf["iterations"][off].min()) if off.any() else None)
print("borderline flagged:", int(df["borderline"].sum()), "of which offenders", int((df["borderline"].astype(bool) & off).sum()))
for r_ in (1, 2, 3, 4):
    m = df["n_rise"] >= r_
    print("n_rise >= %d: %d items, covers %d of %d offenders" % (r_, int(m.sum()), int((m & off).sum()), int(off.sum())))
os.makedirs("gpurun_out", exist_ok=True)
np.savez("gpurun_out/r02_hybrid_items.npz", it_fast=rf["iterations"], it_strict=rs["iterations"], conv_fast=rf["converged"], conv_strict=rs["converged"],
         dt=dt, dr=dr, borderline=df["borderline"], n_rise=df["n_rise"], last_norm=df["last_norm"], score_fast=rf["score"], score_strict=rs["score"])
rows = []
for K in (8, 10, 12, 14, 16, 20):
    rh, sh, dh = run(-1, {"NDTB_REPLAY_ITERS": K})
    dth, drh = diff(rh, rs)
    offh = (rh["iterations"] != rs["iterations"]) | (rh["converged"] != rs["converged"]) | (dth >= 1e-4) | (drh >= 1e-5)
    rows.append(dict(K=K, total_ms=sh["total_ms"], deriv_ms=sh["deriv_ms"], replay_ms=sh["replay_ms"], replay_deriv_ms=sh["replay_deriv_ms"],
                     replayed=int(sh["replayed"]), outside_gate=int(offh.sum()), aligns_per_s=n / (sh["total_ms"] / 1e3)))
    print(json.dumps(rows[-1]))
json.dump(dict(items=n, strict=ss, fast=sf, fast_outside_gate=int(off.sum()), hybrid=rows), open("gpurun_out/r02_hybrid_sweep.json", "w"), indent=1)
