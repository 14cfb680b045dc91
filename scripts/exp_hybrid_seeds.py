"""Experiment: does the hybrid mode's replay criterion (tuned on the default C5 job) hold on other jobs?  For each world seed:
the whole 4096-registration job with the strict kernel, with the factorised kernel alone (no abort, no replay) and in the default
hybrid mode; counts the registrations of each outside the north_star gate (iteration count / convergence flag equal to the strict
kernel's, final transform within 1e-4 m / 1e-5 rad) and keeps the per-item loop diagnostics and |delta_p| histories of the strict
and the factorised run (what a replay criterion can be studied on offline: scripts/study_replay_criterion.py).
Writes gpurun_out/r02_hybrid_seeds.json and gpurun_out/r02_hybrid_seed_<offset>.npz.
    python scripts/exp_hybrid_seeds.py [seed_offset ...]
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xchu_slam_b200 import ndt, synth  # noqa: E402

offsets = [int(a) for a in sys.argv[1:]] or [0, 1000, 2000, 3000, 4000, 5000]
nsub = int(os.environ.get("EXP_SUBMAPS", 64))
tag = os.environ.get("EXP_TAG", "")       # output-file suffix (kernel-variant runs: NDTB_LIB=variants/lib_x.so EXP_TAG=_x)


def run(w, strict):
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2, strict_math=strict)
    b.clear()
    for t in w["targets"]:
        b.add_target(t)
    for s in w["sources"]:
        b.add_source(s)
    r = b.run_raw(w["item_target"], w["item_source"], w["guesses"]).copy()
    n = len(w["item_target"])
    return r, b.stats(), b.diag(n), b.step_history(n)


def outside(a, s):
    Ta, Ts = a["T"].reshape(-1, 4, 4), s["T"].reshape(-1, 4, 4)
    dt = np.max(np.abs(Ta[:, 3, :3] - Ts[:, 3, :3]), axis=1)
    dr = np.max(np.abs(Ta[:, :3, :3] - Ts[:, :3, :3]), axis=(1, 2))
    off = (a["iterations"] != s["iterations"]) | (a["converged"] != s["converged"]) | (dt >= 1e-4) | (dr >= 1e-5)
    return off, dt, dr


rows = []
os.makedirs("gpurun_out", exist_ok=True)
for o in offsets:
    w = synth.batch_workload(n_submaps=nsub, scans_per_submap=8, hyps_per_scan=8, seed=synth.BASE_SEED + o)
    n = len(w["item_target"])
    rs, ss, ds, hs = run(w, 1)
    rf, sf, df, hf = run(w, 0)
    rh, sh, dh, _ = run(w, -1)
    of, dtf, drf = outside(rf, rs)
    oh, dth, drh = outside(rh, rs)
    row = dict(seed_offset=o, items=n, factorised_outside=int(of.sum()), hybrid_outside=int(oh.sum()), replayed=int(sh["replayed"]),
               hybrid_max_dt=float(dth.max()), hybrid_max_dr=float(drh.max()), hybrid_total_ms=sh["total_ms"], strict_total_ms=ss["total_ms"],
               factorised_total_ms=sf["total_ms"], factorised_deriv_ms=sf["deriv_ms"], strict_deriv_ms=ss["deriv_ms"],
               hybrid_deriv_ms=sh["deriv_ms"], hybrid_replay_deriv_ms=sh["replay_deriv_ms"], mean_iterations=float(rs["iterations"].mean()),
               hybrid_offenders=[dict(item=int(i), it_hybrid=int(rh["iterations"][i]), it_strict=int(rs["iterations"][i]), dt=float(dth[i]), dr=float(drh[i]))
                                 for i in np.flatnonzero(oh)])
    rows.append(row)
    print(json.dumps(row), flush=True)
    json.dump(rows, open("gpurun_out/r02_hybrid_seeds%s.json" % tag, "w"), indent=1)
    np.savez_compressed("gpurun_out/r02_hybrid_seed_%d%s.npz" % (o, tag), it_fast=rf["iterations"], it_strict=rs["iterations"], conv_fast=rf["converged"],
                        conv_strict=rs["converged"], dt=dtf.astype(np.float32), dr=drf.astype(np.float32), dt_hybrid=dth.astype(np.float32),
                        dr_hybrid=drh.astype(np.float32), replayed=dh["replayed"], borderline=df["borderline"], n_rise=df["n_rise"],
                        last_norm=df["last_norm"], norms_fast=hf, norms_strict=hs, score_fast=rf["score"], score_strict=rs["score"])
