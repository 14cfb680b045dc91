"""Experiment: how far are the factorised kernel's score / gradient / Hessian from the term-by-term kernel's AT IDENTICAL INPUTS
(one computeDerivatives at a perturbed pose, both kernels, several scan / map cases)?  Run with NDTB_LIB pointing at a kernel
variant to compare (e.g. variants/lib_fast_exactcore.so: the scalar core of every pair in the reference's float operations).
    python scripts/exp_fast_terms.py [n_cases]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xchu_slam_b200 import ndt, synth  # noqa: E402

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 6
rs, rg, rh = [], [], []
for c in range(n_cases):
    src, tgt, gt, guess = synth.make_case(n_source=20000, n_target=300000, seed=synth.BASE_SEED + 40 + c)
    rng = np.random.default_rng(c)
    hs = []
    for strict in (True, False):
        g = ndt.NormalDistributionsTransform()
        g.setResolution(1.0)
        g.setNeighborhoodSearchMethod(ndt.DIRECT7)
        g.setStrictMath(strict)
        g.setInputTarget(tgt)
        g.setInputSource(src)
        hs.append(g)
    for k in range(4):
        p = np.array(gt, dtype=np.float64) + np.concatenate([rng.uniform(-0.3, 0.3, 3), np.deg2rad(rng.uniform(-1, 1, 3))])
        T = ndt.NormalDistributionsTransform.convertTransform(p)
        (s1, g1, H1), (s0, g0, H0) = hs[0].evalDerivatives(T, p, True), hs[1].evalDerivatives(T, p, True)
        rs.append(abs(s0 - s1) / abs(s1))
        rg.append(np.max(np.abs(g0 - g1)) / np.max(np.abs(g1)))
        rh.append(np.max(np.abs(H0 - H1)) / np.max(np.abs(H1)))
        # the Newton step both would take
        d1, d0 = np.linalg.solve(H1, -g1), np.linalg.solve(H0, -g0)
        print("case %d pose %d: score %.2e  grad %.2e  hess %.2e  |dp| %.3e  rel diff of dp %.2e" %
              (c, k, rs[-1], rg[-1], rh[-1], np.linalg.norm(d1), np.linalg.norm(d0 - d1) / np.linalg.norm(d1)), flush=True)
print("lib=%s  median / max relative difference: score %.2e / %.2e   gradient %.2e / %.2e   Hessian %.2e / %.2e" %
      (os.environ.get("NDTB_LIB", "in-tree"), np.median(rs), max(rs), np.median(rg), max(rg), np.median(rh), max(rh)))
