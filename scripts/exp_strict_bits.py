"""Experiment: are two builds of the term-by-term kernel BIT-identical?  One computeDerivatives (score, gradient, Hessian as the
FP64 sums the kernel forms, in the same order) at several poses of several scan / map cases; run once per library (NDTB_LIB),
then compare the dumps.
    python scripts/exp_strict_bits.py dump out.npz        (with NDTB_LIB set, or the in-tree library)
    python scripts/exp_strict_bits.py compare a.npz b.npz
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if sys.argv[1] == "compare":
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    same = all(np.array_equal(a[k], b[k]) for k in a.files)
    worst = max(float(np.max(np.abs(a[k] - b[k]) / np.maximum(np.abs(a[k]), 1e-300))) for k in a.files)
    print("%s vs %s: %s (%d arrays, %d values, largest relative difference %.3e)" %
          (sys.argv[2], sys.argv[3], "BIT-IDENTICAL" if same else "DIFFERENT", len(a.files), sum(a[k].size for k in a.files), worst))
    sys.exit(0 if same else 1)

from xchu_slam_b200 import ndt, synth  # noqa: E402

out = {}
for c in range(3):
    src, tgt, gt, guess = synth.make_case(n_source=20000, n_target=300000, seed=synth.BASE_SEED + 60 + c)
    rng = np.random.default_rng(c)
    for method in (ndt.DIRECT7, ndt.DIRECT1, ndt.KDTREE):
        g = ndt.NormalDistributionsTransform()
        g.setResolution(1.0)
        g.setNeighborhoodSearchMethod(method)
        g.setStrictMath(True)
        g.setInputTarget(tgt)
        g.setInputSource(src)
        for k in range(3):
            p = np.array(gt, dtype=np.float64) + np.concatenate([rng.uniform(-0.3, 0.3, 3), np.deg2rad(rng.uniform(-1, 1, 3))])
            T = ndt.NormalDistributionsTransform.convertTransform(p)
            s, gr, H = g.evalDerivatives(T, p, True)
            out["c%d_m%d_p%d" % (c, method, k)] = np.concatenate([[s], gr, H.reshape(-1)])
        # and one whole registration
        g.setTransformationEpsilon(0.01)
        g.setMaximumIterations(30)
        g.align(guess)
        r = g.result()
        out["c%d_m%d_align" % (c, method)] = np.concatenate([[r["iterations"], r["score"]], np.asarray(r["T"], dtype=np.float64).reshape(-1)])
np.savez(sys.argv[2], **out)
print("wrote", sys.argv[2], len(out), "arrays; lib =", os.environ.get("NDTB_LIB", "in-tree"))
