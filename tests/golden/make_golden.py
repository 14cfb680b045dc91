#!/usr/bin/env python
"""Generates tests/golden/ndt_tiny.npz: a small registration problem (inputs included) and the
ORACLE's outputs for it at every stage of the hot path.

The reference ships no golden vectors for this path and cannot be built here, so
these fixtures pin the oracle's behaviour (a regression anchor for oracle and GPU alike), not the
reference's -- parity stays "unpinned" apart from the Gauss-constant known answers.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import oracle  # noqa: E402
from xchu_slam_b200 import synth  # noqa: E402


def main():
    # a dense local crop (+-16 m around the sensor) of a 5k-pt scan / 60k-pt localmap problem keeps the file small
    src, tgt, gt, guess = synth.make_case(n_source=5000, n_target=60000, seed=synth.BASE_SEED + 42)
    world_src = synth.transform(src, gt)
    keep_s = np.max(np.abs(world_src[:, :2] - gt[:2]), axis=1) < 15.0
    keep_t = np.max(np.abs(tgt[:, :2] - gt[:2]), axis=1) < 16.0
    src, tgt = np.ascontiguousarray(src[keep_s]), np.ascontiguousarray(tgt[keep_t])
    out = dict(source=src, target=tgt, gt=gt, guess=guess)
    for name, method in (("d7", oracle.DIRECT7), ("d1", oracle.DIRECT1), ("kd", oracle.KDTREE), ("d26", oracle.DIRECT26)):
        o = oracle.OracleNDT()
        o.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=method, num_threads=1)
        o.set_target(tgt)
        o.set_source(src)
        if name == "d7":
            L = o.dump_leaves()
            for k in ("ijk", "nr_points", "n_raw", "mean", "icov", "centroid"):
                out["vox_" + k] = L[k]
            out["gauss"] = np.array(o.gauss())
        p = oracle.matrix_to_pose(guess)
        xt = o.transform_source(guess)
        cnt, ijk = o.neighbors(xt, method)
        s, g, H, npairs = o.eval_derivatives(guess, p, True)
        out[name + "_nbr_cnt"], out[name + "_nbr_ijk"] = cnt, ijk
        out[name + "_p0"], out[name + "_score0"], out[name + "_g0"], out[name + "_H0"] = p, s, g, H
        out[name + "_npairs0"] = npairs
        r = o.align(guess)
        out[name + "_T"], out[name + "_iterations"], out[name + "_converged"] = r["T"], r["iterations"], int(r["converged"])
        out[name + "_trace_p"] = np.stack([t["p"] for t in r["trace"]])
        out[name + "_trace_score"] = np.array([t["score"] for t in r["trace"]])
        out[name + "_trace_g"] = np.stack([t["g"] for t in r["trace"]])
        out[name + "_trace_step"] = np.array([t["step"] for t in r["trace"]])
        out[name + "_fitness"] = o.fitness()
        if name == "d7":
            out["hess64"] = o.compute_hessian(guess, p)
            out["calc_score"] = o.calculate_score(guess)
    path = os.path.join(HERE, "ndt_tiny.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", {k: (out[k + "_iterations"], float(out[k + "_score0"])) for k in ("d7", "d1", "kd", "d26")})


if __name__ == "__main__":
    main()
