"""Committed fixtures (tests/golden/ndt_tiny.npz, made by tests/golden/make_golden.py).

CPU: the oracle still reproduces them (any change of the oracle's behaviour is visible in review).
GPU: the CUDA path through the C ABI reproduces them -- neighbour sets / voxel membership bit-exact,
score / gradient / Hessian within 1e-6, iteration counts equal, final transform 1e-4 m / 1e-5 rad.
The fixtures hold the ORACLE's outputs; the reference has none for this path."""
import os

import numpy as np
import pytest

from conftest import rel_err

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ndt_tiny.npz")
METHODS = {"d7": 2, "d1": 3, "kd": 0, "d26": 1}


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def test_gauss_constants_fixture(gold):
    assert tuple(gold["gauss"]) == (-2.217225244042889, 0.43312300470355464, 0.5978370007556204)


@pytest.mark.parametrize("name", list(METHODS))
def test_oracle_reproduces_golden(gold, name):
    from oracle import oracle
    o = oracle.OracleNDT()
    o.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=METHODS[name], num_threads=2)
    o.set_target(gold["target"])
    o.set_source(gold["source"])
    if name == "d7":
        L = o.dump_leaves()
        for k in ("ijk", "nr_points", "n_raw", "mean", "centroid"):
            assert np.array_equal(L[k], gold["vox_" + k]), k
        assert np.allclose(L["icov"], gold["vox_icov"], rtol=1e-12, atol=0)
    guess = gold["guess"]
    cnt, ijk = o.neighbors(o.transform_source(guess), METHODS[name])
    assert np.array_equal(cnt, gold[name + "_nbr_cnt"]) and np.array_equal(ijk, gold[name + "_nbr_ijk"])
    s, g, H, npairs = o.eval_derivatives(guess, gold[name + "_p0"], True)
    assert npairs == gold[name + "_npairs0"]
    assert abs(s - gold[name + "_score0"]) <= 1e-12 * abs(s)      # thread count only changes the f64 summation order
    assert rel_err(g, gold[name + "_g0"]) < 1e-12 and rel_err(H, gold[name + "_H0"]) < 1e-12
    r = o.align(guess)
    assert r["iterations"] == gold[name + "_iterations"] and int(r["converged"]) == gold[name + "_converged"]
    assert np.max(np.abs(r["T"] - gold[name + "_T"])) < 1e-6
    assert abs(o.fitness() - gold[name + "_fitness"]) <= 1e-9 * gold[name + "_fitness"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(METHODS))
def test_gpu_reproduces_golden(gold, name):
    from xchu_slam_b200 import ndt
    g = ndt.NormalDistributionsTransform()
    g.setResolution(1.0)
    g.setStepSize(0.1)
    g.setTransformationEpsilon(0.01)
    g.setMaximumIterations(30)
    g.setNeighborhoodSearchMethod(METHODS[name])
    g.setInputTarget(gold["target"])
    g.setInputSource(gold["source"])
    guess = gold["guess"]
    if name == "d7":
        L = g.dumpVoxels()
        so = np.lexsort((gold["vox_ijk"][:, 0], gold["vox_ijk"][:, 1], gold["vox_ijk"][:, 2]))
        sg = np.lexsort((L["ijk"][:, 0], L["ijk"][:, 1], L["ijk"][:, 2]))
        for k in ("ijk", "nr_points", "n_raw", "mean", "centroid"):
            assert np.array_equal(L[k][sg], gold["vox_" + k][so]), k
        assert rel_err(g.computeHessian(guess, gold["d7_p0"]), gold["hess64"]) < 1e-6
        assert abs(g.calculateScore(guess) - gold["calc_score"]) <= 1e-6 * abs(gold["calc_score"])
    # neighbour sets of the source moved by the guess: pcl::transformPointCloud in float, as the oracle does it
    T = guess.astype(np.float32)
    s3 = gold["source"][:, :3].astype(np.float32)
    xt = np.stack([((T[r, 0] * s3[:, 0] + T[r, 1] * s3[:, 1]) + T[r, 2] * s3[:, 2]) + T[r, 3] for r in range(3)], axis=1)
    cnt, ijk = g.neighbors(xt, METHODS[name])
    assert np.array_equal(cnt, gold[name + "_nbr_cnt"]) and np.array_equal(ijk, gold[name + "_nbr_ijk"])
    s, gr, H = g.evalDerivatives(guess, gold[name + "_p0"], True)
    assert abs(s - gold[name + "_score0"]) <= 1e-6 * abs(gold[name + "_score0"])
    assert rel_err(gr, gold[name + "_g0"]) < 1e-6 and rel_err(H, gold[name + "_H0"]) < 1e-6
    g.align(guess)
    r = g.result()
    assert r["iterations"] == gold[name + "_iterations"] and int(r["converged"]) == gold[name + "_converged"]
    assert np.max(np.abs(r["T"][:3, 3] - gold[name + "_T"][:3, 3])) < 1e-4
    assert np.max(np.abs(r["T"][:3, :3] - gold[name + "_T"][:3, :3])) < 1e-5
    assert abs(g.getFitnessScore() - gold[name + "_fitness"]) <= 1e-6 * gold[name + "_fitness"]
