"""Pins the oracle (oracle/ndt_oracle.cpp) against the REFERENCE'S OWN NDT SOURCES.

oracle/_ref/libndt_ref.so is /root/reference/xchu_mapping/include/pclomp/{ndt_omp.h, ndt_omp_impl.hpp,
voxel_grid_covariance_omp.h, voxel_grid_covariance_omp_impl.hpp} compiled unmodified (oracle/Makefile
target `ref`) against stand-in Eigen/PCL headers (oracle/ref_stubs/) whose arithmetic is the shared
restatement oracle/thirdparty_math.h.  Both sides therefore use the same linear-algebra primitives and
differ in who wrote the NDT code -- so with one thread the comparison is BIT-EXACT, for every stage the
path has: constructor defaults, Gauss constants, voxel grid (keys, counts, mean, covariance, eigenvalues,
inverse covariance, f32 centroid), neighbour sets of all four search methods, computeDerivatives,
computeHessian, calculateScore, the Newton / More-Thuente loop (iteration count, final transform,
probability, output cloud) and getFitnessScore.

Built where /root/reference exists (this container); the prebuilt library travels to the GPU box.
"""
import numpy as np
import pytest

from oracle import oracle, ref

pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libndt_ref.so absent and /root/reference not present to build it")

METHODS = {"KDTREE": 0, "DIRECT26": 1, "DIRECT7": 2, "DIRECT1": 3}


def _pair(case, res=1.0, step=0.1, eps=0.01, max_iter=30, method=2, threads=1):
    src, tgt = case[0], case[1]
    o, r = oracle.OracleNDT(), ref.RefNDT()
    for m in (o, r):
        m.set_params(resolution=res, step_size=step, trans_eps=eps, max_iter=max_iter, search_method=method, num_threads=threads)
        m.set_target(tgt)
        m.set_source(src)
    return o, r


def test_ctor_defaults_are_the_references():
    d = ref.RefNDT().ctor_defaults()      # NDT_I:46-69, VGC_H:209-212
    assert d == dict(resolution=1.0, step_size=0.1, outlier_ratio=0.55, trans_eps=0.1, max_iter=35, search_method=2, min_points=6, eig_mult=0.01)
    o = oracle.OracleNDT()
    assert (o.resolution, o.step_size, o.outlier_ratio, o.trans_eps, o.max_iter, o.search_method) == (1.0, 0.1, 0.55, 0.1, 35, 2)


@pytest.mark.parametrize("res", [0.5, 1.0, 2.0])
def test_gauss_constants(case_tiny, res):
    o, r = _pair(case_tiny, res=res)
    assert o.gauss() == r.gauss()
    if res == 1.0:      # the known answer embedded in the reference's lib/libndt_cpu.so
        assert r.gauss()[:2] == (-2.217225244042889, 0.43312300470355464)


@pytest.mark.parametrize("res", [0.5, 1.0, 2.0])
def test_voxel_grid_bit_exact(case_small, res):
    o, r = _pair(case_small, res=res)
    lo, lr = o.dump_leaves(), r.dump_leaves()
    assert len(lr["ijk"]) > 1000
    for k in ("ijk", "nr_points", "mean", "cov", "icov", "centroid", "evals"):
        assert np.array_equal(lo[k], lr[k], equal_nan=True), k
    mn, mx = r.grid_bounds()
    assert np.array_equal(lo["ijk"].min(0) >= mn, [True] * 3) and np.array_equal(lo["ijk"].max(0) <= mx, [True] * 3)
    # voxel_centroids_ (the radius-search cloud) holds exactly the leaves with >= min_points points, flagged ones included
    assert r.num_centroids() == int(np.count_nonzero((lo["n_raw"] >= 6)))


def test_voxel_params(case_tiny):
    o, r = _pair(case_tiny)
    for m in (o, r):
        m.set_voxel_params(min_points=3, eig_mult=0.05)
        m.set_target(case_tiny[1])
    lo, lr = o.dump_leaves(), r.dump_leaves()
    for k in ("nr_points", "icov", "evals"):
        assert np.array_equal(lo[k], lr[k], equal_nan=True), k


@pytest.mark.parametrize("name", list(METHODS))
def test_neighbour_sets_bit_exact(case_small, name):
    o, r = _pair(case_small, method=METHODS[name])
    pts = o.transform_source(case_small[3])
    rng = np.random.default_rng(7)
    far = (pts[:200] + rng.uniform(-80, 80, (200, 3))).astype(np.float32)     # points outside the grid box too
    pts = np.concatenate([pts, far])
    co, io = o.neighbors(pts, METHODS[name])
    cr, ir = r.neighbors(pts, METHODS[name])
    assert np.array_equal(co, cr)
    if name == "KDTREE":
        # equal squared distances may be ordered either way by the two sorts: compare as sets per point
        for i in np.flatnonzero(co):
            assert set(map(tuple, io[i, :co[i]])) == set(map(tuple, ir[i, :cr[i]]))
        assert np.mean(np.all(io == ir, axis=(1, 2))) > 0.999
    else:
        assert np.array_equal(io, ir)
    assert co.max() > 1 or name == "DIRECT1"


@pytest.mark.parametrize("name", list(METHODS))
@pytest.mark.parametrize("ch", [True, False])
def test_compute_derivatives_bit_exact(case_small, name, ch):
    o, r = _pair(case_small, method=METHODS[name])
    guess = case_small[3]
    p = oracle.matrix_to_pose(guess)
    T = oracle.pose_to_matrix(p)
    assert np.array_equal(T, ref.pose_to_matrix(p))
    so, go, Ho, npairs = o.eval_derivatives(T, p, ch)
    sr, gr, Hr = r.eval_derivatives(T, p, ch)
    assert npairs > 1000
    assert so == sr and np.array_equal(go, gr) and np.array_equal(Ho, Hr)
    if ch:
        assert np.abs(Hr).max() > 0
    else:
        assert np.abs(Hr).max() == 0


def test_compute_derivatives_threads_change_only_summation_order(case_small):
    o, r = _pair(case_small, threads=4)
    guess = case_small[3]
    p = oracle.matrix_to_pose(guess)
    so, go, Ho, _ = o.eval_derivatives(guess, p, True)
    sr, gr, Hr = r.eval_derivatives(guess, p, True)
    assert abs(so - sr) <= 1e-12 * abs(sr)
    assert np.max(np.abs(go - gr)) <= 1e-12 * np.max(np.abs(gr)) and np.max(np.abs(Ho - Hr)) <= 1e-12 * np.max(np.abs(Hr))


def test_small_angle_snap(case_tiny):
    """computeAngleDerivatives snaps |angle| < 10e-5 to sin 0 / cos 1 (NDT_I:292-311)."""
    o, r = _pair(case_tiny)
    p = np.array([0.3, -0.2, 0.05, 5e-5, -9.9e-5, 2e-4])
    T = oracle.pose_to_matrix(p)
    so, go, Ho, _ = o.eval_derivatives(T, p, True)
    sr, gr, Hr = r.eval_derivatives(T, p, True)
    assert so == sr and np.array_equal(go, gr) and np.array_equal(Ho, Hr)


def test_compute_hessian_f64_and_calculate_score(case_tiny):
    o, r = _pair(case_tiny)
    guess = case_tiny[3]
    p = oracle.matrix_to_pose(guess)
    assert np.array_equal(o.compute_hessian(guess, p), r.compute_hessian(guess, p))
    assert o.calculate_score(guess) == r.calculate_score(guess)


def _same_align(o, r, guess):
    ro, rr = o.align(guess), r.align(guess)
    assert ro["iterations"] == rr["iterations"] and ro["converged"] == rr["converged"]
    assert np.array_equal(ro["T"], rr["T"])
    assert ro["trans_probability"] == rr["trans_probability"]
    assert np.array_equal(o.get_output(), r.get_output())
    return ro, rr


@pytest.mark.parametrize("name", list(METHODS))
def test_align_bit_exact(case_small, name):
    o, r = _pair(case_small, method=METHODS[name])
    ro, _ = _same_align(o, r, case_small[3])
    assert ro["iterations"] >= 2
    assert o.fitness() == r.fitness()
    assert o.fitness(0.5) == r.fitness(0.5)


def test_align_c1(case_c1):
    """BASELINE configs[0]: 20k-pt scan vs 300k-pt localmap, res 1.0, DIRECT7, <= 30 iterations."""
    o, r = _pair(case_c1, threads=1)
    ro, _ = _same_align(o, r, case_c1[3])
    assert ro["converged"]


def test_align_identity_guess(case_tiny):
    """guess == Identity skips the pre-transform (NDT_I:89-94)."""
    o, r = _pair(case_tiny)
    _same_align(o, r, np.eye(4, dtype=np.float32))


@pytest.mark.parametrize("res", [0.5, 2.0])
def test_align_resolutions(case_small, res):
    o, r = _pair(case_small, res=res)
    _same_align(o, r, case_small[3])


@pytest.mark.parametrize("step,eps,max_iter,trials", [(0.004, 0.01, 6, False), (0.5, 2.0, 4, True), (0.3, 1.0, 4, True)])
def test_align_more_thuente_live(case_tiny, step, eps, max_iter, trials):
    """step_size <= eps/2 makes the inverted predicate (NDT_I:807) run the More-Thuente loop and computeHessian.  The
    metre-sized steps fail the Wolfe test at the first trial: ten trialValueSelectionMT / updateIntervalMT rounds (kind 2),
    then computeHessian (kind 3, NDT_I:912-913)."""
    o, r = _pair(case_tiny, step=step, eps=eps, max_iter=max_iter)
    ro, _ = _same_align(o, r, case_tiny[3])
    kinds = [t["kind"] for t in ro["trace"]]
    assert any(k != 0 for k in kinds[1:])
    if trials:
        assert 2 in kinds and 3 in kinds


def test_align_iteration_cap(case_tiny):
    o, r = _pair(case_tiny, max_iter=2, eps=1e-9)
    ro, _ = _same_align(o, r, case_tiny[3])
    assert ro["iterations"] == 4          # nr_iterations_ > max_iterations_ is tested before the increment (NDT_I:152-157)


def test_more_thuente_pieces():
    r = ref.RefNDT()
    rng = np.random.default_rng(3)
    for _ in range(400):
        a = rng.uniform(0, 1, 3)
        f = rng.normal(0, 1, 3)
        g = rng.normal(0, 1, 3)
        args = (a[0], f[0], g[0], a[1], f[1], g[1], a[2], f[2], g[2])
        vo, vr = oracle.mt_trial(*args), r.mt_trial(*args)
        assert vo == vr or (np.isnan(vo) and np.isnan(vr))
        st = [a[0], f[0], g[0], a[1], f[1], g[1]]
        co, so = oracle.mt_update(st, a[2], f[2], g[2])
        cr, sr = r.mt_update(st, a[2], f[2], g[2])
        assert co == cr and np.array_equal(so, sr)


def test_golden_fixture_is_what_the_reference_sources_produce():
    """tests/golden/ndt_tiny.npz (made from the oracle) equals the reference build's outputs."""
    import os
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ndt_tiny.npz"))
    for name, method in (("d7", 2), ("d1", 3), ("kd", 0), ("d26", 1)):
        r = ref.RefNDT()
        r.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=method, num_threads=1)
        r.set_target(gold["target"])
        r.set_source(gold["source"])
        if name == "d7":
            L = r.dump_leaves()
            for k in ("ijk", "nr_points", "mean", "centroid"):
                assert np.array_equal(L[k], gold["vox_" + k]), k
            assert np.allclose(L["icov"], gold["vox_icov"], rtol=1e-12, atol=0, equal_nan=True)
        s, g, H = r.eval_derivatives(gold["guess"], gold[name + "_p0"], True)
        assert abs(s - gold[name + "_score0"]) <= 1e-12 * abs(s)
        assert np.max(np.abs(g - gold[name + "_g0"])) <= 1e-12 * np.max(np.abs(g))
        assert np.max(np.abs(H - gold[name + "_H0"])) <= 1e-12 * np.max(np.abs(H))
        a = r.align(gold["guess"])
        assert a["iterations"] == gold[name + "_iterations"] and int(a["converged"]) == gold[name + "_converged"]
        assert np.max(np.abs(a["T"] - gold[name + "_T"])) < 1e-6
        assert abs(r.fitness() - gold[name + "_fitness"]) <= 1e-9 * gold[name + "_fitness"]


def _edge_cases(case):
    src, tgt, gt, guess = case
    far = src.copy()
    far[:, :3] += 5000.0
    eye = np.eye(4, dtype=np.float32)
    return {"empty source": (np.zeros((0, 4), np.float32), tgt, guess),          # score 0/0: trans_probability NaN, converged (NDT_I:126-130)
            "source outside the grid": (far, tgt, eye),                          # no neighbour anywhere: H = 0, zero step
            "one-point source": (src[:1], tgt, guess),
            "no voxel reaches 6 points": (src, tgt[::40], guess)}


@pytest.mark.parametrize("name", ["empty source", "source outside the grid", "one-point source", "no voxel reaches 6 points"])
def test_edge_cases_bit_exact(case_tiny, name):
    s, t, G = _edge_cases(case_tiny)[name]
    o, r = _pair((s, t))
    ro, rr = o.align(G), r.align(G)
    assert ro["iterations"] == rr["iterations"] and ro["converged"] == rr["converged"]
    assert np.array_equal(ro["T"], rr["T"])
    assert ro["trans_probability"] == rr["trans_probability"] or (np.isnan(ro["trans_probability"]) and np.isnan(rr["trans_probability"]))
