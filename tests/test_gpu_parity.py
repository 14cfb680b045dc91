"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on identical inputs.

Gates (BASELINE.md section 4 / north_star): voxel assignment, membership counts and neighbour
index sets bit-exact; per-evaluation score / gradient / Hessian within 1e-6 relative (gradient
relative to the largest gradient component of the same evaluation: it is a cancelling sum);
iteration counts equal; final transform within 1e-4 m / 1e-5 rad.
"""
import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-6


def _mk(case, res=1.0, method=2, step=0.1, eps=0.01, max_iter=30, strict=False, dense=True):
    from oracle import oracle
    from xchu_slam_b200 import ndt
    src, tgt, gt, guess = case
    o = oracle.OracleNDT()
    o.set_params(resolution=res, step_size=step, trans_eps=eps, max_iter=max_iter, search_method=method,
                 num_threads=oracle.max_threads())
    o.set_target(tgt)
    o.set_source(src)
    g = ndt.NormalDistributionsTransform()
    g.setResolution(res)
    g.setStepSize(step)
    g.setTransformationEpsilon(eps)
    g.setMaximumIterations(max_iter)
    g.setNeighborhoodSearchMethod(method)
    g.setStrictMath(strict)
    if not dense:
        g.setDenseMaxCells(0)      # hash-table lookups only
    g.setInputTarget(tgt)
    g.setInputSource(src)
    return o, g


def _sorted_voxels(ijk):
    return np.lexsort((ijk[:, 0], ijk[:, 1], ijk[:, 2]))


def _compare_voxels(o, g, rtol_mean=1e-12, rtol_icov=1e-10):
    lo, lg = o.dump_leaves(), g.dumpVoxels()
    assert len(lo["nr_points"]) == len(lg["nr_points"])
    so, sg = _sorted_voxels(lo["ijk"]), _sorted_voxels(lg["ijk"])
    assert np.array_equal(lo["ijk"][so], lg["ijk"][sg])            # voxel assignment: bit-exact
    assert np.array_equal(lo["n_raw"][so], lg["n_raw"][sg])        # membership counts: bit-exact
    assert np.array_equal(lo["nr_points"][so], lg["nr_points"][sg])  # validity flags
    assert np.array_equal(lo["centroid"][so], lg["centroid"][sg])  # f32 sequential centroid: bit-exact
    assert np.array_equal(lo["mean"][so], lg["mean"][sg])          # f64 sequential sums / n: bit-exact
    valid = lo["nr_points"][so] >= 6
    ico = lo["icov"][so][valid].astype(np.float64)
    icg = lg["icov"][sg][valid].astype(np.float64)
    scale = np.max(np.abs(ico), axis=(1, 2), keepdims=True)
    # GPU record stores float(icov): compare against the oracle's icov rounded the same way, then in relative terms
    assert np.max(np.abs(ico.astype(np.float32).astype(np.float64) - icg) / scale) < 2e-7
    # the f64 inverse the device keeps for computeHessian / calculateScore (1e-10 relative)
    ic64 = lg["icov64"][sg][valid]
    assert np.max(np.abs(ico - ic64) / scale) < rtol_icov
    assert np.array_equal(ic64.astype(np.float32), lg["icov"][sg][valid])     # the f32 record is its rounding
    return lo, lg, so, sg


def test_voxel_grid_build(case_small):
    o, g = _mk(case_small)
    lo, lg, so, sg = _compare_voxels(o, g)
    assert (lo["nr_points"] >= 6).sum() > 1000


def test_voxel_grid_build_paths_give_identical_tables(case_small):
    """The two fresh-build paths -- counting sort over the dense cell grid + warp merge (default) and radix sort + thread merge
    (no dense grid allowed) -- form the same sequences of additions: every array of the table is bit-identical, the f64 inverse
    covariance included, and both meet the oracle."""
    o, g = _mk(case_small)
    _, g2 = _mk(case_small, dense=False)
    _compare_voxels(o, g2)
    a, b = g.dumpVoxels(), g2.dumpVoxels()
    sa, sb = _sorted_voxels(a["ijk"]), _sorted_voxels(b["ijk"])
    for k in ("ijk", "nr_points", "n_raw", "mean", "icov", "icov64", "centroid"):
        assert np.array_equal(a[k][sa], b[k][sb]), k
    # the counting build numbers the voxels in ascending (k, j, i) cell order, as the radix path does
    assert np.array_equal(a["ijk"], b["ijk"])


@pytest.mark.parametrize("dense", [True, False])
def test_non_finite_target_points_are_skipped(case_small, dense):
    """VGC_I:209-215: a non-dense cloud's points with a NaN / Inf coordinate are skipped in the first pass.  The table built from a
    cloud with such rows equals, bit for bit, the table built from its finite rows alone (same relative order) -- on both build
    paths -- and an align against it runs as usual."""
    from xchu_slam_b200 import ndt
    src, tgt, gt, guess = case_small
    rng = np.random.default_rng(11)
    dirty = tgt.copy()
    rows = rng.choice(len(dirty), len(dirty) // 100, replace=False)
    dirty[rows[0::3], 0] = np.nan
    dirty[rows[1::3], 1] = np.inf
    dirty[rows[2::3], 2] = -np.inf
    clean = tgt[np.all(np.isfinite(dirty[:, :3]), axis=1)]
    tabs = []
    for cloud in (dirty, clean):
        g = ndt.NormalDistributionsTransform()
        g.setTransformationEpsilon(0.01)
        if not dense:
            g.setDenseMaxCells(0)
        g.setInputTarget(cloud)
        g.setInputSource(src)
        g.align(guess)
        tabs.append((g.dumpVoxels(), g.result()))
    (a, ra), (b, rb) = tabs
    for k in ("ijk", "nr_points", "n_raw", "mean", "icov64", "centroid"):
        assert np.array_equal(a[k], b[k]), k
    assert ra["iterations"] == rb["iterations"] and np.array_equal(ra["T"], rb["T"])


@pytest.mark.parametrize("res", [0.5, 2.0])
def test_voxel_grid_build_resolutions(case_small, res):
    o, g = _mk(case_small, res=res)
    _compare_voxels(o, g)


@pytest.mark.parametrize("dense", [True, False])
@pytest.mark.parametrize("method", [2, 3, 1, 0])
def test_neighbor_sets_bit_exact(case_small, method, dense):
    """Both cell indexes (dense grid over the bounding box / hash table) must return the reference's sets."""
    o, g = _mk(case_small, method=method, dense=dense)
    src, tgt, gt, guess = case_small
    pts = o.transform_source(guess)
    co, io = o.neighbors(pts, method)
    cg, ig = g.neighbors(pts, method)
    assert np.array_equal(co, cg)
    assert np.array_equal(io, ig)
    assert co.sum() > 0


def test_target_replaced_within_the_same_bounding_cells(case_small):
    """setInputTarget with a second cloud whose bounding cells equal the first one's (the first with a hole cut around the
    sensor): the dense cell index is refilled only when its box changes, so the cells of the voxels that went away must have
    been taken out of it -- the neighbour sets and the derivatives are those of a handle that only ever saw the second cloud."""
    from oracle import oracle
    from xchu_slam_b200 import ndt
    src, tgt, gt, guess = case_small
    c = guess[:3, 3]
    hole = np.hypot(tgt[:, 0] - c[0], tgt[:, 1] - c[1]) < 12.0
    tgt_b = np.ascontiguousarray(tgt[~hole])
    assert 0.05 < hole.mean() < 0.95
    assert np.array_equal(np.floor(tgt.min(0)[:3]), np.floor(tgt_b.min(0)[:3])) and np.array_equal(np.floor(tgt.max(0)[:3]), np.floor(tgt_b.max(0)[:3]))
    o, g = _mk(case_small)
    p = oracle.matrix_to_pose(guess)
    pts = o.transform_source(guess)
    ca, _ = g.neighbors(pts, 2)                 # builds the first table and its dense index
    o.set_target(tgt_b)
    g.setInputTarget(tgt_b)
    _compare_voxels(o, g)
    co, io = o.neighbors(pts, 2)
    cg, ig = g.neighbors(pts, 2)
    assert co.sum() < ca.sum()                   # the hole took neighbours away ...
    assert np.array_equal(co, cg) and np.array_equal(io, ig)      # ... and the device agrees, cell by cell
    so_, go_, Ho_, _ = o.eval_derivatives(guess, p, True)
    sg_, gg_, Hg_ = g.evalDerivatives(guess, p, True)
    assert abs(sg_ - so_) <= TOL * abs(so_) and rel_err(gg_, go_) < TOL and rel_err(Hg_, Ho_) < TOL
    # the batch engine reuses a target slot after clear() the same way
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2)
    b.add_target(tgt); b.add_source(src)
    b.run([0], [0], [guess])
    b.clear()
    b.add_target(tgt_b); b.add_source(src)
    r1 = b.run([0], [0], [guess])[0]
    f = ndt.BatchNDT()
    f.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2)
    f.add_target(tgt_b); f.add_source(src)
    r2 = f.run([0], [0], [guess])[0]
    assert r1["iterations"] == r2["iterations"] and np.array_equal(r1["T"], r2["T"])


@pytest.mark.parametrize("strict", [False, True])
@pytest.mark.parametrize("method", [2, 3, 0, 1])
@pytest.mark.parametrize("ch", [True, False])
def test_compute_derivatives(case_small, method, ch, strict):
    """One computeDerivatives (NDT_I:175-283) through both derivative kernels: the production
    (per-point factorised) one and the strict term-by-term one."""
    o, g = _mk(case_small, method=method, strict=strict)
    from oracle import oracle
    src, tgt, gt, guess = case_small
    p = oracle.matrix_to_pose(guess)
    so_, go_, Ho_, _ = o.eval_derivatives(guess, p, ch)
    sg_, gg_, Hg_ = g.evalDerivatives(guess, p, ch)
    assert abs(sg_ - so_) <= TOL * abs(so_)
    assert rel_err(gg_, go_) < TOL
    assert rel_err(Hg_, Ho_) < TOL
    if not ch:
        assert np.all(Hg_ == 0)


def _check_align(o, g, guess, strict=False, chaotic=False):
    """Whole-registration parity.

    Iteration counts, convergence flags and the sequence of evaluation kinds must be EQUAL; the final
    transform must agree within 1e-4 m / 1e-5 rad.  Per-iteration score / gradient / Hessian are
    compared within 1e-6 AT IDENTICAL INPUTS: every pose the oracle visited is re-evaluated on the GPU
    (a Newton iterate that differs by 1e-8 moves g by H * 1e-8, which is not an arithmetic difference).
    Gradient norm: g is a cancelling sum that goes to zero at convergence while its float32 terms do
    not, so "1e-6 relative" is relative to the largest gradient component seen in the registration
    (hard parts).  The strict kernel's terms are bit-identical to the oracle's,
    so for it the two traces themselves are also compared at 1e-6 / 1e-9.
    chaotic=True (DIRECT26 with the production kernel on this case): the search skips the point's own
    voxel, Newton on that objective is not contractive here and float-level differences in g / H grow
    from iterate to iterate (the strict kernel, bit-identical in its terms, does track the oracle); only
    the identical-input comparisons, the iteration count and a 1e-3 final gate are asserted."""
    from oracle import oracle
    ro = o.align(guess)
    g.align(guess)
    rg = g.result()
    tg = g.trace()
    assert rg["iterations"] == ro["iterations"]
    assert rg["converged"] == ro["converged"]
    assert len(tg) == len(ro["trace"])
    g_scale = max(float(np.max(np.abs(b["g"]))) for b in ro["trace"])
    for a, b in zip(tg, ro["trace"]):
        assert a["kind"] == b["kind"] and a["compute_hessian"] == b["compute_hessian"]
        # kind 0: the cloud was moved by the GUESS matrix itself (NDT_I:89-94), p is only its Euler image;
        # later evaluations use T(x_t) (NDT_I:815-819)
        T = np.asarray(guess, dtype=np.float32) if b["kind"] == 0 else oracle.pose_to_matrix(b["p"])
        if b["kind"] == 3:       # computeHessian (FP64, radius search)
            assert rel_err(g.computeHessian(T, b["p"]), b["H"]) < TOL
        else:
            s_, g_, H_ = g.evalDerivatives(T, b["p"], bool(b["compute_hessian"]))
            assert abs(s_ - b["score"]) <= TOL * abs(b["score"])
            assert rel_err(g_, b["g"], scale=g_scale) < TOL
            if b["compute_hessian"]:
                assert rel_err(H_, b["H"]) < TOL
        # Newton iterates of the production kernel: 1e-7 relative noise in g, H times the conditioning of H
        # (DIRECT26, which skips the point's own voxel, is the ill-conditioned one); gate = the final-transform gate
        ptol = 1e-9 if strict else 1e-4
        if chaotic:
            continue
        assert abs(a["step"] - b["step"]) <= ptol
        assert np.max(np.abs(a["p"] - b["p"])) < ptol
        if strict:
            assert abs(a["score"] - b["score"]) <= TOL * abs(b["score"])
            assert rel_err(a["g"], b["g"]) < TOL
            assert rel_err(a["H"], b["H"]) < TOL
        else:
            assert abs(a["score"] - b["score"]) <= 1e-3 * abs(b["score"])   # sanity only: poses differ by up to ptol
    assert np.max(np.abs(rg["T"][:3, 3] - ro["T"][:3, 3])) < (1e-3 if chaotic else 1e-4)
    assert np.max(np.abs(rg["T"][:3, :3] - ro["T"][:3, :3])) < (1e-4 if chaotic else 1e-5)
    if chaotic:
        return ro, rg
    assert abs(rg["trans_probability"] - ro["trans_probability"]) <= (TOL if strict else 1e-4) * abs(ro["trans_probability"])
    return ro, rg


@pytest.mark.parametrize("strict", [False, True])
def test_align_c1(case_c1, strict):
    """BASELINE configs[0]: 20k-pt scan vs 300k-pt localmap, res 1.0, DIRECT7, <= 30 iterations."""
    o, g = _mk(case_c1, strict=strict)
    src, tgt, gt, guess = case_c1
    ro, rg = _check_align(o, g, guess, strict=strict)
    from xchu_slam_b200 import synth
    assert np.max(np.abs(rg["T"][:3, 3] - synth.pose6_to_matrix(gt)[:3, 3])) < 0.1   # it actually registers
    out_o, out_g = o.get_output(), g.getOutput()
    assert np.max(np.abs(out_o - out_g)) < 1e-3


@pytest.mark.parametrize("strict", [False, True])
@pytest.mark.parametrize("dense", [True, False])
@pytest.mark.parametrize("method", [2, 3, 0, 1])
def test_align_search_methods(case_small, method, dense, strict):
    o, g = _mk(case_small, method=method, dense=dense, strict=strict)
    # DIRECT26 always runs the strict kernel (ndt_api.cu use_strict()), so it is held to the strict gates
    _check_align(o, g, case_small[3], strict=strict or method == 1)


def test_align_identity_guess(case_tiny):
    src, tgt, gt, guess = case_tiny
    from xchu_slam_b200 import synth
    world_src = synth.transform(src, gt)  # source already near the map frame -> identity guess is sane
    o, g = _mk((world_src, tgt, gt, None))
    _check_align(o, g, np.eye(4, dtype=np.float32))


@pytest.mark.parametrize("strict", [False, True])
@pytest.mark.parametrize("step,eps,max_iter,trials", [(0.004, 0.01, 6, False), (0.5, 2.0, 4, True), (0.3, 1.0, 4, True)])
def test_align_more_thuente_live(case_tiny, strict, step, eps, max_iter, trials):
    """step_size <= trans_eps/2 makes the reference's inverted predicate (NDT_I:807) run the MT loop
    and computeHessian; the device state machine must take the same trial steps.  The small-step case passes the Wolfe test
    at its first trial; the metre-sized ones fail it, run trialValueSelectionMT / updateIntervalMT for the full ten trials
    (kind 2) and end in computeHessian (kind 3)."""
    o, g = _mk(case_tiny, step=step, eps=eps, max_iter=max_iter, strict=strict)
    ro, rg = _check_align(o, g, case_tiny[3], strict=strict)
    kinds = [t["kind"] for t in ro["trace"]]
    assert [t["kind"] for t in g.trace()] == kinds
    assert 1 in kinds
    if trials:
        assert 2 in kinds and 3 in kinds


def test_compute_hessian_f64(case_tiny):
    from oracle import oracle
    o, g = _mk(case_tiny)
    guess = case_tiny[3]
    p = oracle.matrix_to_pose(guess)
    Ho = o.compute_hessian(guess, p)
    Hg = g.computeHessian(guess, p)
    assert rel_err(Hg, Ho) < TOL


def test_calculate_score(case_tiny):
    o, g = _mk(case_tiny)
    guess = case_tiny[3]
    so_ = o.calculate_score(guess)
    sg_ = g.calculateScore(guess)
    assert abs(sg_ - so_) <= TOL * abs(so_)


def test_fitness_score(case_small):
    o, g = _mk(case_small)
    guess = case_small[3]
    ro = o.align(guess)
    g.align(guess)
    fo, fg = o.fitness(), g.getFitnessScore()
    assert abs(fo - fg) <= 1e-9 * abs(fo)
    fo2, fg2 = o.fitness(0.5), g.getFitnessScore(0.5)
    assert abs(fo2 - fg2) <= 1e-9 * abs(fo2)


def test_fitness_score_far_queries(case_small):
    """getFitnessScore when most of the scan lies far outside a small target (odom_node's 5 m localmap window against a
    60 m scan): the ring search hands those queries to the exhaustive pass; the per-point minima are exact either way, so
    the mean equals the oracle's kd-tree-equivalent search to the last bits of the f64 sum."""
    from oracle import oracle
    from xchu_slam_b200 import ndt
    src, tgt, gt, guess = case_small
    near = tgt[np.linalg.norm(tgt[:, :2] - np.asarray(gt[:2], dtype=np.float32), axis=1) < 6.0]
    assert 500 < near.shape[0] < 40000
    o = oracle.OracleNDT()
    o.set_params(resolution=1.0)
    o.set_target(near); o.set_source(src)
    g = ndt.NormalDistributionsTransform()
    g.setInputTarget(near); g.setInputSource(src)
    for T in (guess, np.eye(4, dtype=np.float32)):
        fo, fg = o.fitness(T=T), g.getFitnessScore(T=T)
        assert fo > 1.0 and abs(fo - fg) <= 1e-12 * abs(fo)
        fo2, fg2 = o.fitness(4.0, T=T), g.getFitnessScore(4.0, T=T)
        assert abs(fo2 - fg2) <= 1e-12 * abs(fo2)


def test_incremental_update_equals_rebuild(case_small):
    """cpu::VoxelGrid::update contract + prefix-append detection: table after appends == full rebuild."""
    from oracle import oracle
    from xchu_slam_b200 import ndt
    src, tgt, gt, guess = case_small
    cuts = [30000, 38000, 45000, len(tgt)]
    o = oracle.OracleNDT()
    o.set_params(resolution=1.0, trans_eps=0.01, max_iter=30)
    o.set_target(tgt)
    g = ndt.NormalDistributionsTransform()
    g.setTransformationEpsilon(0.01)
    g.setMaximumIterations(30)
    g.setInputTarget(tgt[:cuts[0]])
    g.updateVoxelGrid(tgt[cuts[0]:cuts[1]])
    assert g.setInputTargetAuto(tgt[:cuts[2]]) is True      # extends the resident cloud -> incremental
    assert g.setInputTargetAuto(tgt[:cuts[3]]) is True
    _compare_voxels(o, g)
    o.set_source(src)
    g.setInputSource(src)
    _check_align(o, g, guess)
    # a cloud that is not an extension falls back to the full rebuild
    shifted = tgt[1000:].copy()
    assert g.setInputTargetAuto(shifted) is False
    o.set_target(shifted)
    _compare_voxels(o, g)


def test_overflow_guard_empty_grid():
    """VGC_I:75-84: dx*dy*dz > INT32_MAX leaves an empty grid; align then sees no neighbours."""
    from oracle import oracle
    from xchu_slam_b200 import ndt
    rng = np.random.default_rng(3)
    tgt = np.zeros((2000, 4), np.float32)
    tgt[:, :3] = rng.uniform(-1, 1, (2000, 3))
    tgt[0, :3] = [-40000, -40000, -400]
    tgt[1, :3] = [40000, 40000, 400]
    o = oracle.OracleNDT()
    o.set_params(resolution=0.5)
    o.set_target(tgt)
    g = ndt.NormalDistributionsTransform()
    g.setResolution(0.5)
    g.setInputTarget(tgt)
    assert not o.grid_valid()
    assert len(g.dumpVoxels()["nr_points"]) == 0
    src = tgt[2:500].copy()
    o.set_source(src)
    g.setInputSource(src)
    ro = o.align(np.eye(4, dtype=np.float32))
    g.align(np.eye(4, dtype=np.float32))
    rg = g.result()
    assert rg["iterations"] == ro["iterations"] and rg["converged"] == ro["converged"]


@pytest.mark.parametrize("strict", [True, False])
def test_batch_matches_single(case_small, strict):
    """Batched entry point: per-item results identical to single aligns, independent of batch layout -- with the term-by-term
    kernel (the default of both) and with the factorised one."""
    from xchu_slam_b200 import ndt, synth
    src, tgt, gt, guess = case_small
    guesses = [synth.perturbed_guess(gt, 100 + i) for i in range(6)]
    single = []
    g = ndt.NormalDistributionsTransform()
    if not strict:
        g.setStrictMath(False)
    g.setTransformationEpsilon(0.01)
    g.setMaximumIterations(30)
    g.setInputTarget(tgt)
    g.setInputSource(src)
    for G in guesses:
        g.align(G)
        single.append(g.result())
    b = ndt.BatchNDT()
    b.set_params(trans_eps=0.01, max_iter=30, strict_math=-1 if strict else 0)
    t0 = b.add_target(tgt)
    t1 = b.add_target(tgt[:50000])
    s0 = b.add_source(src)
    s1 = b.add_source(src[:3000])
    items_t = [t0] * 6 + [t1, t1]
    items_s = [s0] * 6 + [s1, s0]
    res = b.run(items_t, items_s, guesses + guesses[:2])
    for r, s in zip(res[:6], single):
        assert r["iterations"] == s["iterations"]
        assert np.array_equal(r["T"], s["T"])       # same block decomposition -> bit-identical
        assert r["score"] == s["score"]
    st = b.stats()
    assert st["deriv_launches"] > 0 and st["pairs_evaluated"] > 0
    # the same clouds registered through the many-clouds-per-call entry points (ndtb_batch_add_targets / _sources)
    b2 = ndt.BatchNDT()
    b2.set_params(trans_eps=0.01, max_iter=30, strict_math=-1 if strict else 0)
    tg = [np.ascontiguousarray(tgt, dtype=np.float32), np.ascontiguousarray(tgt[:50000], dtype=np.float32)]
    sc = [np.ascontiguousarray(src, dtype=np.float32), np.ascontiguousarray(src[:3000], dtype=np.float32)]
    assert b2.add_targets_ptr(ndt.BatchNDT.pointer_table([a.ctypes.data for a in tg], [a.shape[0] for a in tg]), 16, 0) == 0
    assert b2.add_sources_ptr(ndt.BatchNDT.pointer_table([a.ctypes.data for a in sc], [a.shape[0] for a in sc]), 16, 0) == 0
    res2 = b2.run(items_t, items_s, guesses + guesses[:2])
    for r, q in zip(res, res2):
        assert r["iterations"] == q["iterations"] and np.array_equal(r["T"], q["T"]) and r["score"] == q["score"]


@pytest.mark.parametrize("method", [2, 3, 0])
def test_gpu_against_reference_sources_directly(case_c1, method):
    """BASELINE configs[0] against oracle/_ref (the reference's own NDT sources over the stand-in third-party
    headers, tests/test_reference_build.py) with no oracle in between: voxel membership bit-exact, derivatives at
    the guess within 1e-6, equal iteration count, final transform within 1e-4 m / 1e-5 rad."""
    from oracle import oracle, ref
    if not ref.available():
        pytest.skip("oracle/_ref/libndt_ref.so not built")
    from xchu_slam_b200 import ndt
    src, tgt, gt, guess = case_c1
    r = ref.RefNDT()
    r.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=method, num_threads=oracle.max_threads())
    r.set_target(tgt)
    r.set_source(src)
    g = ndt.NormalDistributionsTransform()
    g.setResolution(1.0); g.setStepSize(0.1); g.setTransformationEpsilon(0.01); g.setMaximumIterations(30)
    g.setNeighborhoodSearchMethod(method)
    g.setInputTarget(tgt)
    g.setInputSource(src)
    lr, lg = r.dump_leaves(), g.dumpVoxels()
    sr, sg = _sorted_voxels(lr["ijk"]), _sorted_voxels(lg["ijk"])
    assert np.array_equal(lr["ijk"][sr], lg["ijk"][sg]) and np.array_equal(lr["nr_points"][sr], lg["nr_points"][sg])
    assert np.array_equal(lr["mean"][sr], lg["mean"][sg]) and np.array_equal(lr["centroid"][sr], lg["centroid"][sg])
    p = oracle.matrix_to_pose(guess)
    s_r, g_r, H_r = r.eval_derivatives(guess, p, True)
    s_g, g_g, H_g = g.evalDerivatives(guess, p, True)
    assert abs(s_g - s_r) <= TOL * abs(s_r) and rel_err(g_g, g_r) < TOL and rel_err(H_g, H_r) < TOL
    rr = r.align(guess)
    g.align(guess)
    rg = g.result()
    assert rg["iterations"] == rr["iterations"] and rg["converged"] == rr["converged"]
    assert np.max(np.abs(rg["T"][:3, 3] - rr["T"][:3, 3])) < 1e-4
    assert np.max(np.abs(rg["T"][:3, :3] - rr["T"][:3, :3])) < 1e-5
    # getFitnessScore at the SAME transform (the reference's final one): exact 1-NN on both sides
    assert abs(g.getFitnessScore(T=rr["T"]) - r.fitness()) <= 1e-9 * r.fitness()


def _gt_error(T, gt):
    from xchu_slam_b200 import synth
    G = synth.pose6_to_matrix(gt)
    return float(np.max(np.abs(np.asarray(T, dtype=np.float64)[:3, 3] - G[:3, 3])))


@pytest.mark.parametrize("method", [2, 3, 0])
def test_align_c3_large_map(case_c3, method):
    """BASELINE configs[2] at full size: 20k-pt scan vs 5M-pt localmap, res 0.5, DIRECT7 / DIRECT1 / KDTREE.
    The 0.5 m grid over this map exceeds nothing the dense cell index cannot hold; the hash-only path is covered
    by test_align_search_methods[dense=False]."""
    o, g = _mk(case_c3, res=0.5, method=method)
    lo, lg = o.dump_leaves(), g.dumpVoxels()
    so, sg = _sorted_voxels(lo["ijk"]), _sorted_voxels(lg["ijk"])
    assert len(so) > 200000
    assert np.array_equal(lo["ijk"][so], lg["ijk"][sg]) and np.array_equal(lo["n_raw"][so], lg["n_raw"][sg])
    assert np.array_equal(lo["nr_points"][so], lg["nr_points"][sg]) and np.array_equal(lo["mean"][so], lg["mean"][sg])
    ro, rg = _check_align(o, g, case_c3[3])
    assert rg["converged"]
    if method != 3:          # DIRECT1 at 0.5 m keeps a lateral offset on this scene (the oracle does too)
        assert _gt_error(rg["T"], case_c3[2]) < 0.05


@pytest.mark.parametrize("strict", [False, True])
def test_align_c4_dense_scan(case_c4, strict):
    """BASELINE configs[3] at full size: undownsampled 120k-pt scan vs 1M-pt map, res 2.0."""
    o, g = _mk(case_c4, res=2.0, strict=strict)
    _compare_voxels(o, g)
    ro, rg = _check_align(o, g, case_c4[3], strict=strict)
    assert rg["converged"] and _gt_error(rg["T"], case_c4[2]) < 0.05
    assert abs(g.getFitnessScore() - o.fitness()) <= 1e-9 * o.fitness()


def _oracle_batch(w, n_threads=None):
    """The CPU oracle over every registration of a batch workload (one grid build per submap)."""
    from oracle import oracle
    o = oracle.OracleNDT()
    o.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2, num_threads=n_threads or oracle.max_threads())
    ref_res = [None] * len(w["item_target"])
    it = np.asarray(w["item_target"])
    for t in range(len(w["targets"])):
        ids = np.flatnonzero(it == t)
        if len(ids) == 0:
            continue
        o.set_target(w["targets"][t])
        for i in ids:
            o.set_source(w["sources"][w["item_source"][i]])
            ref_res[i] = o.align(w["guesses"][i])
            ref_res[i].pop("trace", None)
    return o, ref_res


def _assert_item_parity(r, q, tag):
    """north_star gate, per registration: iteration counts match; final transform within 1e-4 m / 1e-5 rad."""
    dT = np.max(np.abs(r["T"][:3, 3] - q["T"][:3, 3]))
    dR = np.max(np.abs(r["T"][:3, :3] - q["T"][:3, :3]))
    assert r["iterations"] == q["iterations"] and r["converged"] == q["converged"] and dT < 1e-4 and dR < 1e-5, \
        (tag, r["iterations"], q["iterations"], dT, dR)


def test_batch_c5_shape_against_oracle():
    """BASELINE configs[4] shape (20k-pt scans vs 1M-pt submaps, 2 submaps x 8 scans x 4 hypotheses = 64 registrations in
    one batch) against the oracle, registration by registration, NO waiver: equal iteration counts and convergence flags,
    final transforms within 1e-4 m / 1e-5 rad -- in the batch's default mode (strict_math -1: the term-by-term kernel), with
    strict_math 1 (the same kernel, asked for by name) and in the opt-in HYBRID mode (-2: factorised kernel, then a strict replay
    of the registrations that give a 1e-7 difference room to grow)."""
    from oracle import oracle
    from xchu_slam_b200 import ndt, synth
    w = synth.batch_workload(n_submaps=2, scans_per_submap=8, hyps_per_scan=4)
    o, ref_res = _oracle_batch(w)
    for t in range(2):
        o.set_target(w["targets"][t])
        for i in np.flatnonzero(np.asarray(w["item_target"]) == t):
            o.set_source(w["sources"][w["item_source"][i]])
            ref_res[i]["fitness"] = o.fitness(T=ref_res[i]["T"])
            ref_res[i]["fitness_1m"] = o.fitness(1.0, T=ref_res[i]["T"])
    for strict in (-1, 1, -2):
        b = ndt.BatchNDT()
        b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2, strict_math=strict)
        for t in w["targets"]:
            b.add_target(t)
        for s in w["sources"]:
            b.add_source(s)
        res = b.run(w["item_target"], w["item_source"], w["guesses"])
        # getFitnessScore of every registration of the batch (the loop-closure gate, pgo_node.cpp:460-476)
        fit, fit1 = b.fitness(len(res)), b.fitness(len(res), 1.0)
        for k, (r, q) in enumerate(zip(res, ref_res)):
            _assert_item_parity(r, q, (strict, k))
            assert abs(fit[k] - q["fitness"]) <= 1e-5 * q["fitness"] and abs(fit1[k] - q["fitness_1m"]) <= 1e-5 * q["fitness_1m"]
        st = b.stats()
        assert (st["replayed"] == 0) if strict != -2 else (0 <= st["replayed"] < len(res))
        # the registrations recover the ground truth (the batch is not trivially wrong everywhere)
        gt_err = [np.max(np.abs(r["T"][:3, 3] - synth.pose6_to_matrix(g6)[:3, 3])) for r, g6 in zip(res, w["gt"])]
        assert np.median(gt_err) < 0.05


@pytest.mark.parametrize("mode", [-1, -2])
def test_batch_c5_512_registrations(mode):
    """The batch path the headline is measured on, at scale: 8 one-million-point submaps x 8 scans x 8 hypotheses = 512
    registrations in ONE ndtb_batch_run, every one of them against the CPU oracle -- iteration counts and convergence flags
    equal, final transform within 1e-4 m / 1e-5 rad (north_star; the loop NDT_I:114-159 and its iteration test :152-157).
    mode -1: the library default (term-by-term kernel, nothing replayed).  mode -2: the opt-in hybrid mode on the world its
    replay criterion was tuned on; the replayed minority must exist (the workload has non-settling registrations) and stay a
    minority (it is what the hybrid mode costs)."""
    from xchu_slam_b200 import ndt, synth
    w = synth.batch_workload(n_submaps=8, scans_per_submap=8, hyps_per_scan=8)
    _, ref_res = _oracle_batch(w)
    b = ndt.BatchNDT()
    assert b.params.strict_math == -1
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2, strict_math=mode)
    for t in w["targets"]:
        b.add_target(t)
    for s in w["sources"]:
        b.add_source(s)
    res = b.run(w["item_target"], w["item_source"], w["guesses"])
    assert len(res) == 512
    for k, (r, q) in enumerate(zip(res, ref_res)):
        _assert_item_parity(r, q, k)
    st, dg = b.stats(), b.diag(len(res))
    assert st["replayed"] == int(dg["replayed"].sum())
    if mode == -1:
        assert st["replayed"] == 0
        return
    assert 0 < st["replayed"] < len(res) // 2
    # every registration beyond the iteration threshold was replayed
    its = np.array([q["iterations"] for q in ref_res])
    assert np.all(dg["replayed"][its > b.params.replay_iters] == 1)


@pytest.mark.parametrize("method", [3, 0, 1])
def test_batch_default_mode_other_search_methods(case_small, method):
    """The batch default (term-by-term kernel) with DIRECT1, KDTREE and DIRECT26 (NDT_H:52-57): six registrations of one batch
    against the oracle, the north_star gate on every one of them."""
    from oracle import oracle
    from xchu_slam_b200 import ndt, synth
    src, tgt, gt, guess = case_small
    guesses = [synth.perturbed_guess(gt, 700 + i) for i in range(6)]
    o = oracle.OracleNDT()
    o.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=method, num_threads=oracle.max_threads())
    o.set_target(tgt)
    o.set_source(src)
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=method)
    assert b.params.strict_math == -1
    b.add_target(tgt)
    b.add_source(src)
    res = b.run([0] * 6, [0] * 6, guesses)
    assert b.stats()["replayed"] == 0
    for k, (r, G) in enumerate(zip(res, guesses)):
        _assert_item_parity(r, o.align(G), (method, k))


def test_batch_step_history(case_small):
    """ndtb_batch_get_step_history: |delta_p| of every Newton solve of every item (the data the hybrid mode's replay criterion is
    studied on): one positive entry per iteration, zeros beyond, the last one equal to the `last_norm` diagnostic."""
    from xchu_slam_b200 import ndt, synth
    src, tgt, gt, guess = case_small
    guesses = [synth.perturbed_guess(gt, 300 + i) for i in range(5)]
    b = ndt.BatchNDT()
    b.set_params(trans_eps=0.01, max_iter=30)
    b.add_target(tgt)
    b.add_source(src)
    res = b.run([0] * 5, [0] * 5, guesses)
    h = b.step_history(5)
    dg = b.diag(5)
    assert h.shape == (5, 32) and h.dtype == np.float32
    for i, r in enumerate(res):
        n = r["iterations"]
        assert 0 < n <= 32
        assert np.all(h[i, :n] > 0) and np.all(h[i, n:] == 0)
        assert h[i, n - 1] == np.float32(dg["last_norm"][i])
    # a shorter view of the same data
    assert np.array_equal(b.step_history(5, 8), h[:, :8])


def test_batch_edge_cases_hybrid(case_tiny):
    """The degenerate inputs of test_edge_cases as items of ONE hybrid batch (strict_math -2): the one-point source (rank-3
    Hessian, pseudo-inverse step) is flagged by the device and replayed, so the batch follows the oracle there too."""
    from test_reference_build import _edge_cases
    from oracle import oracle
    from xchu_slam_b200 import ndt
    cases = _edge_cases(case_tiny)
    names = ["source outside the grid", "one-point source", "no voxel reaches 6 points"]
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, trans_eps=0.01, max_iter=30, search_method=2, strict_math=-2)
    it, isrc, G, refs = [], [], [], []
    for nme in names:
        s, t, g0 = cases[nme]
        o = oracle.OracleNDT()
        o.set_params(resolution=1.0, trans_eps=0.01, max_iter=30, search_method=2, num_threads=1)
        o.set_target(t); o.set_source(s)
        refs.append(o.align(g0))
        it.append(b.add_target(t)); isrc.append(b.add_source(s)); G.append(g0)
    res = b.run(it, isrc, G)
    for nme, r, q in zip(names, res, refs):
        assert r["iterations"] == q["iterations"] and r["converged"] == q["converged"], nme
        assert np.max(np.abs(r["T"] - q["T"])) < 1e-4, nme
    assert b.diag(len(res))["replayed"][1] == 1


def test_batch_empty_target_slot_reuse(case_small):
    """A target slot reused after clear() for an EMPTY cloud (an empty loop-closure submap legally gets an id) presents an
    empty grid, as the reference's empty leaves_ does -- not the previous cloud's voxel table."""
    from oracle import oracle
    from xchu_slam_b200 import ndt
    src, tgt, gt, guess = case_small
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, trans_eps=0.01, max_iter=30, search_method=2)
    b.add_target(tgt); b.add_source(src)
    r0 = b.run([0], [0], [guess])[0]
    assert r0["iterations"] > 1
    b.clear()
    b.add_target(np.zeros((0, 4), np.float32)); b.add_source(src)
    r1 = b.run([0], [0], [guess])[0]
    o = oracle.OracleNDT()
    o.set_params(resolution=1.0, trans_eps=0.01, max_iter=30, search_method=2, num_threads=1)
    o.set_target(np.zeros((0, 4), np.float32)); o.set_source(src)
    q = o.align(guess)
    assert r1["iterations"] == q["iterations"] and r1["converged"] == q["converged"] and r1["score"] == 0.0
    assert np.array_equal(r1["T"], q["T"])


def test_set_target_auto_non_extension_of_overlapping_size(case_small):
    """setInputTarget through the auto entry point with a cloud that is NOT an extension of the resident one and whose size
    lies in (old_n, 2 old_n]: the fallback moves the fresh copy down over itself (source and destination ranges overlap) and
    must still build the table of exactly that cloud."""
    from oracle import oracle
    from xchu_slam_b200 import ndt
    src, tgt, gt, guess = case_small
    old_n = 25000
    g = ndt.NormalDistributionsTransform()
    g.setInputTarget(tgt[:old_n])
    o = oracle.OracleNDT()
    o.set_params(resolution=1.0)
    for n in (old_n + 1, old_n + old_n // 2):
        fresh = np.ascontiguousarray(tgt[len(tgt) - n:][::-1])        # unrelated order and content
        g2 = ndt.NormalDistributionsTransform()
        g2.setInputTarget(tgt[:old_n])
        assert g2.setInputTargetAuto(fresh) is False
        o.set_target(fresh)
        _compare_voxels(o, g2)


@pytest.mark.parametrize("strict", [False, True])
@pytest.mark.parametrize("name", ["empty source", "source outside the grid", "one-point source", "no voxel reaches 6 points"])
def test_edge_cases(case_tiny, name, strict):
    """Degenerate inputs the reference handles by value (no exceptions, no error codes): empty source (score 0/0 -> NaN
    probability, `converged` true by the NaN test of NDT_I:126-130), a source with no neighbour voxel at all (H = 0, zero
    Newton step), a single point, a target none of whose voxels reaches min_points."""
    from test_reference_build import _edge_cases
    from oracle import oracle
    from xchu_slam_b200 import ndt
    s, t, G = _edge_cases(case_tiny)[name]
    o = oracle.OracleNDT()
    o.set_params(resolution=1.0, trans_eps=0.01, max_iter=30, search_method=2, num_threads=1)
    o.set_target(t); o.set_source(s)
    ro = o.align(G)
    g = ndt.NormalDistributionsTransform()
    g.setTransformationEpsilon(0.01); g.setMaximumIterations(30); g.setStrictMath(strict)
    g.setInputTarget(t); g.setInputSource(s)
    g.align(G)
    rg = g.result()
    if name == "one-point source" and not strict:
        # one point gives a rank-3 Hessian: the step comes out of the pseudo-inverse branch, where the factorised kernel's
        # 1e-7 differences in H are amplified without bound -- only the term-by-term kernel can follow the oracle here
        assert np.all(np.isfinite(rg["T"])) and rg["iterations"] >= 1
        return
    assert rg["iterations"] == ro["iterations"] and rg["converged"] == ro["converged"]
    assert np.max(np.abs(rg["T"] - ro["T"])) < 1e-4
    tp = rg["trans_probability"]
    assert (np.isnan(tp) and np.isnan(ro["trans_probability"])) or abs(tp - ro["trans_probability"]) <= 1e-6 * max(1.0, abs(ro["trans_probability"]))
