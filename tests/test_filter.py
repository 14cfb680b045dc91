"""filter_node's front end (reference src/filter_node.cpp:236-262 -- NaN removal, range gate
1 < r_xy < 60, 0.5 m pcl::VoxelGrid, pcl::StatisticalOutlierRemoval k = 30 / 1 sigma) on the device (ndtb_filter_scan)
against the oracle's restatement (oracle/thirdparty_math.h).  Everything is float / double arithmetic in a fixed order and
an exact k-NN, so the gate is BIT-EXACT: same points, same order, same bytes."""
import numpy as np
import pytest


def _raw_scan(k=55, seed=123):
    from xchu_slam_b200 import synth
    w = synth.World(synth.BASE_SEED)
    poses = w.trajectory(70, seed=synth.BASE_SEED & 0xFFFF)
    raw = w.scan(poses[k], seed).copy()
    raw[:, 3] = np.random.default_rng(seed).uniform(0, 255, raw.shape[0]).astype(np.float32)
    return raw


def test_oracle_filter_scan_against_scipy():
    """CPU: the restated gate / VoxelGrid / SOR chain against numpy + scipy.cKDTree in float64."""
    from scipy.spatial import cKDTree
    from oracle import oracle
    raw = _raw_scan()
    raw[::1000, 0] = np.nan                      # removeNaNFromPointCloud
    r = oracle.filter_scan(raw)
    ok = np.isfinite(raw[:, :3]).all(1)
    rr = np.hypot(raw[:, 0].astype(np.float64), raw[:, 1].astype(np.float64))
    gated = raw[ok & (rr > 1.0) & (rr < 60.0)]
    assert r["counts"][0] == gated.shape[0]
    ds = oracle.voxelgrid_downsample(gated, 0.5)
    assert r["counts"][1] == ds.shape[0]
    d, _ = cKDTree(ds[:, :3].astype(np.float64)).query(ds[:, :3].astype(np.float64), k=31)
    md = d[:, 1:].mean(1)
    assert np.max(np.abs(md - r["distances"])) < 1e-5
    keep = md <= md.mean() + md.std(ddof=1)
    # the threshold is a float32-vs-float64 knife edge for at most a point or two
    assert abs(int(keep.sum()) - r["counts"][2]) <= 2
    assert 0.5 * ds.shape[0] < r["counts"][2] < ds.shape[0]
    # kept points are a subsequence of the downsampled cloud, in order
    idx = [i for i in range(ds.shape[0]) if (r["distances"][i] <= np.float64(md.mean() + md.std(ddof=1)) + 1e-4)]
    assert r["cloud"].shape[0] == r["counts"][2] and np.all(np.isin(r["cloud"].view([("", np.float32)] * 4), ds.view([("", np.float32)] * 4)))
    # fewer points than mean_k + 1: returned unchanged; empty in, empty out
    tiny = oracle.filter_scan(gated[:20], leaf=0.0001)
    assert tiny["counts"][2] == tiny["counts"][1] == 20
    assert oracle.filter_scan(np.zeros((0, 4), np.float32))["counts"] == (0, 0, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("stride32", [False, True])
def test_device_filter_scan_bit_exact(stride32):
    from oracle import oracle
    from xchu_slam_b200 import ndt
    g = ndt.NormalDistributionsTransform()
    for k, seed in ((55, 123), (60, 7)):
        raw = _raw_scan(k, seed)
        raw[::977, 1] = np.inf
        raw[5::1013, 2] = np.nan
        ro = oracle.filter_scan(raw)
        if stride32:                                # pcl::PointXYZI records
            rec = np.zeros((raw.shape[0], 8), np.float32)
            rec[:, :3] = raw[:, :3]; rec[:, 3] = 1.0; rec[:, 4] = raw[:, 3]
            counts = g.filterScan(rec)
        else:
            counts = g.filterScan(raw)
        assert counts == ro["counts"]
        out = g.getSource()
        assert out.shape == ro["cloud"].shape and np.array_equal(out, ro["cloud"])
    # the filtered scan IS the source: a registration runs on it directly
    from xchu_slam_b200 import synth
    src, tgt, gt, guess = synth.make_case(n_source=4000, n_target=50000, seed=synth.BASE_SEED + 5)
    g.setInputTarget(tgt)
    g.setTransformationEpsilon(0.01)
    g.filterScan(src, r_min=0.0, r_max=1e9, leaf=0.0, mean_k=0)        # pass-through
    assert np.array_equal(g.getSource(), src)
    g.align(guess)
    h = ndt.NormalDistributionsTransform()
    h.setInputTarget(tgt); h.setTransformationEpsilon(0.01); h.setInputSource(src); h.align(guess)
    assert np.array_equal(g.result()["T"], h.result()["T"]) and g.result()["iterations"] == h.result()["iterations"]


@pytest.mark.gpu
def test_device_filter_scan_edges():
    from oracle import oracle
    from xchu_slam_b200 import ndt
    g = ndt.NormalDistributionsTransform()
    raw = _raw_scan()
    # gate only / gate + downsample / different k and sigma / isolated far points (the exhaustive k-NN pass)
    for kw in (dict(leaf=0.0, mean_k=0), dict(leaf=1.0, mean_k=0), dict(leaf=0.5, mean_k=10, stddev_mul=0.5), dict(leaf=2.0, mean_k=30),
               dict(leaf=0.1, mean_k=8)):      # > 32768 points after the downsample: cell-grid ring search + sweep for the leftovers
        ro = oracle.filter_scan(raw, r_min=1.0, r_max=60.0, leaf=kw["leaf"] if kw["leaf"] > 0 else 1e-4, mean_k=kw["mean_k"] if kw["mean_k"] > 0 else 100000,
                                std_mul=kw.get("stddev_mul", 1.0)) if kw["leaf"] > 0 else None
        counts = g.filterScan(raw, **kw)
        if ro is not None:
            assert counts == ro["counts"] and np.array_equal(g.getSource(), ro["cloud"])
        else:
            assert counts[0] == counts[1] == counts[2] > 100000
    lonely = np.concatenate([raw[:40000], np.array([[45.0, 45.0, 30.0, 1.0], [-50.0, 20.0, 25.0, 2.0], [10.0, -55.0, 40.0, 3.0]], np.float32)])
    ro = oracle.filter_scan(lonely)
    assert g.filterScan(lonely) == ro["counts"] and np.array_equal(g.getSource(), ro["cloud"])
    assert g.filterScan(np.zeros((0, 4), np.float32)) == (0, 0, 0)
    few = raw[:25]
    ro = oracle.filter_scan(few)
    assert g.filterScan(few) == ro["counts"] and np.array_equal(g.getSource(), ro["cloud"])
    with pytest.raises(RuntimeError):
        g.filterScan(raw, mean_k=64)
