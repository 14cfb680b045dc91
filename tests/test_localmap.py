"""odom_node's localmap maintenance on the device (ndtb_localmap_*), against the oracle's
restatement of what LidarOdom::OdomEstimate does to its clouds either side of align (reference src/odom_node.cpp:218-233,
289-290, 329-356): pcl::transformPointCloud, pcl::VoxelGrid centroid downsample (downSizeFilterLocalmap), localmap /
tmp_map appends, the pc_target_ snapshot that lags one keyframe, the window roll.  Everything is float arithmetic in a
fixed order, so the gate is BIT-EXACT; the voxel table built from the device localmap must equal the one built from
the same cloud handed over from the host."""
import numpy as np
import pytest


def _scans(n_scans=7, n_points=6000):
    from xchu_slam_b200 import synth
    w = synth.World(synth.BASE_SEED)
    poses = w.trajectory(60 + n_scans, seed=synth.BASE_SEED & 0xFFFF)
    out = []
    rng = np.random.default_rng(11)
    for k in range(n_scans):
        s = synth.source_scan(w, poses[50 + k], synth.BASE_SEED + 50 + k, n_points)
        s = s.copy()
        s[:, 3] = rng.uniform(0, 255, s.shape[0]).astype(np.float32)     # intensities
        out.append((s, synth.pose6_to_matrix(synth.sensor_pose(poses[50 + k])).astype(np.float32)))
    return out


def test_oracle_voxelgrid_downsample_properties():
    """CPU: the restated pcl::VoxelGrid -- one point per occupied leaf, ascending (k, j, i), float centroids of x y z intensity."""
    from oracle import oracle
    rng = np.random.default_rng(5)
    p = rng.uniform(-30, 30, (20000, 4)).astype(np.float32)
    o = oracle.voxelgrid_downsample(p, 1.0)
    cells = np.floor(p[:, :3] * np.float32(1.0)).astype(np.int64)
    uniq = np.unique(cells, axis=0)
    assert o.shape[0] == uniq.shape[0]
    oc = np.floor(o[:, :3]).astype(np.int64)
    order = np.lexsort((oc[:, 0], oc[:, 1], oc[:, 2]))
    assert np.array_equal(order, np.arange(len(oc)))                     # ascending leaf index
    assert np.array_equal(oc, uniq[np.lexsort((uniq[:, 0], uniq[:, 1], uniq[:, 2]))])
    key = {tuple(c): i for i, c in enumerate(oc)}
    acc = np.zeros((len(oc), 4)); cnt = np.zeros(len(oc))
    for c, q in zip(cells, p):
        acc[key[tuple(c)]] += q; cnt[key[tuple(c)]] += 1
    assert np.max(np.abs(acc / cnt[:, None] - o)) < 1e-4
    # idempotent on cell occupancy; a leaf much larger than the cloud gives one point; empty in, empty out
    assert oracle.voxelgrid_downsample(o, 1.0).shape[0] == o.shape[0]
    assert oracle.voxelgrid_downsample(p, 1000.0).shape[0] <= 8
    assert oracle.voxelgrid_downsample(np.zeros((0, 4), np.float32), 1.0).shape[0] == 0
    # overflow guard: PCL returns the input unchanged
    far = np.array([[0, 0, 0, 1], [3e6, 3e6, 3e6, 2]], np.float32)
    assert np.array_equal(oracle.voxelgrid_downsample(far, 0.001), far)


@pytest.mark.gpu
def test_device_localmap_matches_oracle_bit_exact():
    from oracle import oracle
    from xchu_slam_b200 import ndt
    scans = _scans()
    g = ndt.NormalDistributionsTransform()
    g.setResolution(1.0)
    localmap = np.zeros((0, 4), np.float32)
    tmp_map = np.zeros((0, 4), np.float32)
    # first scan: raw transformed scan into the localmap only (ODOM:218-231)
    s0, T0 = scans[0]
    g.setInputSource(s0)
    n = g.localmapAppendSource(T0, leaf=0.0, also_tmp=False)
    localmap = np.concatenate([localmap, oracle.transform_cloud(s0, T0)])
    assert n == s0.shape[0] and np.array_equal(g.localmapDownload(0), localmap)
    g.localmapSnapshot()
    assert g.localmapSetTarget() is False           # first build is a full build
    host = ndt.NormalDistributionsTransform()
    host.setResolution(1.0)
    incs = []
    for k, (s, T) in enumerate(scans[1:], 1):
        g.localmapSnapshot()                        # pc_target_ = copy of the localmap (ODOM:233)
        snapshot = localmap.copy()
        g.setInputSource(s)
        n = g.localmapAppendSource(T, leaf=1.0, also_tmp=True)            # ODOM:290, 334-337
        ds = oracle.voxelgrid_downsample(oracle.transform_cloud(s, T), 1.0)
        assert n == ds.shape[0] and np.array_equal(g.localmapDownload(2), ds)
        localmap = np.concatenate([localmap, ds]); tmp_map = np.concatenate([tmp_map, ds])
        incs.append(g.localmapSetTarget())          # setInputTarget(pc_target_): lags the append by one keyframe (ODOM:349)
        host.setInputTarget(snapshot)
        a, b = g.dumpVoxels(), host.dumpVoxels()
        ia, ib = np.lexsort(a["ijk"].T[::-1]), np.lexsort(b["ijk"].T[::-1])
        for key in ("ijk", "nr_points", "n_raw", "mean", "icov", "centroid"):
            assert np.array_equal(a[key][ia], b[key][ib], equal_nan=True), (k, key)
        if k == 4:                                   # window roll (ODOM:352-356)
            g.localmapRoll()
            localmap, tmp_map = tmp_map, np.zeros((0, 4), np.float32)
        assert g.localmapSizes()[:2] == (localmap.shape[0], tmp_map.shape[0])
        assert np.array_equal(g.localmapDownload(0), localmap) and np.array_equal(g.localmapDownload(1), tmp_map)
    # every refresh merges only the new tail, except the one after the roll
    assert incs == [True, True, True, True, False, True]
    # a registration against the device-built target equals one against the host-built target
    s, T = scans[-1]
    g.setTransformationEpsilon(0.01); host.setTransformationEpsilon(0.01)
    g.setInputSource(s); host.setInputSource(s)
    g.align(T); host.align(T)
    assert g.result()["iterations"] == host.result()["iterations"] and np.array_equal(g.result()["T"], host.result()["T"])


@pytest.mark.gpu
def test_device_localmap_errors_and_edges():
    from xchu_slam_b200 import ndt
    g = ndt.NormalDistributionsTransform()
    with pytest.raises(RuntimeError):
        g.localmapAppendSource(np.eye(4, dtype=np.float32))              # no source yet
    with pytest.raises(RuntimeError):
        g.localmapSetTarget()                                            # no snapshot
    s, T = _scans(1, 3000)[0]
    g.setInputSource(s)
    assert g.localmapAppendSource(T, leaf=1000.0) <= 8                   # one huge leaf
    g.localmapClear()
    assert g.localmapSizes() == (0, 0, 0)
    # 32-byte PointXYZI records: intensity at byte offset 16 survives the trip
    rec = np.zeros((s.shape[0], 8), np.float32)
    rec[:, :3] = s[:, :3]; rec[:, 3] = 1.0; rec[:, 4] = np.arange(s.shape[0], dtype=np.float32)
    g.setInputSource(rec)
    g.localmapAppendSource(np.eye(4, dtype=np.float32), leaf=0.0)
    out = g.localmapDownload(0)
    assert np.array_equal(out[:, :3], s[:, :3]) and np.array_equal(out[:, 3], rec[:, 4])


def _surround_target_oracle(scans, ids, leaf_kf=1.0, leaf_target=0.5):
    """Oracle restatement of LidarOdom::ExtractSurroundKeyframes (reference src/odom_node.cpp:737-802): each keyframe moved by its
    pose (pcl::transformPointCloud, :749-751) and downsampled with downSizeFilterLocalmap (:752-753), concatenated in deque
    order (:778-781), downsampled with downSizeFilterKeyFrames (:783-784)."""
    from oracle import oracle
    clouds = [oracle.voxelgrid_downsample(oracle.transform_cloud(scans[i][0], scans[i][1]), leaf_kf) for i in ids]
    cat = np.concatenate(clouds, axis=0) if clouds else np.zeros((0, 4), np.float32)
    return clouds, oracle.voxelgrid_downsample(cat, leaf_target)


def test_oracle_surround_target_properties():
    """CPU: the window target is a function of the SET of keyframes in deque order -- newest first -- and one point per
    occupied 0.5 m leaf of the stacked 1.0 m-downsampled keyframes."""
    scans = _scans(n_scans=5, n_points=3000)
    clouds, tgt = _surround_target_oracle(scans, [4, 3, 2, 1, 0])
    assert all(len(c) <= len(s[0]) for c, s in zip(clouds, scans[::-1]))
    cells = np.unique(np.floor(np.concatenate(clouds)[:, :3] * np.float32(2.0)).astype(np.int64), axis=0)
    assert tgt.shape[0] == cells.shape[0]
    _, tgt2 = _surround_target_oracle(scans, [0, 1, 2, 3, 4])
    assert tgt2.shape == tgt.shape and not np.array_equal(tgt2, tgt)      # float sums follow the concatenation order


@pytest.mark.gpu
def test_device_surround_window_matches_oracle_bit_exact():
    """ndtb_window_*: the keyframe-count window kept in HBM.  Window clouds, the target cloud and the voxel table built from it
    must equal the oracle's restatement bit for bit, through pushes, a pop of the oldest keyframe and a skipped keyframe."""
    from oracle import oracle
    from xchu_slam_b200 import ndt
    scans = _scans(n_scans=8, n_points=6000)
    g = ndt.NormalDistributionsTransform()
    g.setResolution(2.0)
    ids = []
    for s, T in scans:
        g.setInputSource(s)
        i, n = g.windowAddSource(T, 1.0)
        ids.append(i)
        assert n == oracle.voxelgrid_downsample(oracle.transform_cloud(s, T), 1.0).shape[0]
    assert ids == list(range(8))
    o = oracle.OracleNDT()
    o.set_params(resolution=2.0)
    for order in ([2, 1, 0], [5, 4, 3, 2, 1], [7, 6, 4, 3, 2]):          # growing window; oldest popped; keyframe 5 skipped
        _, want = _surround_target_oracle(scans, order)
        n = g.windowSetTarget(order, 0.5)
        got = g.getTarget(n + 8)
        assert n == want.shape[0] and got.tobytes() == want.tobytes()
        o.set_target(want)
        lo, lg = o.dump_leaves(), g.dumpVoxels()
        so, sg = np.lexsort((lo["ijk"][:, 0], lo["ijk"][:, 1], lo["ijk"][:, 2])), np.lexsort((lg["ijk"][:, 0], lg["ijk"][:, 1], lg["ijk"][:, 2]))
        assert np.array_equal(lo["ijk"][so], lg["ijk"][sg]) and np.array_equal(lo["n_raw"][so], lg["n_raw"][sg])
        assert np.array_equal(lo["nr_points"][so], lg["nr_points"][sg]) and np.array_equal(lo["mean"][so], lg["mean"][sg])
        assert np.array_equal(lo["centroid"][so], lg["centroid"][sg])
    g.windowDrop(0)
    with pytest.raises(ndt.NdtError):
        g.windowSetTarget([0, 1], 0.5)
    g.windowClear()
    with pytest.raises(ndt.NdtError):
        g.windowSetTarget([1], 0.5)
