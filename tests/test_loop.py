"""loop-closure refinement (reference src/pgo_node.cpp:404-476, 530-554, 568-590) with the
keyframe store, the submap assembly, the registration of every candidate and the fitness gate on the device.

Parity bars: the clouds LoopFindNearKeyframesCloud produces (TransformCloud2Map per keyframe, stack, pcl::VoxelGrid) are
float arithmetic in a fixed order -> BIT-EXACT against the oracle's restatement; the registrations are the batched NDT of
the hot path (ICP is PCL's, outside /root/reference, and is what this row replaces) -> oracle NDT on the same clouds, final
transform within 1e-4, equal iteration counts; getFitnessScore within 1e-6 relative; accept / reject decisions equal."""
import numpy as np
import pytest


def _keyframes(n_kf=14, n_points=5000, drift=True):
    """Keyframe clouds in the sensor frame + noisy odometry poses (x y z roll pitch yaw), seeded."""
    from xchu_slam_b200 import synth
    w = synth.World(synth.BASE_SEED)
    poses = w.trajectory(60 + 2 * n_kf, seed=synth.BASE_SEED & 0xFFFF)
    rng = np.random.default_rng(23)
    clouds, pose6 = [], []
    for k in range(n_kf):
        vp = poses[50 + 2 * k]
        s = synth.source_scan(w, vp, synth.BASE_SEED + 300 + k, n_points).copy()
        s[:, 3] = rng.uniform(0, 255, s.shape[0]).astype(np.float32)
        p = synth.sensor_pose(vp)
        if drift:
            p = p + np.array([rng.normal(0, 0.05), rng.normal(0, 0.05), rng.normal(0, 0.01), rng.normal(0, 0.002),
                              rng.normal(0, 0.002), rng.normal(0, 0.004)])
        clouds.append(s)
        pose6.append(p)
    return clouds, pose6


def test_get_transformation_host_mirror_matches_oracle_bitwise():
    """CPU: the product's pcl::getTransformation restatement and the oracle's are written independently and agree bit for bit;
    the matrix is Rz(yaw) Ry(pitch) Rx(roll) to float accuracy."""
    from oracle import oracle
    from xchu_slam_b200 import loop, synth
    rng = np.random.default_rng(3)
    for _ in range(200):
        p = np.concatenate([rng.uniform(-500, 500, 3), rng.uniform(-np.pi, np.pi, 3)])
        a, b = loop.get_transformation(*p), oracle.get_transformation(*p)
        assert a.dtype == np.float32 and a.tobytes() == b.tobytes()
        assert np.max(np.abs(a.astype(np.float64) - synth.pose6_to_matrix(p))) < 1e-4    # translation rounds at ~3e-5 near 500 m
        assert np.max(np.abs(a[:3, :3].astype(np.float64) - synth.pose6_to_matrix(p)[:3, :3])) < 3e-7
    z = loop.get_transformation(0, 0, 0, 0, 0, 0)
    assert np.array_equal(z, np.eye(4, dtype=np.float32))


def test_oracle_submap_window_and_ignored_key():
    """CPU: LoopFindNearKeyframesCloud's window is clipped at both ends of the keyframe list, `key` is ignored (pgo_node.cpp:537),
    a window with no keyframe gives an empty cloud, leaf <= 0 returns the plain stack."""
    from oracle import oracle
    rng = np.random.default_rng(9)
    kfs = [rng.uniform(-20, 20, (300 + 10 * i, 4)).astype(np.float32) for i in range(6)]
    P = [oracle.get_transformation(i, -i, 0.1 * i, 0.01 * i, -0.02 * i, 0.3 * i) for i in range(6)]
    full = oracle.loop_find_near_keyframes_cloud(kfs, P, 0, 2, 1, 0.0)
    want = np.concatenate([oracle.transform_cloud(kfs[k], P[k]) for k in (0, 1, 2, 3)])       # -1 is out of range
    assert full.tobytes() == want.tobytes()
    a = oracle.loop_find_near_keyframes_cloud(kfs, P, 5, 1, 4, 0.5)
    b = oracle.loop_find_near_keyframes_cloud(kfs, P, 0, 1, 4, 0.5)
    assert a.tobytes() == b.tobytes() and a.shape[0] > 0
    stack = np.concatenate([oracle.transform_cloud(kfs[k], P[k]) for k in (3, 4, 5)])
    assert a.tobytes() == oracle.voxelgrid_downsample(stack, 0.5).tobytes()
    assert oracle.loop_find_near_keyframes_cloud(kfs, P, 0, 1, 40, 0.5).shape == (0, 4)
    assert oracle.loop_find_near_keyframes_cloud(kfs, P, 0, 3, -2, 0.0).tobytes() == np.concatenate(
        [oracle.transform_cloud(kfs[k], P[k]) for k in (0, 1)]).tobytes()


def _oracle_refine(kfs, pose6, pairs, leaf, half_window, threshold, mode, ndt_params):
    from oracle import oracle
    P = [oracle.get_transformation(*p) for p in pose6]
    out = []
    for loop_idx, curr_idx in pairs:
        tgt = oracle.loop_find_near_keyframes_cloud(kfs, P, loop_idx, half_window, loop_idx, leaf)
        if mode == "reference":
            src = oracle.loop_find_near_keyframes_cloud(kfs, P, curr_idx, 0, loop_idx, leaf)
        elif mode == "current":
            src = oracle.loop_find_near_keyframes_cloud(kfs, P, curr_idx, 0, curr_idx, leaf)
        else:
            ok = 0 <= curr_idx < len(kfs) and 0 <= loop_idx < len(kfs)
            src = (oracle.voxelgrid_downsample(oracle.transform_cloud(kfs[curr_idx], P[loop_idx]), leaf) if ok
                   else np.zeros((0, 4), np.float32))
        r = dict(source=src, target=tgt, accepted=False, converged=False, iterations=0, fitness=np.finfo(np.float64).max,
                 T=np.eye(4, dtype=np.float32))
        if src.shape[0] and tgt.shape[0]:
            o = oracle.OracleNDT()
            o.set_params(**ndt_params)
            o.set_target(tgt)
            o.set_source(src)
            a = o.align(np.eye(4, dtype=np.float32))
            fit = o.fitness()
            r.update(converged=a["converged"], iterations=a["iterations"], T=a["T"], fitness=fit,
                     accepted=bool(a["converged"] and fit <= threshold))
        out.append(r)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["reference", "current", "current_at_loop_pose"])
def test_loop_refiner_against_oracle(mode):
    from xchu_slam_b200 import loop
    kfs, pose6 = _keyframes()
    params = dict(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30)
    ref = loop.LoopRefiner(0, filter_size=0.5, history_keyframe_search_num=4, loop_fitness_score_threshold=0.3, source_mode=mode,
                           **params)
    for c, p in zip(kfs, pose6):
        ref.add_keyframe(c, p)
    assert ref.num_keyframes() == len(kfs) and ref.batch.num_keyframes() == len(kfs)
    #        interior      window clipped at 0   clipped at the end   same loop twice   loop keyframe out of range
    pairs = [(5, 9), (1, 8), (12, 3), (5, 11), (40, 2), (-3, 5)]
    got = ref.refine(pairs, download_clouds=True)
    want = _oracle_refine(kfs, pose6, pairs, 0.5, 4, 0.3, mode, dict(params, max_iter=30))
    n_live = 0
    for (lp, cu), g, w in zip(pairs, got, want):
        assert g["source"].tobytes() == w["source"].tobytes(), (lp, cu, "source cloud")
        assert g["target"].tobytes() == w["target"].tobytes(), (lp, cu, "target cloud")
        if w["source"].shape[0] == 0 or w["target"].shape[0] == 0:
            assert not g["accepted"] and not g["converged"]
            continue
        n_live += 1
        assert g["iterations"] == w["iterations"] and g["converged"] == w["converged"], (lp, cu)
        assert np.max(np.abs(g["T"] - w["T"])) < 1e-4, (lp, cu)
        assert abs(g["fitness"] - w["fitness"]) <= 1e-6 * max(abs(w["fitness"]), 1e-3), (lp, cu, g["fitness"], w["fitness"])
        assert g["accepted"] == w["accepted"], (lp, cu)
    assert n_live >= 4
    if mode == "reference":
        # the reference's source is the loop keyframe itself, already inside the target: accepted, and the correction stays small
        # (NDT settles a few cm to dm off identity: the source is one of nine noisy-posed keyframes of the target)
        for g in got[:4]:
            assert g["accepted"] and np.max(np.abs(g["T"][:3, 3])) < 0.5 and np.max(np.abs(g["T"][:3, :3] - np.eye(3))) < 1e-2


@pytest.mark.gpu
def test_pose_update_changes_submaps_and_keyframes_survive_clear():
    """keyframePosesUpdated changes after a graph update: the same stored keyframes give the submap of the new poses; the
    keyframe store outlives ndtb_batch_clear; leaf <= 0 returns the undownsampled stack."""
    from oracle import oracle
    from xchu_slam_b200 import loop
    kfs, pose6 = _keyframes(n_kf=6, n_points=3000)
    ref = loop.LoopRefiner(0, filter_size=0.4, history_keyframe_search_num=2)
    for c, p in zip(kfs, pose6):
        ref.add_keyframe(c, p)
    b = ref.batch
    for rnd in range(2):
        P = [oracle.get_transformation(*p) for p in pose6]
        b.clear()
        tid, tn = b.add_submap(3, 2, np.stack(ref.poses), 0.4, as_target=True)
        assert b.download(tid, tn).tobytes() == oracle.loop_find_near_keyframes_cloud(kfs, P, 0, 2, 3, 0.4).tobytes()
        sid, sn = b.add_submap(0, 1, np.stack(ref.poses), 0.0, as_target=False)
        assert sn == kfs[0].shape[0] + kfs[1].shape[0]
        assert b.download(sid, sn, is_target=False).tobytes() == oracle.loop_find_near_keyframes_cloud(kfs, P, 0, 1, 0, 0.0).tobytes()
        pose6 = [p + np.array([0.3, -0.2, 0.05, 0.001, -0.002, 0.01]) * (i + 1) for i, p in enumerate(pose6)]
        ref.update_poses(pose6)
    b.clear_keyframes()
    assert b.num_keyframes() == 0
    _, n = b.add_submap(3, 2, np.zeros((0, 4, 4), np.float32), 0.4)
    assert n == 0
