"""The two loop-closure rows of SURVEY.md section 8(f) working together the way pgo_node drives them (reference
src/pgo_node.cpp:236, 340-362, 404-476): every keyframe gets a ScanContext descriptor, detectLoopClosureID proposes
(loop, current) pairs, the refinement registers each pair and gates it on the fitness score -- all on the device, every
decision equal to the CPU oracles'."""
import numpy as np
import pytest


def _session(n_kf=64):
    """A drive of n_kf keyframes followed by two revisits (keyframes 5 and 12 seen again with fresh sensor noise) whose
    odometry poses have drifted by a few decimetres and about a degree."""
    from xchu_slam_b200 import synth
    w = synth.World(synth.BASE_SEED)
    poses = w.trajectory(60 + 3 * n_kf, seed=synth.BASE_SEED & 0xFFFF)
    clouds, pose6, truth = [], [], []
    for k in range(n_kf):
        vp = poses[50 + 3 * k]
        clouds.append(synth.voxel_downsample(synth.range_gate(w.scan(vp, 5000 + k), 1.0, 80.0), 0.5))
        pose6.append(synth.sensor_pose(vp))
        truth.append(k)
    drift = np.array([0.30, -0.25, 0.05, 0.0, 0.0, np.deg2rad(1.0)])
    for j, k in enumerate((5, 12)):
        vp = poses[50 + 3 * k]
        clouds.append(synth.voxel_downsample(synth.range_gate(w.scan(vp, 9000 + j), 1.0, 80.0), 0.5))
        pose6.append(synth.sensor_pose(vp) + drift)
        truth.append(k)
    return clouds, pose6, drift


@pytest.mark.gpu
def test_scancontext_candidates_refined_by_batched_ndt():
    from oracle import sc
    from xchu_slam_b200 import loop, scancontext
    from test_loop import _oracle_refine
    clouds, pose6, drift = _session()
    scm, sco = scancontext.SCManager(0), sc.OracleSC()
    params = dict(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30)
    ref = loop.LoopRefiner(0, filter_size=0.5, history_keyframe_search_num=3, source_mode="current", **params)
    pairs = []
    for k, (c, p) in enumerate(zip(clouds, pose6)):
        ref.add_keyframe(c, p)
        scm.makeAndSaveScancontextAndKeys(c)
        sco.add(c)
        got, want = scm.detectLoopClosureID(), sco.detect()
        assert got[0] == want[0] and np.float32(got[1]).tobytes() == np.float32(want[1]).tobytes(), (k, got, want)
        if got[0] >= 0:
            pairs.append((got[0], k))                      # pgo_node.cpp:352-360: loop_queue_.push(prev_node_idx, curr_node_idx)
    assert pairs == [(5, 64), (12, 65)], pairs              # both revisits proposed, nothing else
    res = ref.refine(pairs)
    want = _oracle_refine(clouds, pose6, pairs, 0.5, 3, 0.3, "current", params)
    for g, o in zip(res, want):
        assert g["accepted"] and o["accepted"]
        assert g["iterations"] == o["iterations"] and np.max(np.abs(g["T"] - o["T"])) < 1e-4
        assert abs(g["fitness"] - o["fitness"]) <= 1e-6 * abs(o["fitness"])
        # the correction undoes the injected drift (0.3 m, 1 degree): applying T to the drifted pose puts the sensor back within
        # centimetres in the plane (NDT against a seven-keyframe submap; the oracle lands on the same pose)
        P = loop.get_transformation(*pose6[g["curr"]]).astype(np.float64)
        fixed = g["T"].astype(np.float64) @ P
        true = loop.get_transformation(*(np.array(pose6[g["curr"]]) - drift)).astype(np.float64)
        assert np.max(np.abs(fixed[:2, 3] - true[:2, 3])) < 0.08 and abs(fixed[2, 3] - true[2, 3]) < 0.2, fixed[:3, 3] - true[:3, 3]
        assert np.max(np.abs(fixed[:3, :3] - true[:3, :3])) < 8e-3
