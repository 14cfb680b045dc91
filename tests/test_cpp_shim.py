"""The C++ host side above the C ABI: include/pclomp/ndt_omp.h (drop-in for the reference header) driven
by examples/odom_sequence.cpp, the headless mirror of odom_node's loop (reference
src/odom_node.cpp:208-356).  CPU: it compiles, links and fails loudly without a device.  GPU: a short
sequence tracks ground truth, localmap refreshes take the incremental path, and the incremental voxel
rebuild gives the same trajectory as full rebuilds."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def exe():
    from xchu_slam_b200 import build
    build.build_cuda()
    return build.build_examples()


def test_shim_compiles_and_fails_loudly_without_gpu(exe, tmp_path):
    import torch
    assert os.path.exists(exe)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from xchu_slam_b200 import synth
    seq = synth.write_sequence(str(tmp_path / "s.bin"), n_scans=2, n_points=2000)
    r = subprocess.run([exe, seq, str(tmp_path / "o.txt")], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr


def _run(exe, seq, out, *flags):
    r = subprocess.run([exe, seq, out] + list(flags), capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rows = [l.split() for l in open(out) if not l.startswith("#")]
    summary = [l for l in open(out) if l.startswith("#")][0].split()
    return np.array(rows, dtype=np.float64), dict(zip(summary[1::2], summary[2::2]))


@pytest.mark.gpu
def test_odometry_sequence_tracks_ground_truth_and_refreshes_incrementally(exe, tmp_path):
    from xchu_slam_b200 import synth
    seq = synth.write_sequence(str(tmp_path / "seq.bin"), n_scans=40, n_points=20000)
    a, sa = _run(exe, seq, str(tmp_path / "inc.txt"), "--res", "1.0")
    b, sb = _run(exe, seq, str(tmp_path / "full.txt"), "--res", "1.0", "--no-incremental")
    assert a.shape[0] == 40
    err = np.linalg.norm(a[:, 4:6] - a[:, 14:16], axis=1)     # xy drift against ground truth
    assert err.max() < 0.5 and np.all(a[:, 2] == 1)           # converged every scan, no divergence
    assert int(sa["target_sets"]) >= 30
    assert int(sa["incremental"]) >= int(sa["target_sets"]) - 8   # every refresh except window resets (every ~6 scans) appends
    assert int(sb["incremental"]) == 0
    # incremental voxel rebuild == full rebuild: identical tables -> identical trajectories
    assert np.array_equal(a[:, 1:3], b[:, 1:3])
    assert np.array_equal(a[:, 4:10], b[:, 4:10])


def _pcl_mode_exe():
    from oracle import ref_examples
    exe = ref_examples.build_shim_pcl_mode()
    if exe is None:
        pytest.skip("neither /root/reference (to build it) nor a prebuilt examples/shim_pcl_mode")
    return exe


def test_header_in_pcl_build_derives_from_registration_and_converts_transforms(exe):
    """include/pclomp/ndt_omp.h compiled in its PCL build (over the stand-in PCL headers): the static_assert in
    tests/cpp/shim_pcl_mode.cpp holds (derived from pcl::Registration, NDT_H:70-71) and both static convertTransform overloads
    (NDT_H:210-229) give the oracle's float chain bit for bit -- no device involved."""
    from oracle import oracle
    r = subprocess.run([_pcl_mode_exe(), "--static-only"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = [float(v) for v in r.stdout.split()[1:5]]
    T = oracle.pose_to_matrix(np.array([1.0, 2.0, 3.0, 0.1, -0.2, 0.3]))
    want = [T[0, 0], T[1, 0], T[2, 1], T[0, 3]]
    assert [np.float32(g) for g in got] == [np.float32(w) for w in want]


def test_convert_transform_host_entry_matches_oracle():
    """ndtb_pose_to_matrix (host only) against the oracle's restatement of the Eigen float chain, bit-exact over random poses."""
    from oracle import oracle
    from xchu_slam_b200 import ndt
    rng = np.random.default_rng(5)
    for _ in range(200):
        p = np.concatenate([rng.uniform(-50, 50, 3), rng.uniform(-3.2, 3.2, 3)])
        assert np.array_equal(ndt.NormalDistributionsTransform.convertTransform(p), oracle.pose_to_matrix(p))


@pytest.mark.gpu
def test_align_through_registration_base_pointer(exe):
    """The PCL build of the header driven through a pcl::Registration pointer: the base's setters and align() reach the device
    (virtual setInputTarget / setInputSource / computeTransformation), and calculateScore(cloud) == calculateScore(T)."""
    r = subprocess.run([_pcl_mode_exe()], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "base-pointer align" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("strategy", ["distance", "surround"])
def test_same_caller_two_backends_same_trajectory(exe, tmp_path, strategy):
    """Drop-in check at the caller level: examples/odom_sequence.cpp (the mirror of odom_node's loop) compiled twice --
    against include/pclomp/ndt_omp.h (libndt_b200.so, GPU) and against the reference's own pclomp headers
    (examples/odom_sequence_ref, CPU, over the stand-in Eigen/PCL headers) -- must produce the SAME odometry on the same
    sequence, for BOTH localmap refresh strategies north_star names: `distance` = OdomEstimate's distance-threshold refresh
    (odom_node.cpp:329-356, res 1.0), `surround` = the keyframe-count sliding window ExtractSurroundKeyframes
    (odom_node.cpp:737-802: newest 20 keyframes, a non-prefix target and a full table build every scan; at the reference's
    default ndt_res 2.0, the resolution at which one 1.0 m-downsampled keyframe can fill a voxel).
    A handle runs the term-by-term derivative kernel by default (strict_math -1), whose per-pair float terms
    are the reference's bit for bit; each pose feeds the next guess and the localmap, so the trajectories can only stay
    identical if every align is: iteration counts, convergence flags, localmap sizes and poses are required to be EQUAL,
    fitness scores within 1e-9 relative (exact 1-NN on both sides, f64 mean).
    The factorised kernel (--fast, 1e-6 per evaluation) is held to the north-star gate per align instead: it tracks the
    reference within 1e-4 m / 1e-5 rad until the first voxel-membership flip of the localmap feedback loop."""
    from oracle import ref_examples
    from xchu_slam_b200 import build, synth
    ref_exe = ref_examples.build_example_reference()
    if ref_exe is None:
        pytest.skip("examples/odom_sequence_ref not built (needs /root/reference)")
    seq = synth.write_sequence(str(tmp_path / "seq.bin"), n_scans=60, n_points=20000)
    common = ["--res", "1.0"] if strategy == "distance" else ["--res", "2.0", "--refresh", "surround"]
    b, sb = _run(ref_exe, seq, str(tmp_path / "ref.txt"), *common, "--threads", str(os.cpu_count() or 1))
    if strategy == "surround":
        assert int(sb["surround_sets"]) == 58      # every scan but the first and the `static latest_frame_id` call (:762-763)
        assert np.max(np.linalg.norm(b[:, 4:6] - b[:, 14:16], axis=1)) < 0.5
    for flags in ([], ["--device-localmap"]):
        a, sa = _run(exe, seq, str(tmp_path / "gpu.txt"), *common, *flags)
        assert a.shape == b.shape and a.shape[0] == 60
        assert sa["surround_sets"] == sb["surround_sets"]
        assert np.array_equal(a[:, 1:3], b[:, 1:3])                       # iterations, converged
        assert np.array_equal(a[:, 10], b[:, 10])                         # localmap sizes
        assert np.array_equal(a[:, 4:10], b[:, 4:10])                     # poses: identical
        assert np.max(np.abs(a[:, 3] - b[:, 3]) / b[:, 3]) < 1e-9
    if strategy == "surround":
        return
    f, sf = _run(exe, seq, str(tmp_path / "fast.txt"), "--res", "1.0", "--fast")
    same = np.flatnonzero(f[:, 10] != b[:, 10])
    upto = int(same[0]) if len(same) else 60                              # scans before the localmaps first differ
    assert upto >= 10
    assert np.array_equal(f[:upto, 1:3], b[:upto, 1:3])
    assert np.max(np.abs(f[:upto, 4:7] - b[:upto, 4:7])) < 1e-4 and np.max(np.abs(f[:upto, 7:10] - b[:upto, 7:10])) < 1e-5
    assert np.max(np.linalg.norm(f[:, 4:6] - b[:, 4:6], axis=1)) < 0.05   # and stays a good odometry afterwards


@pytest.mark.gpu
def test_device_localmap_mode_gives_the_same_odometry(exe, tmp_path):
    """--device-localmap keeps localmap / tmp_map / the pc_target_ snapshot in HBM (ndtb_localmap_*): same clouds bit for bit,
    hence the same voxel tables and the same trajectory as the host-side localmap, with or without incremental refresh."""
    from xchu_slam_b200 import synth
    seq = synth.write_sequence(str(tmp_path / "seq.bin"), n_scans=40, n_points=20000)
    a, sa = _run(exe, seq, str(tmp_path / "host.txt"), "--res", "1.0")
    b, sb = _run(exe, seq, str(tmp_path / "dev.txt"), "--res", "1.0", "--device-localmap")
    c, sc = _run(exe, seq, str(tmp_path / "devfull.txt"), "--res", "1.0", "--device-localmap", "--no-incremental")
    for x, sx in ((b, sb), (c, sc)):
        assert np.array_equal(a[:, 1:3], x[:, 1:3])          # iterations, converged
        assert np.array_equal(a[:, 3:11], x[:, 3:11])        # fitness, pose, target size
    assert int(sb["incremental"]) >= int(sb["target_sets"]) - 8 and int(sc["incremental"]) == 0


# ---- include/scancontext/Scancontext.h: the SCManager drop-in ------------------------------------
def _sc_exes():
    from oracle import ref_examples
    from xchu_slam_b200 import build
    build.build_cuda()
    return build.build_sc_example(), ref_examples.build_sc_example_reference()


def test_sc_shim_compiles_and_fails_loudly_without_gpu(tmp_path):
    import torch
    exe_dev, _ = _sc_exes()
    assert os.path.exists(exe_dev)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from xchu_slam_b200 import synth
    seq = synth.write_sequence(str(tmp_path / "s.bin"), n_scans=2, n_points=2000)
    r = subprocess.run([exe_dev, seq, str(tmp_path / "o.txt")], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr


def test_sc_reference_binary_agrees_with_the_oracle(tmp_path):
    """The reference-compiled caller (examples/sc_sequence_ref) against the oracle port driven from Python on the same file:
    same loop ids, same yaw bits -- what the device binary is held to on the GPU box."""
    _, exe_ref = _sc_exes()
    if exe_ref is None:
        pytest.skip("examples/sc_sequence_ref not built (needs /root/reference)")
    from oracle import sc
    from xchu_slam_b200 import synth
    seq = synth.write_sequence(str(tmp_path / "seq.bin"), n_scans=70, n_points=6000)
    out = str(tmp_path / "ref.txt")
    r = subprocess.run([exe_ref, seq, out, "--revisit", "3,10"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rows = [l.split() for l in open(out) if not l.startswith("#")]
    scans = synth.read_sequence(seq)[0]
    scans = scans + [scans[3], scans[10]]
    o = sc.OracleSC()
    assert len(rows) == len(scans)
    for k, s in enumerate(scans):
        o.add(s)
        lid, yaw = o.detect()
        assert int(rows[k][1]) == lid, k
        assert np.float32(float(rows[k][2])).tobytes() == np.float32(yaw).tobytes(), k
    assert [int(x[1]) for x in rows[-2:]] == [3, 10]


@pytest.mark.gpu
def test_sc_same_caller_two_backends_same_loops(tmp_path):
    """examples/sc_sequence.cpp (pgo_node's three SCManager calls) compiled over include/scancontext/Scancontext.h (device) and over
    the reference's own Scancontext.cpp (CPU): identical loop ids and identical yaw differences for every keyframe."""
    exe_dev, exe_ref = _sc_exes()
    if exe_ref is None:
        pytest.skip("examples/sc_sequence_ref not built (needs /root/reference)")
    from xchu_slam_b200 import synth
    seq = synth.write_sequence(str(tmp_path / "seq.bin"), n_scans=70, n_points=6000)
    outs = []
    for exe, name in ((exe_dev, "dev.txt"), (exe_ref, "ref.txt")):
        r = subprocess.run([exe, seq, str(tmp_path / name), "--revisit", "3,10,20"], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        outs.append([l for l in open(str(tmp_path / name)) if not l.startswith("#")])
    assert len(outs[0]) == 73
    assert outs[0] == outs[1]
    assert [int(l.split()[1]) for l in outs[0][-3:]] == [3, 10, 20]
