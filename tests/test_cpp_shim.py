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
