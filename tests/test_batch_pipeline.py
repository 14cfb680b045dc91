"""examples/batch_pipeline.cpp: the batch entry points driven from C++ over the C ABI (page-locked clouds, two batches in flight).
CPU: it builds and fails loudly without a device.  GPU: its results equal the Python mirror's on the same job, for 16-byte and
packed 12-byte uploads."""
import json
import os
import subprocess

import numpy as np
import pytest


@pytest.fixture(scope="module")
def exe():
    from xchu_slam_b200 import build
    build.build_cuda()
    return build.build_batch_pipeline()


def _job(tmp_path, stride, n_sub=2):
    from xchu_slam_b200 import jobfile, synth
    w = synth.batch_workload(n_submaps=n_sub, scans_per_submap=2, hyps_per_scan=3, n_target=60000, n_source=4000)
    path = str(tmp_path / ("job%d.bin" % stride))
    jobfile.write_job(path, w["targets"], w["sources"], w["item_target"], w["item_source"], w["guesses"], stride=stride)
    return w, path


def test_job_file_layout(tmp_path):
    from xchu_slam_b200 import jobfile
    t = [np.arange(12, dtype=np.float32).reshape(3, 4)]
    s = [np.ones((2, 4), np.float32), np.zeros((1, 4), np.float32)]
    p = jobfile.write_job(str(tmp_path / "j.bin"), t, s, [0, 0], [0, 1], np.tile(np.eye(4, dtype=np.float32).ravel(), (2, 1)), stride=12)
    raw = open(p, "rb").read()
    assert raw[:8] == b"NDTBJOB1"
    assert np.frombuffer(raw[8:24], np.int32).tolist() == [1, 2, 2, 12]
    assert len(raw) == 24 + 8 * 3 + 4 * 4 + 2 * 64 + (3 + 2 + 1) * 12
    assert np.array_equal(np.frombuffer(raw[-72:-36], np.float32).reshape(3, 3), t[0][:, :3])


def test_pipeline_fails_loudly_without_gpu(exe, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    _, path = _job(tmp_path, 16, n_sub=1)
    r = subprocess.run([exe, path], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("stride", [16, 12])
def test_pipeline_results_equal_python_mirror(exe, tmp_path, stride):
    from xchu_slam_b200 import jobfile, ndt
    w, path = _job(tmp_path, stride)
    out = str(tmp_path / "res.bin")
    r = subprocess.run([exe, path, "--steps", "2", "--warmup", "1", "--out", out], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["items"] == len(w["item_target"]) and line["stride_bytes"] == stride
    got = jobfile.read_results(out)
    b = ndt.BatchNDT()
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2)
    for t in w["targets"]:
        b.add_target(t)
    for s in w["sources"]:
        b.add_source(s)
    ref = b.run_raw(w["item_target"], w["item_source"], w["guesses"])
    assert np.array_equal(got[:, 16].astype(np.int32), ref["iterations"])
    assert np.array_equal(got[:, 17].astype(np.int32), ref["converged"])
    assert np.array_equal(got[:, :16].astype(np.float32), ref["T"].reshape(-1, 16))     # same library, same inputs: bit-identical
