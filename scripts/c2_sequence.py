#!/usr/bin/env python
"""Config C2 (BASELINE.json configs[1]): KITTI-00-shaped synthetic odometry sequence, sequential scan-to-localmap NDT
with distance-threshold localmap refresh, 1 GPU -- scans/s of examples/odom_sequence (drop-in header over
libndt_b200.so; incremental voxel rebuild vs full rebuilds) beside examples/odom_sequence_ref (the SAME caller over the
reference's own pclomp sources on the host cores; a bounded prefix of the sequence).  Prints one JSON object."""
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_examples  # noqa: E402
from xchu_slam_b200 import build, synth  # noqa: E402


def run(exe, seq, out, *flags):
    t0 = time.perf_counter()
    r = subprocess.run([exe, seq, out] + list(flags), capture_output=True, text=True)
    wall = time.perf_counter() - t0
    assert r.returncode == 0, r.stderr
    rows = np.array([l.split() for l in open(out) if not l.startswith("#")], dtype=np.float64)
    summ = [l for l in open(out) if l.startswith("#")][0].split()
    return rows, dict(zip(summ[1::2], summ[2::2])), wall


def main():
    n_scans = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    n_cpu = int(sys.argv[2]) if len(sys.argv) > 2 else 100
    res = sys.argv[3] if len(sys.argv) > 3 else "1.0"
    tmp = "/tmp/ndt_c2"
    os.makedirs(tmp, exist_ok=True)
    build.build_cuda()
    exe = build.build_examples()
    ref_exe = ref_examples.build_example_reference()
    seq = synth.write_sequence(os.path.join(tmp, "seq%d.bin" % n_scans), n_scans=n_scans, n_points=20000)
    out = {"config": "C2: %d-scan synthetic sequence (20k-pt scans, ~0.86 m/scan), NDT res %s, DIRECT7, eps 0.01, <= 30 iterations, "
                     "localmap window 5 m, keyframe every >= 0.5 m" % (n_scans, res)}
    run(exe, seq, os.path.join(tmp, "warm.txt"), "--res", res, "--max-scans", "20")
    for name, flags in (("gpu_device_localmap", ["--device-localmap"]), ("gpu_device_localmap_full_rebuild", ["--device-localmap", "--no-incremental"]),
                        ("gpu_incremental", []), ("gpu_full_rebuild", ["--no-incremental"]), ("gpu_device_localmap_fast_kernel", ["--device-localmap", "--fast"])):
        rows, s, wall = run(exe, seq, os.path.join(tmp, name + ".txt"), "--res", res, *flags)
        err = np.linalg.norm(rows[:, 4:6] - rows[:, 14:16], axis=1)
        out[name] = {"scans_per_s": float(s["scans_per_s"]), "align_ms_mean": float(s["align_ms_mean"]), "target_ms_mean": float(s["target_ms_mean"]),
                     "fitness_ms_mean": float(s.get("fitness_ms_mean", 0)), "source_ms_mean": float(s.get("source_ms_mean", 0)),
                     "target_sets": int(s["target_sets"]), "incremental": int(s["incremental"]), "converged_frac": float(np.mean(rows[:, 2])),
                     "mean_iterations": float(np.mean(rows[:, 1])), "xy_drift_max_m": float(err.max()), "xy_drift_end_m": float(err[-1])}
    if ref_exe:
        nthreads = len(os.sched_getaffinity(0))
        for name, th in (("cpu_reference_all_cores", nthreads), ("cpu_reference_1_thread_as_shipped", 1)):
            rows, s, wall = run(ref_exe, seq, os.path.join(tmp, name + ".txt"), "--res", res, "--max-scans", str(n_cpu), "--threads", str(th))
            out[name] = {"scans_per_s": float(s["scans_per_s"]), "align_ms_mean": float(s["align_ms_mean"]), "target_ms_mean": float(s["target_ms_mean"]),
                         "scans": int(rows.shape[0]), "threads": th, "mean_iterations": float(np.mean(rows[:, 1])),
                         "kind": "reference sources over stand-in Eigen/PCL headers (oracle/ref_stubs)"}
        c = np.array([l.split() for l in open(os.path.join(tmp, "cpu_reference_all_cores.txt")) if not l.startswith("#")], dtype=np.float64)
        for name in ("gpu_device_localmap", "gpu_device_localmap_fast_kernel"):
            g = np.array([l.split() for l in open(os.path.join(tmp, name + ".txt")) if not l.startswith("#")][:n_cpu], dtype=np.float64)
            neq = np.flatnonzero(g[:, 1] != c[:, 1])
            out["trajectory_parity_first_%d_scans_%s" % (n_cpu, name)] = {
                "scans_with_equal_iterations": int(n_cpu - len(neq)), "first_scan_with_different_iterations": int(neq[0]) if len(neq) else None,
                "max_abs_pose_diff_m": float(np.max(np.abs(g[:, 4:7] - c[:, 4:7]))),
                "max_abs_angle_diff_rad": float(np.max(np.abs(g[:, 7:10] - c[:, 7:10]))),
                "max_abs_pose_diff_m_before_first_difference": float(np.max(np.abs(g[:neq[0], 4:7] - c[:neq[0], 4:7]))) if len(neq) and neq[0] > 0 else None}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
