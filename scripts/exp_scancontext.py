"""ScanContext on the device timed on one GPU: descriptors of a batch of scans, column-shift
distances of a batch of pairs, loop detection for a batch of queries against a ring-key set; beside them the oracle port and,
where it was built, the reference's own Scancontext.cpp (oracle/_ref/libsc_ref.so over stand-in Eigen headers) on one host
core -- the reference runs this path single-threaded (pgo_node.cpp:236, 344).  Prints one JSON object."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from oracle import sc
    from xchu_slam_b200 import scancontext, synth
    n_base = int(sys.argv[1]) if len(sys.argv) > 1 else 48
    n_scans = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
    w = synth.World(synth.BASE_SEED)
    poses = w.trajectory(60 + 3 * n_base, seed=synth.BASE_SEED & 0xFFFF)
    base = [synth.voxel_downsample(w.scan(poses[50 + 3 * k], 4000 + k), 0.5) for k in range(n_base)]
    rng = np.random.default_rng(7)
    scans = []
    for i in range(n_scans):                      # yaw-rotated, jittered copies: distinct descriptors, realistic occupancy
        s = base[i % n_base].copy()
        a = rng.uniform(0, 2 * np.pi)
        c, s_ = np.float32(np.cos(a)), np.float32(np.sin(a))
        x, y = s[:, 0].copy(), s[:, 1].copy()
        s[:, 0], s[:, 1] = c * x - s_ * y, s_ * x + c * y
        s[:, :3] += rng.normal(0, 0.05, (s.shape[0], 3)).astype(np.float32)
        scans.append(s)
    pts = float(np.mean([s.shape[0] for s in scans]))
    out = {"scans": n_scans, "mean_points_per_scan": pts}

    g = scancontext.SCManager(0)
    g.add_scans(scans[:8]); g.clear()
    t0 = time.perf_counter(); g.add_scans(scans); t_first = time.perf_counter() - t0      # includes growing the store and the buffers
    g.clear()
    t0 = time.perf_counter(); g.add_scans(scans); t_make = time.perf_counter() - t0
    n_pairs = 65536
    i1, i2 = rng.integers(0, n_scans, n_pairs), rng.integers(0, n_scans, n_pairs)
    g.distance(i1[:64], i2[:64])
    t0 = time.perf_counter(); dist, shift = g.distance(i1, i2); t_dist = time.perf_counter() - t0
    q = np.arange(n_scans // 2, n_scans)
    ts = q - 30
    g.detect(q[:8], ts[:8])
    t0 = time.perf_counter(); det = g.detect(q, ts); t_det = time.perf_counter() - t0
    out["gpu"] = {"descriptors_per_s": n_scans / t_make, "make_ms_total": t_make * 1e3, "make_ms_first_call": t_first * 1e3, "pair_distances_per_s": n_pairs / t_dist,
                  "distance_ms_total": t_dist * 1e3, "detect_queries_per_s": len(q) / t_det, "detect_ms_total": t_det * 1e3,
                  "note": "host wall time of one C-ABI call each, host buffers in, results out (uploads and downloads included)"}

    def cpu(m, label):
        n1 = min(n_scans, 256)
        t0 = time.perf_counter()
        D = [m.make(s) for s in scans[:n1]]
        t_make = (time.perf_counter() - t0) / n1
        n2 = 512
        t0 = time.perf_counter()
        ok = True
        for k in range(n2):
            a, b = int(i1[k]) % n1, int(i2[k]) % n1
            d, s = m.distance(D[a], D[b])
        t_d = (time.perf_counter() - t0) / n2
        return {"descriptors_per_s": 1 / t_make, "pair_distances_per_s": 1 / t_d, "cores": 1, "sample": "%d descriptors, %d pairs" % (n1, n2),
                "kind": label}

    o = sc.OracleSC()
    out["cpu_port"] = cpu(o, "port")
    if sc.ref_available():
        out["cpu_reference"] = cpu(sc.RefSC(), "reference sources over stand-in Eigen (oracle/ref_stubs)")
    # spot parity on the timed data
    bad = 0
    Dg = {}
    for k in range(200):
        a, b = int(i1[k]), int(i2[k])
        for x in (a, b):
            if x not in Dg:
                Dg[x] = o.make(scans[x])
        d, s = o.distance(Dg[a], Dg[b])
        bad += int(not (d == dist[k] and s == shift[k]))
    out["parity_spot_check"] = {"pairs": 200, "mismatches": bad}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
