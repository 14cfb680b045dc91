"""Loop-closure refinement timing on one GPU: keyframe store resident in HBM, `n_pairs`
candidates per call, each a (+-25 keyframe submap, 0.5 m VoxelGrid) target and a one-keyframe source, registered by the
batched NDT and gated on getFitnessScore.  Beside it the oracle (CPU restatement: same clouds, NDT, exact 1-NN fitness)
on a bounded sample of the candidates -- a checker timed as a baseline, not a tuned one.  The reference's own arm for
this row is PCL's ICP, which is not in /root/reference: unavailable.  Prints one JSON object."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from xchu_slam_b200 import loop, synth
    n_kf = int(sys.argv[1]) if len(sys.argv) > 1 else 160
    n_pairs = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    cpu_sample = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    w = synth.World(synth.BASE_SEED)
    poses = w.trajectory(60 + 2 * n_kf, seed=synth.BASE_SEED & 0xFFFF)
    rng = np.random.default_rng(5)
    kfs, pose6 = [], []
    for k in range(n_kf):
        vp = poses[50 + 2 * k]
        kfs.append(synth.source_scan(w, vp, synth.BASE_SEED + 900 + k, 8000))
        pose6.append(synth.sensor_pose(vp) + np.array([rng.normal(0, 0.05), rng.normal(0, 0.05), rng.normal(0, 0.01), rng.normal(0, 0.002),
                                                       rng.normal(0, 0.002), rng.normal(0, 0.004)]))
    params = dict(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30)
    out = {"keyframes": n_kf, "points_per_keyframe": 8000, "pairs_per_call": n_pairs, "half_window": 25, "filter_size": 0.5}
    pairs = [(int(rng.integers(25, n_kf - 25)), 0) for _ in range(n_pairs)]
    pairs = [(lp, min(n_kf - 1, lp + int(rng.integers(1, 4)))) for lp, _ in pairs]
    for mode in ("reference", "current"):
        r = loop.LoopRefiner(0, source_mode=mode, **params)
        t0 = time.perf_counter()
        for c, p in zip(kfs, pose6):
            r.add_keyframe(c, p)
        t_store = time.perf_counter() - t0
        r.refine(pairs[:4])
        ts, stages = [], []
        for _ in range(5):
            tm = {}
            t0 = time.perf_counter()
            res = r.refine(pairs, timing=tm)
            ts.append(time.perf_counter() - t0)
            stages.append(tm)
        ms = float(np.median(ts) * 1e3)
        out[mode] = {"gpu_ms_per_call": ms, "gpu_candidates_per_s": n_pairs / (ms * 1e-3), "keyframe_store_upload_s": t_store,
                     "stage_ms": {k: float(np.median([st[k] for st in stages]) * 1e3) for k in stages[0]},
                     "distinct_targets": len({p[0] for p in pairs}),
                     "accepted": int(sum(x["accepted"] for x in res)), "mean_iterations": float(np.mean([x["iterations"] for x in res])),
                     "mean_target_points": float(np.mean([x["n_target"] for x in res])),
                     "mean_source_points": float(np.mean([x["n_source"] for x in res]))}
        # CPU oracle on a bounded sample of the same candidates
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
        from test_loop import _oracle_refine
        t0 = time.perf_counter()
        want = _oracle_refine(kfs, pose6, pairs[:cpu_sample], 0.5, 25, 0.3, mode, params)
        cpu_s = (time.perf_counter() - t0) / cpu_sample
        agree = all(g["accepted"] == o["accepted"] and g["iterations"] == o["iterations"] and np.max(np.abs(g["T"] - o["T"])) < 1e-4
                    for g, o in zip(res[:cpu_sample], want))
        out[mode].update(cpu_oracle_s_per_candidate=cpu_s, cpu_sample=cpu_sample, sample_agrees_with_oracle=bool(agree),
                         cpu_note="oracle restatement (assembly + NDT on all host cores + exact 1-NN fitness); PCL ICP, the reference's "
                                  "own arm for this row, is not in /root/reference")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
