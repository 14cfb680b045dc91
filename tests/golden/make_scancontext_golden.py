"""Generates tests/golden/scancontext_small.npz from the REFERENCE build (oracle/_ref/libsc_ref.so = the reference's own
Scancontext.cpp over stand-in Eigen / PCL headers): six small seeded scans, their descriptors and keys, and the column-shift
distance of every ordered pair.  Run in the container that has /root/reference:  python tests/golden/make_scancontext_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    from oracle import sc
    from xchu_slam_b200 import synth
    r = sc.RefSC()
    w = synth.World(synth.BASE_SEED)
    poses = w.trajectory(120, seed=synth.BASE_SEED & 0xFFFF)
    out = {"n_scans": 6}
    desc, ring, sector = [], [], []
    for i in range(6):
        s = synth.voxel_downsample(w.scan(poses[50 + 9 * i], 7000 + i), 1.5)[:, :3].copy()      # ~2-3k points each
        out["scan_%d" % i] = s.astype(np.float32)
        d = r.make(s)
        rk, sk = r.keys(d)
        desc.append(d); ring.append(rk); sector.append(sk)
    pairs, dist, shift = [], [], []
    for i in range(6):
        for j in range(6):
            d, sh = r.distance(desc[i], desc[j])
            pairs.append((i, j)); dist.append(d); shift.append(sh)
    out.update(desc=np.stack(desc), ring=np.stack(ring), sector=np.stack(sector), pairs=np.array(pairs, np.int32), dist=np.array(dist),
               shift=np.array(shift, np.int32))
    path = os.path.join(ROOT, "tests", "golden", "scancontext_small.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
