"""Deterministic synthetic inputs for the NDT hot path.

Input tooling for tests and bench.py: procedural urban scene, HDL-64E-shaped ray-caster,
filter_node-style gating/downsample (reference src/filter_node.cpp:242-251), localmap
accumulation in the way odom_node builds it (reference src/odom_node.cpp:329-350: every
keyframe is 1.0 m voxel-downsampled, transformed by its pose and appended).  Host-side C++
(lidar_synth.cpp) bound through ctypes; it is neither the product path nor the oracle.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(os.path.dirname(_HERE), "_lidar_synth.so")
BASE_SEED = 20261104
SENSOR_H = 1.73


def build(force=False):
    src = os.path.join(_HERE, "lidar_synth.cpp")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fopenmp", "-fPIC", "-shared", "-o", _LIB, src])
    return _LIB


_lib = None


def _L():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        L.synth_world_create.restype = C.c_void_p
        L.synth_world_create.argtypes = [C.c_uint64, C.c_int]
        L.synth_world_destroy.argtypes = [C.c_void_p]
        L.synth_world_num_prims.argtypes = [C.c_void_p]
        L.synth_trajectory.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_uint64, C.c_void_p]
        L.synth_scan.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
        L.synth_scan.restype = C.c_int
        L.synth_range_gate.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float]
        L.synth_range_gate.restype = C.c_int
        L.synth_voxel_downsample.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_int]
        L.synth_voxel_downsample.restype = C.c_int
        L.synth_transform.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def pose6_to_matrix(p):
    """x, y, z, roll, pitch, yaw -> 4x4 float64 (R = Rz Ry Rx, as common.h:64-71 Pose6D2Matrix)."""
    x, y, z, r, pt, yw = [float(v) for v in p]
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(pt), np.sin(pt), np.cos(yw), np.sin(yw)
    T = np.eye(4)
    T[:3, :3] = [[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                 [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                 [-sp, cp * sr, cp * cr]]
    T[:3, 3] = [x, y, z]
    return T


class World:
    def __init__(self, seed=BASE_SEED, half_blocks=20):
        self.L = _L()
        self.seed = seed
        self.h = C.c_void_p(self.L.synth_world_create(seed, half_blocks))

    def __del__(self):
        try:
            self.L.synth_world_destroy(self.h)
        except Exception:
            pass

    def trajectory(self, n, step=0.86, seed=0):
        """n sensor poses (x, y, z=SENSOR_H... as vehicle pose6 with z=0), KITTI-00-like 0.86 m/scan."""
        out = np.zeros((n, 3))
        self.L.synth_trajectory(self.h, n, step, seed, _p(out))
        poses = np.zeros((n, 6))
        poses[:, 0], poses[:, 1], poses[:, 5] = out[:, 0], out[:, 1], out[:, 2]
        return poses

    def scan(self, vehicle_pose6, scan_seed):
        """Raw revolution in the sensor frame: (n, 4) float32, ~100-130k returns."""
        pose = np.ascontiguousarray(vehicle_pose6, dtype=np.float64)
        buf = np.zeros((64 * 2083, 4), np.float32)
        n = self.L.synth_scan(self.h, _p(pose), int(scan_seed), _p(buf), buf.shape[0])
        return buf[:n].copy()


def sensor_pose(vehicle_pose6):
    """World pose of the sensor frame (what NDT estimates) for a vehicle pose with roll=pitch=0."""
    p = np.array(vehicle_pose6, dtype=np.float64).copy()
    p[2] += SENSOR_H
    return p


def range_gate(pts, lo=1.0, hi=60.0):
    pts = np.ascontiguousarray(pts, dtype=np.float32).copy()
    n = _L().synth_range_gate(_p(pts), pts.shape[0], lo, hi)
    return pts[:n].copy()


def voxel_downsample(pts, leaf):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    out = np.zeros_like(pts)
    n = _L().synth_voxel_downsample(_p(pts), pts.shape[0], leaf, _p(out), out.shape[0])
    return out[:n].copy()


def transform(pts, pose6):
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    out = np.zeros_like(pts)
    pose = np.ascontiguousarray(pose6, dtype=np.float64)
    _L().synth_transform(_p(pts), pts.shape[0], _p(pose), _p(out))
    return out


def exact_count(pts, n, seed):
    """Seeded subsample / pad (by repeating seeded picks) to exactly n points, order preserved."""
    rng = np.random.default_rng(seed)
    m = pts.shape[0]
    if m == n:
        return pts
    if m > n:
        idx = np.sort(rng.choice(m, n, replace=False))
        return np.ascontiguousarray(pts[idx])
    extra = rng.choice(m, n - m, replace=True)
    return np.ascontiguousarray(np.concatenate([pts, pts[extra]], axis=0))


def source_scan(world, vehicle_pose6, scan_seed, n_points=20000, dense=False):
    """filter_node front end: gate 1 < r_xy < 60 m, 0.5 m voxel downsample, exactly n_points.
    dense=True keeps the undownsampled revolution (config C4, 120k points)."""
    raw = world.scan(vehicle_pose6, scan_seed)
    if not dense:
        gated = range_gate(raw, 1.0, 60.0)
        # 0.5 m leaf as filter_node; the procedural scene is sparser than KITTI, so fall back to
        # a finer leaf until the downsampled scan holds n_points (then seeded-subsample to exactly n)
        for leaf in (0.5, 0.4, 0.3, 0.25, 0.2, 0.15, 0.1):
            raw = voxel_downsample(gated, leaf)
            if raw.shape[0] >= n_points:
                break
    return exact_count(raw, n_points, scan_seed ^ 0x5EED)


def keyframe(world, vehicle_pose6, scan_seed, keyframe_leaf=1.0):
    """One localmap keyframe (ODOM:329-336): 1.0 m voxel-downsampled scan moved to the world frame."""
    raw = world.scan(vehicle_pose6, scan_seed)
    ds = voxel_downsample(range_gate(raw, 1.0, 120.0), keyframe_leaf)
    return transform(ds, sensor_pose(vehicle_pose6))


def localmap(world, vehicle_poses, seed0, n_points=None, keyframe_leaf=1.0):
    """Accumulate GT-posed keyframes the way odom_node does (ODOM:329-350): each scan is
    voxel-downsampled at keyframe_leaf in the sensor frame, transformed by its pose, appended.
    Stops as soon as n_points is reached (then seeded-subsampled to exactly n_points)."""
    chunks, total = [], 0
    for i, vp in enumerate(vehicle_poses):
        chunks.append(keyframe(world, vp, seed0 + i, keyframe_leaf))
        total += chunks[-1].shape[0]
        if n_points is not None and total >= n_points:
            break
    pts = np.concatenate(chunks, axis=0)
    if n_points is not None:
        pts = exact_count(pts, n_points, seed0 ^ 0xA11)
    return np.ascontiguousarray(pts)


def perturbed_guess(sensor_pose6, seed):
    """GT o delta: +-0.5 m x,y; +-0.1 m z; +-0.5 deg roll,pitch; +-2 deg yaw.  Returns float32 4x4."""
    rng = np.random.default_rng(seed)
    d = np.array([rng.uniform(-0.5, 0.5), rng.uniform(-0.5, 0.5), rng.uniform(-0.1, 0.1),
                  np.deg2rad(rng.uniform(-0.5, 0.5)), np.deg2rad(rng.uniform(-0.5, 0.5)), np.deg2rad(rng.uniform(-2, 2))])
    T = pose6_to_matrix(sensor_pose6) @ pose6_to_matrix(d)
    return T.astype(np.float32)


def make_case(n_source=20000, n_target=300000, n_keyframes=None, spacing=2, traj_start=50, seed=BASE_SEED,
              world=None, dense=False, keyframe_leaf=0.5):
    """One registration problem: (source (n,4) f32 sensor frame, target (m,4) f32 world frame,
    gt sensor pose6, guess 4x4 f32).  The target is a localmap of GT-posed keyframes every
    `spacing` scans, each voxel-downsampled at keyframe_leaf, accumulated to n_target points."""
    w = world or World(seed)
    if n_keyframes is None:
        n_keyframes = max(8, int(n_target / (9000 if keyframe_leaf <= 0.5 else 4000)) + 4)
    poses = w.trajectory(traj_start + n_keyframes * spacing + 10, seed=seed & 0xFFFF)
    kf = poses[traj_start:traj_start + n_keyframes * spacing:spacing]
    target = localmap(w, kf, seed + 1000, n_target, keyframe_leaf)
    q = poses[traj_start + (n_keyframes * spacing) // 3 + 1]
    src = source_scan(w, q, seed + 7, n_source, dense=dense)
    gt = sensor_pose(q)
    return src, target, gt, perturbed_guess(gt, seed + 13)


def batch_workload(n_submaps=64, scans_per_submap=8, hyps_per_scan=8, n_target=1000000, n_source=20000,
                   seed=BASE_SEED, keyframe_leaf=0.5, kf_every=2, submap_stride=4, cache_dir=None, verbose=False):
    """BASELINE configs[4] (C5): n_submaps distinct localmaps (consecutive GT-posed keyframes along one
    drive, windows shifted by submap_stride keyframes), scans_per_submap scans taken inside each
    submap, hyps_per_scan perturbed initial guesses per scan.
    Returns dict(targets=[(n_target,4) f32], sources=[(n_source,4) f32], item_target, item_source,
    guesses (n_items,4,4) f32, gt (n_items,6)).  Cached as .npz under cache_dir when given."""
    tag = "c5_%d_%d_%d_%d_%d_%d_%s_%d_%d" % (n_submaps, scans_per_submap, hyps_per_scan, n_target, n_source, seed,
                                             keyframe_leaf, kf_every, submap_stride)
    path = os.path.join(cache_dir, tag + ".npz") if cache_dir else None
    if path and os.path.exists(path):
        z = np.load(path)
        return dict(targets=list(z["targets"]), sources=list(z["sources"]), item_target=z["item_target"],
                    item_source=z["item_source"], guesses=z["guesses"], gt=z["gt"])
    w = World(seed)
    est_kf = int(n_target / (9000 if keyframe_leaf <= 0.5 else 4000)) + 40
    n_kf = n_submaps * submap_stride + est_kf * 2
    poses = w.trajectory(60 + n_kf * kf_every + 16, seed=seed & 0xFFFF)
    kf_cache = {}

    def kf(k):
        if k not in kf_cache:
            kf_cache[k] = keyframe(w, poses[50 + k * kf_every], seed + 1000 + k, keyframe_leaf)
        return kf_cache[k]

    targets, sources, it, isrc, guesses, gts = [], [], [], [], [], []
    for s in range(n_submaps):
        k0, total, chunks = s * submap_stride, 0, []
        k = k0
        while total < n_target:
            chunks.append(kf(k))
            total += chunks[-1].shape[0]
            k += 1
        targets.append(exact_count(np.concatenate(chunks, axis=0), n_target, seed + 77 + s))
        for dead in [d for d in kf_cache if d < k0 + submap_stride]:
            if s + 1 < n_submaps and dead < (s + 1) * submap_stride:
                del kf_cache[dead]
        nk = k - k0
        for j in range(scans_per_submap):
            # a scan pose strictly between two keyframe poses, spread over the middle half of the submap
            kq = k0 + nk // 4 + (j * max(1, (nk // 2) // scans_per_submap))
            pi = 50 + kq * kf_every + 1
            src = source_scan(w, poses[pi], seed + 500000 + s * 97 + j, n_source)
            sid = len(sources)
            sources.append(src)
            gt = sensor_pose(poses[pi])
            for hh in range(hyps_per_scan):
                it.append(s)
                isrc.append(sid)
                guesses.append(perturbed_guess(gt, seed + 900000 + len(it)))
                gts.append(gt)
        if verbose:
            print("[synth] submap %d/%d: %d keyframes" % (s + 1, n_submaps, nk), flush=True)
    out = dict(targets=targets, sources=sources, item_target=np.array(it, np.int32), item_source=np.array(isrc, np.int32),
               guesses=np.stack(guesses).astype(np.float32), gt=np.stack(gts))
    if path:
        os.makedirs(cache_dir, exist_ok=True)
        tmp = path + ".tmp%d.npz" % os.getpid()
        np.savez(tmp, targets=np.stack(targets), sources=np.stack(sources), item_target=out["item_target"],
                 item_source=out["item_source"], guesses=out["guesses"], gt=out["gt"])
        os.replace(tmp, path)
    return out


def write_sequence(path, n_scans=1000, n_points=20000, seed=BASE_SEED, traj_start=50, step=0.86):
    """BASELINE configs[1] (C2) input: n_scans consecutive filter_node-style scans along the KITTI-00-like
    drive with their ground-truth sensor poses, in the flat format examples/odom_sequence.cpp reads:
    "NDTSEQ1\0", int32 n_scans, then per scan int32 n, float64 gt[6], float32 xyzi[n][4]."""
    w = World(seed)
    poses = w.trajectory(traj_start + n_scans + 2, step=step, seed=seed & 0xFFFF)
    tmp = path + ".tmp%d" % os.getpid()
    with open(tmp, "wb") as f:
        f.write(b"NDTSEQ1\0")
        f.write(np.int32(n_scans).tobytes())
        for i in range(n_scans):
            vp = poses[traj_start + i]
            pts = source_scan(w, vp, seed + 300000 + i, n_points)
            f.write(np.int32(pts.shape[0]).tobytes())
            f.write(np.asarray(sensor_pose(vp), dtype=np.float64).tobytes())
            f.write(np.ascontiguousarray(pts, dtype=np.float32).tobytes())
    os.replace(tmp, path)
    return path


def read_sequence(path):
    """Inverse of write_sequence: ([(n, 4) float32 scan, ...], (n_scans, 6) float64 ground-truth poses)."""
    with open(path, "rb") as f:
        if f.read(8) != b"NDTSEQ1\0":
            raise ValueError("not an NDTSEQ1 file: " + path)
        n_scans = int(np.frombuffer(f.read(4), np.int32)[0])
        scans, gts = [], []
        for _ in range(n_scans):
            n = int(np.frombuffer(f.read(4), np.int32)[0])
            gts.append(np.frombuffer(f.read(48), np.float64).copy())
            scans.append(np.frombuffer(f.read(16 * n), np.float32).reshape(n, 4).copy())
    return scans, np.stack(gts) if gts else np.zeros((0, 6))
