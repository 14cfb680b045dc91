"""ctypes binding for the CPU oracle (oracle/ndt_oracle.cpp).

TEST INFRASTRUCTURE ONLY -- imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libndt_oracle.so")

KDTREE, DIRECT26, DIRECT7, DIRECT1 = 0, 1, 2, 3


def build(force=False):
    src = os.path.join(_HERE, "ndt_oracle.cpp")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


class _Result(C.Structure):
    _fields_ = [("T", C.c_float * 16), ("iterations", C.c_int), ("converged", C.c_int),
                ("score", C.c_double), ("trans_probability", C.c_double), ("n_evals", C.c_int)]


_lib = None
_lib_native = None
_NATIVE = os.path.join(_HERE, "_native", "libndt_oracle_native.so")
NATIVE_FLAGS = "-O3 -march=native -std=c++17 -fopenmp -ffp-contract=off -fPIC"


def build_native():
    """The oracle compiled ON THIS MACHINE with BASELINE.md section 2's flags (-O3 -march=native -fopenmp -ffp-contract=off) for the
    timed CPU baseline: a -march=native object built in the build container could not be trusted to run on the GPU box's host CPU,
    so this one is never shipped.  Same source, same IEEE arithmetic (no contraction, no fast-math): same results.  Returns the
    path, or None when the compiler is missing or fails."""
    src = os.path.join(_HERE, "ndt_oracle.cpp")
    try:
        if not os.path.exists(_NATIVE) or os.path.getmtime(_NATIVE) < os.path.getmtime(src):
            os.makedirs(os.path.dirname(_NATIVE), exist_ok=True)
            tmp = _NATIVE + ".tmp%d" % os.getpid()
            subprocess.check_call(["g++"] + NATIVE_FLAGS.split() + ["-w", "-shared", "-o", tmp, src], cwd=_HERE)
            os.replace(tmp, _NATIVE)
        return _NATIVE
    except Exception:
        return None


def lib(native=False):
    global _lib, _lib_native
    if native:
        if _lib_native is None:
            path = build_native()
            if path is None:
                return lib(False)
            _lib_native = _bind(C.CDLL(path))
        return _lib_native
    if _lib is None:
        build()
        _lib = _bind(C.CDLL(_LIB))
    return _lib


def _bind(L):
    L.oracle_create.restype = C.c_void_p
    L.oracle_destroy.argtypes = [C.c_void_p]
    L.oracle_set_params.argtypes = [C.c_void_p, C.c_float, C.c_double, C.c_double, C.c_int, C.c_double, C.c_int, C.c_int]
    L.oracle_set_voxel_params.argtypes = [C.c_void_p, C.c_int, C.c_double]
    for name in ("oracle_set_target", "oracle_update_target", "oracle_set_source"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t]
    L.oracle_grid_valid.argtypes = [C.c_void_p]
    L.oracle_grid_bounds.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_num_leaves.argtypes = [C.c_void_p]
    L.oracle_num_leaves.restype = C.c_size_t
    L.oracle_dump_leaves.argtypes = [C.c_void_p] + [C.c_void_p] * 8
    L.oracle_neighbors.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]
    L.oracle_gauss.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_eval_derivatives.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_eval_derivatives.restype = C.c_double
    L.oracle_compute_hessian.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_transform_source.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_align.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(_Result)]
    L.oracle_get_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    L.oracle_get_trace.restype = C.c_int
    L.oracle_get_output.argtypes = [C.c_void_p, C.c_void_p]
    L.oracle_fitness.argtypes = [C.c_void_p, C.c_double]
    L.oracle_fitness.restype = C.c_double
    L.oracle_build_nn.argtypes = [C.c_void_p]
    L.oracle_set_final_transform.argtypes = [C.c_void_p, C.c_void_p]
    L.oracle_calculate_score.argtypes = [C.c_void_p, C.c_void_p]
    L.oracle_calculate_score.restype = C.c_double
    L.oracle_pose_to_matrix.argtypes = [C.c_void_p, C.c_void_p]
    L.oracle_matrix_to_pose.argtypes = [C.c_void_p, C.c_void_p]
    L.oracle_svd_solve6.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.oracle_mt_trial.argtypes = [C.c_double] * 9
    L.oracle_mt_trial.restype = C.c_double
    L.oracle_mt_update.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double]
    L.oracle_max_threads.restype = C.c_int
    L.oracle_voxelgrid_downsample.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_float, C.c_void_p]
    L.oracle_voxelgrid_downsample.restype = C.c_size_t
    L.oracle_filter_scan.argtypes = [C.c_void_p, C.c_size_t, C.c_double, C.c_double, C.c_float, C.c_int, C.c_double, C.c_void_p,
                                     C.c_void_p, C.c_void_p]
    L.oracle_filter_scan.restype = C.c_size_t
    return L


def _pts(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    assert a.ndim == 2 and a.shape[1] >= 3
    return a


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def pose_to_matrix(p):
    """NDT_H:210-229 convertTransform: returns the 4x4 float32 matrix (row/col indexable)."""
    p = np.ascontiguousarray(p, dtype=np.float64)
    T = np.zeros(16, dtype=np.float32)
    lib().oracle_pose_to_matrix(_p(p), _p(T))
    return T.reshape(4, 4).T.copy()


def matrix_to_pose(T):
    Tc = np.ascontiguousarray(np.asarray(T, dtype=np.float32).T).reshape(-1)
    p = np.zeros(6, dtype=np.float64)
    lib().oracle_matrix_to_pose(_p(Tc), _p(p))
    return p


def svd_solve6(H, b):
    H = np.ascontiguousarray(H, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.zeros(6)
    lib().oracle_svd_solve6(_p(H), _p(b), _p(x))
    return x


def mt_trial(*args):
    return lib().oracle_mt_trial(*[float(a) for a in args])


def mt_update(state, a_t, f_t, g_t):
    s = np.ascontiguousarray(state, dtype=np.float64).copy()
    conv = lib().oracle_mt_update(_p(s), a_t, f_t, g_t)
    return bool(conv), s


def voxelgrid_downsample(pts, leaf):
    """pcl::VoxelGrid<PointXYZI> centroid downsample (odom_node's downSizeFilterLocalmap, ODOM:98, 334-335).
    pts: (n, 4) float32 x y z intensity -> (m, 4), leaves in ascending index order."""
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    assert pts.ndim == 2 and pts.shape[1] == 4
    out = np.zeros_like(pts)
    m = lib().oracle_voxelgrid_downsample(_p(pts), pts.shape[0], 4, 3, float(leaf), _p(out))
    return out[:m].copy()


def filter_scan(pts, r_min=1.0, r_max=60.0, leaf=0.5, mean_k=30, std_mul=1.0):
    """filter_node's front end (reference src/filter_node.cpp:236-273): NaN removal, range gate, pcl::VoxelGrid, statistical
    outlier removal.  pts (n, 4) float32 -> dict(cloud (m, 4), counts (after gate, after downsample, after SOR), distances)."""
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    assert pts.ndim == 2 and pts.shape[1] == 4
    out = np.zeros_like(pts)
    counts = np.zeros(3, dtype=np.uint64)
    dist = np.zeros(pts.shape[0], dtype=np.float32)
    m = lib().oracle_filter_scan(_p(pts), pts.shape[0], r_min, r_max, leaf, mean_k, std_mul, _p(out), _p(counts), _p(dist))
    return dict(cloud=out[:m].copy(), counts=tuple(int(c) for c in counts), distances=dist[:int(counts[1])].copy())


def transform_cloud(pts, T):
    """pcl::transformPointCloud: x' = ((m00 x + m01 y) + m02 z) + m03 in float32, other fields copied."""
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    T = np.asarray(T, dtype=np.float32)
    out = pts.copy()
    for r in range(3):
        out[:, r] = ((T[r, 0] * pts[:, 0] + T[r, 1] * pts[:, 1]) + T[r, 2] * pts[:, 2]) + T[r, 3]
    return out


def get_transformation(x, y, z, roll, pitch, yaw):
    """pcl::getTransformation (float overload; PCL common/eigen.h, not in /root/reference -- restated from the published
    formula): R = Rz(yaw) Ry(pitch) Rx(roll) written out entry by entry in float, sinf/cosf per thirdparty_math.h's convention
    ((float) of the double function).  Used by PGO::TransformCloud2Map (pgo_node.cpp:574)."""
    import struct

    def f32(v):
        return struct.unpack("f", struct.pack("f", float(v)))[0]
    import math
    x, y, z, roll, pitch, yaw = (f32(v) for v in (x, y, z, roll, pitch, yaw))
    A, B = f32(math.cos(yaw)), f32(math.sin(yaw))
    Cc, D = f32(math.cos(pitch)), f32(math.sin(pitch))
    E, F = f32(math.cos(roll)), f32(math.sin(roll))
    DE, DF = f32(D * E), f32(D * F)
    rows = [[f32(A * Cc), f32(f32(A * DF) - f32(B * E)), f32(f32(B * F) + f32(A * DE)), x],
            [f32(B * Cc), f32(f32(A * E) + f32(B * DF)), f32(f32(B * DE) - f32(A * F)), y],
            [f32(-D), f32(Cc * F), f32(Cc * E), z],
            [0.0, 0.0, 0.0, 1.0]]
    return np.array(rows, dtype=np.float32)


def loop_find_near_keyframes_cloud(keyframes, pose_mats, key, submap_size, root_idx, leaf):
    """PGO::LoopFindNearKeyframesCloud (pgo_node.cpp:530-554): keyframes root_idx - submap_size .. root_idx + submap_size that
    exist, each through TransformCloud2Map with its own pose (pgo_node.cpp:542, 568-590), stacked, then downSizeFilterICP.
    `key` is accepted and ignored exactly as the reference ignores it (pgo_node.cpp:537).  keyframes: list of (n, 4) float32;
    pose_mats: list of 4x4 float32 (get_transformation of keyframePosesUpdated[i]).  An empty stack is returned as is
    (pgo_node.cpp:547-548)."""
    del key
    parts = []
    for i in range(-submap_size, submap_size + 1):
        k = root_idx + i
        if k < 0 or k >= len(keyframes):
            continue
        parts.append(transform_cloud(keyframes[k], pose_mats[k]))
    if not parts or sum(p.shape[0] for p in parts) == 0:
        return np.zeros((0, 4), dtype=np.float32)
    near = np.ascontiguousarray(np.concatenate(parts, axis=0))
    return voxelgrid_downsample(near, leaf) if leaf > 0 else near


def max_threads():
    return lib().oracle_max_threads()


class OracleNDT:
    """Mirror of pclomp::NormalDistributionsTransform (NDT_H:110-204) over the CPU oracle."""

    def __init__(self, native=False):
        """native=True: the -O3 -march=native build made on this machine (bench.py's timed CPU legs only)."""
        self.L = lib(native)
        self.native = native and self.L is _lib_native
        self.h = C.c_void_p(self.L.oracle_create())
        # NDT_I:46-69 ctor defaults
        self.resolution, self.step_size, self.trans_eps = 1.0, 0.1, 0.1
        self.max_iter, self.outlier_ratio, self.search_method, self.num_threads = 35, 0.55, DIRECT7, 1
        self.n_source = 0
        self._push()

    def __del__(self):
        try:
            self.L.oracle_destroy(self.h)
        except Exception:
            pass

    def _push(self):
        self.L.oracle_set_params(self.h, self.resolution, self.step_size, self.trans_eps, self.max_iter,
                                 self.outlier_ratio, self.search_method, self.num_threads)

    def set_params(self, **kw):
        for k, v in kw.items():
            assert hasattr(self, k), k
            setattr(self, k, v)
        self._push()

    def set_voxel_params(self, min_points=6, eig_mult=0.01):
        self.L.oracle_set_voxel_params(self.h, min_points, eig_mult)

    def set_target(self, pts):
        pts = _pts(pts)
        self.L.oracle_set_target(self.h, _p(pts), pts.shape[0], pts.shape[1])

    def update_target(self, pts):
        pts = _pts(pts)
        self.L.oracle_update_target(self.h, _p(pts), pts.shape[0], pts.shape[1])

    def set_source(self, pts):
        pts = _pts(pts)
        self.n_source = pts.shape[0]
        self.L.oracle_set_source(self.h, _p(pts), pts.shape[0], pts.shape[1])

    def grid_valid(self):
        return bool(self.L.oracle_grid_valid(self.h))

    def dump_leaves(self):
        n = self.L.oracle_num_leaves(self.h)
        out = dict(ijk=np.zeros((n, 3), np.int32), nr_points=np.zeros(n, np.int32), n_raw=np.zeros(n, np.int32),
                   mean=np.zeros((n, 3)), cov=np.zeros((n, 3, 3)), icov=np.zeros((n, 3, 3)),
                   centroid=np.zeros((n, 3), np.float32), evals=np.zeros((n, 3)))
        self.L.oracle_dump_leaves(self.h, _p(out["ijk"]), _p(out["nr_points"]), _p(out["n_raw"]), _p(out["mean"]),
                                  _p(out["cov"]), _p(out["icov"]), _p(out["centroid"]), _p(out["evals"]))
        return out

    def neighbors(self, pts_xyz, method=None):
        pts = np.ascontiguousarray(pts_xyz, dtype=np.float32)[:, :3].copy()
        n = pts.shape[0]
        cnt = np.zeros(n, np.int32)
        ijk = np.zeros((n, 27, 3), np.int32)
        self.L.oracle_neighbors(self.h, _p(pts), n, self.search_method if method is None else method, _p(cnt), _p(ijk))
        return cnt, ijk

    def gauss(self):
        d = [C.c_double(), C.c_double(), C.c_double()]
        self.L.oracle_gauss(self.h, C.byref(d[0]), C.byref(d[1]), C.byref(d[2]))
        return tuple(x.value for x in d)

    def eval_derivatives(self, T, p, compute_hessian=True):
        Tc = np.ascontiguousarray(np.asarray(T, dtype=np.float32).T).reshape(-1)
        p = np.ascontiguousarray(p, dtype=np.float64)
        g, H = np.zeros(6), np.zeros((6, 6))
        npairs = C.c_longlong()
        s = self.L.oracle_eval_derivatives(self.h, _p(Tc), _p(p), int(compute_hessian), _p(g), _p(H), C.byref(npairs))
        return s, g, H, npairs.value

    def compute_hessian(self, T, p):
        Tc = np.ascontiguousarray(np.asarray(T, dtype=np.float32).T).reshape(-1)
        p = np.ascontiguousarray(p, dtype=np.float64)
        H = np.zeros((6, 6))
        self.L.oracle_compute_hessian(self.h, _p(Tc), _p(p), _p(H))
        return H

    def transform_source(self, T):
        Tc = np.ascontiguousarray(np.asarray(T, dtype=np.float32).T).reshape(-1)
        out = np.zeros((self.n_source, 3), np.float32)
        self.L.oracle_transform_source(self.h, _p(Tc), _p(out))
        return out

    def align(self, guess=None):
        G = np.eye(4, dtype=np.float32) if guess is None else np.asarray(guess, dtype=np.float32)
        Gc = np.ascontiguousarray(G.T).reshape(-1)
        r = _Result()
        self.L.oracle_align(self.h, _p(Gc), C.byref(r))
        T = np.array(r.T, dtype=np.float32).reshape(4, 4).T.copy()
        tr = np.zeros((max(r.n_evals, 1), 52))
        self.L.oracle_get_trace(self.h, _p(tr), tr.shape[0])
        tr = tr[:r.n_evals]
        trace = [dict(p=t[0:6].copy(), score=t[6], g=t[7:13].copy(), H=t[13:49].reshape(6, 6).copy(), step=t[49],
                      compute_hessian=int(t[50]), kind=int(t[51])) for t in tr]
        return dict(T=T, iterations=r.iterations, converged=bool(r.converged), score=r.score,
                    trans_probability=r.trans_probability, trace=trace)

    def get_output(self):
        out = np.zeros((self.n_source, 3), np.float32)
        self.L.oracle_get_output(self.h, _p(out))
        return out

    def fitness(self, max_range=float(np.finfo(np.float64).max), T=None):
        if T is not None:
            Tc = np.ascontiguousarray(np.asarray(T, dtype=np.float32).T).reshape(-1)
            self.L.oracle_set_final_transform(self.h, _p(Tc))
        return self.L.oracle_fitness(self.h, max_range)

    def build_nn(self):
        self.L.oracle_build_nn(self.h)

    def calculate_score(self, T):
        Tc = np.ascontiguousarray(np.asarray(T, dtype=np.float32).T).reshape(-1)
        return self.L.oracle_calculate_score(self.h, _p(Tc))
