"""ctypes binding for oracle/_ref/libndt_ref.so: the reference's OWN NDT sources
(/root/reference/xchu_mapping/include/pclomp/*), compiled unmodified against the stand-in
Eigen/PCL headers of oracle/ref_stubs/ (recipe: oracle/Makefile target `ref`).

TEST INFRASTRUCTURE ONLY -- used by tests/ and tests/golden/make_golden.py to pin oracle/ndt_oracle.cpp
against the reference's code.  The library is built only where /root/reference exists (this
container); the prebuilt file travels to the GPU box.  Same method names as oracle.OracleNDT.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_ref", "libndt_ref.so")
REFERENCE_INCLUDE = "/root/reference/xchu_mapping/include"

KDTREE, DIRECT26, DIRECT7, DIRECT1 = 0, 1, 2, 3


def can_build():
    return os.path.exists(os.path.join(REFERENCE_INCLUDE, "pclomp", "ndt_omp_impl.hpp"))


def build(force=False):
    """Returns the library path, or None when neither the reference sources nor a prebuilt library exist."""
    if can_build():
        subprocess.check_call(["make", "-C", _HERE, "-s", "ref"] + (["-B"] if force else []))
    return _LIB if os.path.exists(_LIB) else None


def available():
    return build() is not None


class _Result(C.Structure):
    _fields_ = [("T", C.c_float * 16), ("iterations", C.c_int), ("converged", C.c_int),
                ("score", C.c_double), ("trans_probability", C.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = build()
        if path is None:
            raise RuntimeError("oracle/_ref/libndt_ref.so is absent and /root/reference is not here to build it")
        L = C.CDLL(path)
        L.ref_create.restype = C.c_void_p
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_set_params.argtypes = [C.c_void_p, C.c_float, C.c_double, C.c_double, C.c_int, C.c_double, C.c_int, C.c_int]
        L.ref_set_voxel_params.argtypes = [C.c_void_p, C.c_int, C.c_double]
        for name in ("ref_set_target", "ref_set_source"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t]
        L.ref_grid_bounds.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_num_leaves.argtypes = [C.c_void_p]
        L.ref_num_leaves.restype = C.c_size_t
        L.ref_num_centroids.argtypes = [C.c_void_p]
        L.ref_num_centroids.restype = C.c_size_t
        L.ref_dump_leaves.argtypes = [C.c_void_p] + [C.c_void_p] * 7
        L.ref_neighbors.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_gauss.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_eval_derivatives.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_eval_derivatives.restype = C.c_double
        L.ref_compute_hessian.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_calculate_score.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_calculate_score.restype = C.c_double
        L.ref_align.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(_Result)]
        L.ref_get_output.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_fitness.argtypes = [C.c_void_p, C.c_double]
        L.ref_fitness.restype = C.c_double
        L.ref_pose_to_matrix.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_mt_trial.argtypes = [C.c_void_p] + [C.c_double] * 9
        L.ref_mt_trial.restype = C.c_double
        L.ref_mt_update.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_double]
        L.ref_ctor_defaults.argtypes = [C.c_void_p] + [C.c_void_p] * 8
        _lib = L
    return _lib


def _pts(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    assert a.ndim == 2 and a.shape[1] >= 3
    return a


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _colmajor(T):
    return np.ascontiguousarray(np.asarray(T, dtype=np.float32).T).reshape(-1)


def pose_to_matrix(p):
    """NDT_H:210-229 convertTransform, the reference's own code over the stand-in Eigen geometry."""
    p = np.ascontiguousarray(p, dtype=np.float64)
    T = np.zeros(16, dtype=np.float32)
    lib().ref_pose_to_matrix(_p(p), _p(T))
    return T.reshape(4, 4).T.copy()


class RefNDT:
    """pclomp::NormalDistributionsTransform<PointXYZI, PointXYZI> of the reference, through oracle/ref_shim.cpp."""

    def __init__(self):
        self.L = lib()
        self.h = C.c_void_p(self.L.ref_create())
        d = self.ctor_defaults()
        self.resolution, self.step_size, self.trans_eps = d["resolution"], d["step_size"], d["trans_eps"]
        self.max_iter, self.outlier_ratio, self.search_method, self.num_threads = d["max_iter"], d["outlier_ratio"], d["search_method"], 1
        self.n_source = 0
        self._push()

    def __del__(self):
        try:
            self.L.ref_destroy(self.h)
        except Exception:
            pass

    def ctor_defaults(self):
        res, mi, me, mp = C.c_float(), C.c_int(), C.c_int(), C.c_int()
        st, ou, ep, em = C.c_double(), C.c_double(), C.c_double(), C.c_double()
        self.L.ref_ctor_defaults(self.h, C.byref(res), C.byref(st), C.byref(ou), C.byref(ep), C.byref(mi), C.byref(me), C.byref(mp), C.byref(em))
        return dict(resolution=res.value, step_size=st.value, outlier_ratio=ou.value, trans_eps=ep.value, max_iter=mi.value,
                    search_method=me.value, min_points=mp.value, eig_mult=em.value)

    def _push(self):
        self.L.ref_set_params(self.h, self.resolution, self.step_size, self.trans_eps, self.max_iter, self.outlier_ratio,
                              self.search_method, self.num_threads)

    def set_params(self, **kw):
        for k, v in kw.items():
            assert hasattr(self, k), k
            setattr(self, k, v)
        self._push()

    def set_voxel_params(self, min_points=6, eig_mult=0.01):
        self.L.ref_set_voxel_params(self.h, min_points, eig_mult)

    def set_target(self, pts):
        pts = _pts(pts)
        self.L.ref_set_target(self.h, _p(pts), pts.shape[0], pts.shape[1])

    def set_source(self, pts):
        pts = _pts(pts)
        self.n_source = pts.shape[0]
        self.L.ref_set_source(self.h, _p(pts), pts.shape[0], pts.shape[1])

    def grid_bounds(self):
        a, b = np.zeros(3, np.int32), np.zeros(3, np.int32)
        self.L.ref_grid_bounds(self.h, _p(a), _p(b))
        return a, b

    def num_centroids(self):
        return self.L.ref_num_centroids(self.h)

    def dump_leaves(self):
        n = self.L.ref_num_leaves(self.h)
        out = dict(ijk=np.zeros((n, 3), np.int32), nr_points=np.zeros(n, np.int32), mean=np.zeros((n, 3)), cov=np.zeros((n, 3, 3)),
                   icov=np.zeros((n, 3, 3)), centroid=np.zeros((n, 3), np.float32), evals=np.zeros((n, 3)))
        self.L.ref_dump_leaves(self.h, _p(out["ijk"]), _p(out["nr_points"]), _p(out["mean"]), _p(out["cov"]), _p(out["icov"]),
                               _p(out["centroid"]), _p(out["evals"]))
        return out

    def neighbors(self, pts_xyz, method=None):
        pts = np.ascontiguousarray(pts_xyz, dtype=np.float32)[:, :3].copy()
        n = pts.shape[0]
        cnt = np.zeros(n, np.int32)
        ijk = np.zeros((n, 27, 3), np.int32)
        self.L.ref_neighbors(self.h, _p(pts), n, self.search_method if method is None else method, _p(cnt), _p(ijk))
        return cnt, ijk

    def gauss(self):
        d = [C.c_double(), C.c_double(), C.c_double()]
        self.L.ref_gauss(self.h, C.byref(d[0]), C.byref(d[1]), C.byref(d[2]))
        return tuple(x.value for x in d)

    def eval_derivatives(self, T, p, compute_hessian=True):
        Tc, p = _colmajor(T), np.ascontiguousarray(p, dtype=np.float64)
        g, H = np.zeros(6), np.zeros((6, 6))
        s = self.L.ref_eval_derivatives(self.h, _p(Tc), _p(p), int(compute_hessian), _p(g), _p(H))
        return s, g, H

    def compute_hessian(self, T, p):
        Tc, p = _colmajor(T), np.ascontiguousarray(p, dtype=np.float64)
        H = np.zeros((6, 6))
        self.L.ref_compute_hessian(self.h, _p(Tc), _p(p), _p(H))
        return H

    def calculate_score(self, T):
        return self.L.ref_calculate_score(self.h, _p(_colmajor(T)))

    def align(self, guess=None):
        G = np.eye(4, dtype=np.float32) if guess is None else np.asarray(guess, dtype=np.float32)
        r = _Result()
        self.L.ref_align(self.h, _p(_colmajor(G)), C.byref(r))
        T = np.array(r.T, dtype=np.float32).reshape(4, 4).T.copy()
        return dict(T=T, iterations=r.iterations, converged=bool(r.converged), score=r.score, trans_probability=r.trans_probability)

    def get_output(self):
        out = np.zeros((self.n_source, 3), np.float32)
        self.L.ref_get_output(self.h, _p(out))
        return out

    def fitness(self, max_range=float(np.finfo(np.float64).max)):
        return self.L.ref_fitness(self.h, max_range)

    def mt_trial(self, *args):
        return self.L.ref_mt_trial(self.h, *[float(a) for a in args])

    def mt_update(self, state, a_t, f_t, g_t):
        s = np.ascontiguousarray(state, dtype=np.float64).copy()
        conv = self.L.ref_mt_update(self.h, _p(s), a_t, f_t, g_t)
        return bool(conv), s
