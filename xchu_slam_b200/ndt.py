"""Python mirror of pclomp::NormalDistributionsTransform over the C ABI of libndt_b200.so.

Method names and argument meaning follow the reference class (reference
xchu_mapping/include/pclomp/ndt_omp.h:110-204 and the pcl::Registration members odom_node uses,
src/odom_node.cpp:55-83, 263-284), so the parity tests read like reference usage.  All compute
happens in the CUDA library; there is no CPU fallback -- a missing library or GPU raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NDTB_LIB") or os.path.join(_HERE, "libndt_b200.so")   # NDTB_LIB: kernel-variant experiments

KDTREE, DIRECT26, DIRECT7, DIRECT1 = 0, 1, 2, 3


NOALLOC_MAX_SOURCE = 262144     # NDTB_NOALLOC_MAX_SOURCE (ndt_c_api.h)


def fitness_noalloc(device, source_ptr, n_source, target_ptr, n_target, T, max_range, out_ptr, stream=0):
    """ndtb_fitness_noalloc: getFitnessScore with every buffer the caller's (DEVICE pointers as integers; `out_ptr` -> one double on
    the device; `stream` a cudaStream_t as integer) and its scratch in a static device array: no allocation, no synchronisation.
    No counterpart in the reference.  Raises NdtError with the library's text on rejection."""
    L = load_library()
    Tc = _colmajor(T)
    rc = L.ndtb_fitness_noalloc(int(device), C.c_void_p(source_ptr), int(n_source), C.c_void_p(target_ptr), int(n_target), _p(Tc),
                                float(max_range), C.c_void_p(out_ptr), C.c_void_p(stream))
    if rc != 0:
        raise NdtError(L.ndtb_noalloc_last_error().decode())


class NdtError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [("resolution", C.c_float), ("step_size", C.c_double), ("trans_eps", C.c_double), ("max_iter", C.c_int),
                ("outlier_ratio", C.c_double), ("search_method", C.c_int), ("min_points", C.c_int), ("eig_mult", C.c_double),
                ("strict_math", C.c_int), ("dense_max_cells", C.c_longlong), ("replay_iters", C.c_int), ("replay_min_norm", C.c_double)]


class Result(C.Structure):
    _fields_ = [("T", C.c_float * 16), ("iterations", C.c_int), ("converged", C.c_int), ("score", C.c_double),
                ("trans_probability", C.c_double), ("n_evals", C.c_int)]


class FilterParams(C.Structure):
    _fields_ = [("r_min", C.c_double), ("r_max", C.c_double), ("leaf", C.c_float), ("mean_k", C.c_int), ("stddev_mul", C.c_double)]


RESULT_DTYPE = np.dtype([("T", np.float32, 16), ("iterations", np.int32), ("converged", np.int32), ("score", np.float64),
                         ("trans_probability", np.float64), ("n_evals", np.int32)], align=True)
assert RESULT_DTYPE.itemsize == C.sizeof(Result)


class BatchStats(C.Structure):
    _fields_ = [("build_ms", C.c_double), ("deriv_ms", C.c_double), ("total_ms", C.c_double),
                ("kernel_launches", C.c_longlong), ("deriv_launches", C.c_longlong),
                ("points_evaluated", C.c_longlong), ("pairs_evaluated", C.c_longlong), ("probes", C.c_longlong),
                ("replayed", C.c_longlong), ("replay_ms", C.c_double), ("replay_deriv_ms", C.c_double),
                ("replay_deriv_launches", C.c_longlong), ("replay_points", C.c_longlong), ("replay_pairs", C.c_longlong)]


# every symbol include/ndt_b200/ndt_c_api.h declares
SYMBOLS = ["ndtb_create", "ndtb_destroy", "ndtb_last_error", "ndtb_default_params", "ndtb_set_params", "ndtb_get_params",
           "ndtb_set_target", "ndtb_set_target_auto", "ndtb_update_target", "ndtb_set_source", "ndtb_align", "ndtb_get_output",
           "ndtb_fitness", "ndtb_calculate_score", "ndtb_num_voxels", "ndtb_dump_voxels", "ndtb_dump_icov64", "ndtb_pose_to_matrix", "ndtb_host_alloc", "ndtb_host_free", "ndtb_fitness_noalloc", "ndtb_noalloc_last_error", "ndtb_neighbors",
           "ndtb_eval_derivatives", "ndtb_compute_hessian", "ndtb_get_trace", "ndtb_last_timing", "ndtb_batch_create",
           "ndtb_batch_destroy", "ndtb_batch_last_error", "ndtb_batch_set_params", "ndtb_batch_clear", "ndtb_batch_add_target",
           "ndtb_batch_add_source", "ndtb_batch_wait_uploads", "ndtb_batch_run", "ndtb_batch_get_stats", "ndtb_batch_set_stream", "ndtb_batch_fitness",
           "ndtb_localmap_clear", "ndtb_localmap_append_source", "ndtb_localmap_snapshot", "ndtb_localmap_set_target",
           "ndtb_localmap_roll", "ndtb_localmap_sizes", "ndtb_localmap_download", "ndtb_default_filter_params", "ndtb_filter_scan",
           "ndtb_get_source", "ndtb_batch_add_keyframe", "ndtb_batch_clear_keyframes", "ndtb_batch_num_keyframes",
           "ndtb_batch_add_submap", "ndtb_batch_download", "ndtb_batch_add_targets", "ndtb_batch_add_sources",
           "ndtb_batch_add_submaps", "ndtb_batch_get_diag", "ndtb_batch_get_step_history", "ndtb_multi_create", "ndtb_multi_destroy", "ndtb_multi_last_error",
           "ndtb_multi_set_params", "ndtb_multi_run", "ndtb_multi_get_stats", "ndtb_window_add_source", "ndtb_window_drop",
           "ndtb_window_clear", "ndtb_window_set_target", "ndtb_get_target"]

_lib = None


def load_library():
    """dlopen the in-tree CUDA library.  Raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NdtError("%s is missing: run `python -m xchu_slam_b200.build` (nvcc, sm_100a). "
                       "There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, sz = C.c_void_p, C.c_size_t
    L.ndtb_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.ndtb_destroy.argtypes = [vp]
    L.ndtb_last_error.argtypes = [vp]
    L.ndtb_last_error.restype = C.c_char_p
    L.ndtb_default_params.argtypes = [C.POINTER(Params)]
    L.ndtb_default_params.restype = None
    L.ndtb_set_params.argtypes = [vp, C.POINTER(Params)]
    L.ndtb_get_params.argtypes = [vp, C.POINTER(Params)]
    L.ndtb_set_target.argtypes = [vp, vp, sz, sz]
    L.ndtb_set_target_auto.argtypes = [vp, vp, sz, sz, C.POINTER(C.c_int)]
    L.ndtb_update_target.argtypes = [vp, vp, sz, sz]
    L.ndtb_set_source.argtypes = [vp, vp, sz, sz]
    L.ndtb_align.argtypes = [vp, vp, C.POINTER(Result)]
    L.ndtb_get_output.argtypes = [vp, vp, sz]
    L.ndtb_fitness.argtypes = [vp, C.c_double, vp, C.POINTER(C.c_double)]
    L.ndtb_calculate_score.argtypes = [vp, vp, C.POINTER(C.c_double)]
    L.ndtb_num_voxels.argtypes = [vp, C.POINTER(sz)]
    L.ndtb_dump_voxels.argtypes = [vp] + [vp] * 6
    L.ndtb_dump_icov64.argtypes = [vp, vp]
    L.ndtb_pose_to_matrix.argtypes = [vp, vp]
    L.ndtb_pose_to_matrix.restype = None
    L.ndtb_host_alloc.argtypes = [sz, C.POINTER(vp)]
    L.ndtb_host_free.argtypes = [vp]
    L.ndtb_fitness_noalloc.argtypes = [C.c_int, vp, sz, vp, sz, vp, C.c_double, vp, vp]
    L.ndtb_noalloc_last_error.argtypes = []
    L.ndtb_noalloc_last_error.restype = C.c_char_p
    L.ndtb_neighbors.argtypes = [vp, vp, sz, C.c_int, vp, vp]
    L.ndtb_eval_derivatives.argtypes = [vp, vp, vp, C.c_int, C.POINTER(C.c_double), vp, vp]
    L.ndtb_compute_hessian.argtypes = [vp, vp, vp, vp]
    L.ndtb_get_trace.argtypes = [vp, vp, C.c_int, C.POINTER(C.c_int)]
    L.ndtb_last_timing.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.ndtb_batch_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.ndtb_batch_destroy.argtypes = [vp]
    L.ndtb_batch_last_error.argtypes = [vp]
    L.ndtb_batch_last_error.restype = C.c_char_p
    L.ndtb_batch_set_params.argtypes = [vp, C.POINTER(Params)]
    L.ndtb_batch_clear.argtypes = [vp]
    L.ndtb_batch_add_target.argtypes = [vp, vp, sz, sz, C.c_int, C.POINTER(C.c_int)]
    L.ndtb_batch_add_source.argtypes = [vp, vp, sz, sz, C.c_int, C.POINTER(C.c_int)]
    L.ndtb_batch_add_targets.argtypes = [vp, C.c_int, vp, vp, sz, C.c_int, C.POINTER(C.c_int)]
    L.ndtb_batch_add_sources.argtypes = [vp, C.c_int, vp, vp, sz, C.c_int, C.POINTER(C.c_int)]
    L.ndtb_batch_wait_uploads.argtypes = [vp]
    L.ndtb_batch_run.argtypes = [vp, C.c_int, vp, vp, vp, vp]
    L.ndtb_batch_get_stats.argtypes = [vp, C.POINTER(BatchStats)]
    L.ndtb_batch_get_diag.argtypes = [vp, vp, vp, vp, vp]
    L.ndtb_batch_get_step_history.argtypes = [vp, vp, C.c_int]
    L.ndtb_window_add_source.argtypes = [vp, vp, C.c_float, C.POINTER(C.c_int), C.POINTER(sz)]
    L.ndtb_window_drop.argtypes = [vp, C.c_int]
    L.ndtb_window_clear.argtypes = [vp]
    L.ndtb_window_set_target.argtypes = [vp, vp, C.c_int, C.c_float, C.POINTER(sz)]
    L.ndtb_get_target.argtypes = [vp, vp, sz, sz, C.POINTER(sz)]
    L.ndtb_multi_create.argtypes = [C.c_int, vp, C.POINTER(vp)]
    L.ndtb_multi_destroy.argtypes = [vp]
    L.ndtb_multi_last_error.argtypes = [vp]
    L.ndtb_multi_last_error.restype = C.c_char_p
    L.ndtb_multi_set_params.argtypes = [vp, C.POINTER(Params)]
    L.ndtb_multi_run.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp, vp, sz, C.c_int, vp, vp, vp, vp]
    L.ndtb_multi_get_stats.argtypes = [vp, vp, vp, vp, vp]
    L.ndtb_batch_set_stream.argtypes = [vp, vp]
    L.ndtb_batch_fitness.argtypes = [vp, C.c_double, vp]
    L.ndtb_batch_add_keyframe.argtypes = [vp, vp, sz, sz, C.POINTER(C.c_int)]
    L.ndtb_batch_clear_keyframes.argtypes = [vp]
    L.ndtb_batch_num_keyframes.argtypes = [vp]
    L.ndtb_batch_add_submap.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_float, C.POINTER(C.c_int), C.POINTER(sz)]
    L.ndtb_batch_add_submaps.argtypes = [vp, C.c_int, vp, vp, vp, vp, C.c_int, C.c_float, vp, vp]
    L.ndtb_batch_download.argtypes = [vp, C.c_int, C.c_int, vp, sz, sz, C.POINTER(sz)]
    L.ndtb_default_filter_params.argtypes = [C.POINTER(FilterParams)]
    L.ndtb_filter_scan.argtypes = [vp, vp, sz, sz, C.POINTER(FilterParams), vp]
    L.ndtb_get_source.argtypes = [vp, vp, sz, sz, C.POINTER(sz)]
    L.ndtb_localmap_clear.argtypes = [vp]
    L.ndtb_localmap_append_source.argtypes = [vp, vp, C.c_float, C.c_int, C.POINTER(sz)]
    L.ndtb_localmap_snapshot.argtypes = [vp]
    L.ndtb_localmap_set_target.argtypes = [vp, C.c_int, C.POINTER(C.c_int)]
    L.ndtb_localmap_roll.argtypes = [vp]
    L.ndtb_localmap_sizes.argtypes = [vp, vp]
    L.ndtb_localmap_download.argtypes = [vp, C.c_int, vp, sz, sz, C.POINTER(sz)]
    _lib = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _cloud(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if a.ndim != 2 or a.shape[1] < 3:
        raise ValueError("point cloud must be (n, >=3) float32")
    return a


def _colmajor(T):
    return np.ascontiguousarray(np.asarray(T, dtype=np.float32).reshape(4, 4).T).reshape(-1)


class NormalDistributionsTransform:
    """pclomp::NormalDistributionsTransform<PointXYZI, PointXYZI> drop-in (reference NDT_H:70-204)."""

    def __init__(self, device=0):
        self._L = load_library()
        self._h = C.c_void_p()
        if self._L.ndtb_create(device, C.byref(self._h)):
            raise NdtError(self._L.ndtb_last_error(None).decode())
        self._params = Params()
        self._L.ndtb_get_params(self._h, C.byref(self._params))
        self._result = None
        self._n_source = 0
        self._src_stride = 16

    def __del__(self):
        try:
            if self._h:
                self._L.ndtb_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise NdtError(self._L.ndtb_last_error(self._h).decode())

    def _push(self):
        self._ck(self._L.ndtb_set_params(self._h, C.byref(self._params)))

    # ---- setters / getters of the reference class ------------------------------------------
    def setNumThreads(self, n):          # NDT_H:110-112 -- CPU threads have no meaning here
        pass

    def setResolution(self, r):          # NDT_H:127-137
        self._params.resolution = float(r)
        self._push()

    def getResolution(self):
        return self._params.resolution

    def setStepSize(self, s):            # NDT_H:160-164
        self._params.step_size = float(s)
        self._push()

    def getStepSize(self):
        return self._params.step_size

    def setOulierRatio(self, o):         # NDT_H:178-182 (sic)
        self._params.outlier_ratio = float(o)
        self._push()

    def getOulierRatio(self):
        return self._params.outlier_ratio

    def setTransformationEpsilon(self, e):
        self._params.trans_eps = float(e)
        self._push()

    def setMaximumIterations(self, n):
        self._params.max_iter = int(n)
        self._push()

    def setNeighborhoodSearchMethod(self, m):   # NDT_H:184-186
        self._params.search_method = int(m)
        self._push()

    def setVoxelParams(self, min_points=6, eig_mult=0.01):
        self._params.min_points, self._params.eig_mult = int(min_points), float(eig_mult)
        self._push()

    def setStrictMath(self, on):
        """1: term-by-term derivative kernel in the reference's float operation order; 0: fast kernel."""
        self._params.strict_math = int(bool(on))
        self._push()

    def setDenseMaxCells(self, n):
        """0: hash-table lookups only; otherwise the dense cell index is used when the box fits."""
        self._params.dense_max_cells = int(n)
        self._push()

    def setInputTarget(self, cloud):     # NDT_H:117-122
        cloud = _cloud(cloud)
        self._ck(self._L.ndtb_set_target(self._h, _p(cloud), cloud.shape[0], cloud.strides[0]))

    def setInputTargetAuto(self, cloud):
        """setInputTarget that merges only the appended tail when the cloud extends the current one."""
        cloud = _cloud(cloud)
        inc = C.c_int()
        self._ck(self._L.ndtb_set_target_auto(self._h, _p(cloud), cloud.shape[0], cloud.strides[0], C.byref(inc)))
        return bool(inc.value)

    def updateVoxelGrid(self, cloud):    # cpu::NormalDistributionsTransform::updateVoxelGrid
        cloud = _cloud(cloud)
        self._ck(self._L.ndtb_update_target(self._h, _p(cloud), cloud.shape[0], cloud.strides[0]))

    def setInputSource(self, cloud):
        cloud = _cloud(cloud)
        self._n_source = cloud.shape[0]
        self._ck(self._L.ndtb_set_source(self._h, _p(cloud), cloud.shape[0], cloud.strides[0]))

    def align(self, guess=None):
        """pcl::Registration::align(output, guess).  Returns the transformed source (n, 3)."""
        G = np.eye(4, dtype=np.float32) if guess is None else np.asarray(guess, dtype=np.float32)
        g = _colmajor(G)
        r = Result()
        self._ck(self._L.ndtb_align(self._h, _p(g), C.byref(r)))
        self._result = r
        return self.getOutput()

    def getOutput(self):
        out = np.zeros((self._n_source, 3), np.float32)
        self._ck(self._L.ndtb_get_output(self._h, _p(out), 12))
        return out

    def getFinalTransformation(self):
        return np.array(self._result.T, dtype=np.float32).reshape(4, 4).T.copy()

    def hasConverged(self):
        return bool(self._result.converged)

    def getFinalNumIteration(self):      # NDT_H:200-204
        return int(self._result.iterations)

    def getTransformationProbability(self):  # NDT_H:191-195
        return float(self._result.trans_probability)

    def getFitnessScore(self, max_range=float(np.finfo(np.float64).max), T=None):
        out = C.c_double()
        Tp = None if T is None else _p(_colmajor(T))
        if T is not None:
            self._Tkeep = _colmajor(T)
            Tp = _p(self._Tkeep)
        self._ck(self._L.ndtb_fitness(self._h, float(max_range), Tp, C.byref(out)))
        return out.value

    def calculateScore(self, T):         # NDT_I:919-952 with cloud = T * source
        out = C.c_double()
        Tc = _colmajor(T)
        self._ck(self._L.ndtb_calculate_score(self._h, _p(Tc), C.byref(out)))
        return out.value

    # ---- inspection hooks -----------------------------------------------------------------
    def result(self):
        r = self._result
        return dict(T=self.getFinalTransformation(), iterations=r.iterations, converged=bool(r.converged), score=r.score,
                    trans_probability=r.trans_probability, n_evals=r.n_evals)

    def trace(self):
        n = C.c_int()
        self._ck(self._L.ndtb_get_trace(self._h, None, 0, C.byref(n)))
        buf = np.zeros((max(n.value, 1), 52))
        self._ck(self._L.ndtb_get_trace(self._h, _p(buf), buf.shape[0], C.byref(n)))
        return [dict(p=t[0:6].copy(), score=t[6], g=t[7:13].copy(), H=t[13:49].reshape(6, 6).copy(), step=t[49],
                     compute_hessian=int(t[50]), kind=int(t[51])) for t in buf[:n.value]]

    def dumpVoxels(self):
        n = C.c_size_t()
        self._ck(self._L.ndtb_num_voxels(self._h, C.byref(n)))
        n = n.value
        out = dict(ijk=np.zeros((n, 3), np.int32), nr_points=np.zeros(n, np.int32), n_raw=np.zeros(n, np.int32),
                   mean=np.zeros((n, 3)), icov=np.zeros((n, 3, 3), np.float32), centroid=np.zeros((n, 3), np.float32),
                   icov64=np.zeros((n, 3, 3)))
        if n:
            self._ck(self._L.ndtb_dump_voxels(self._h, _p(out["ijk"]), _p(out["nr_points"]), _p(out["n_raw"]), _p(out["mean"]),
                                              _p(out["icov"]), _p(out["centroid"])))
            self._ck(self._L.ndtb_dump_icov64(self._h, _p(out["icov64"])))
        return out

    @staticmethod
    def convertTransform(x):
        """NDT_H:210-229 (static in the reference too): [x y z roll pitch yaw] -> 4x4 float, row/column indexable."""
        x = np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(6))
        T = np.zeros(16, np.float32)
        load_library().ndtb_pose_to_matrix(_p(x), _p(T))
        return T.reshape(4, 4).T.copy()

    def neighbors(self, xyz, method=None):
        xyz = np.ascontiguousarray(np.asarray(xyz, dtype=np.float32)[:, :3])
        n = xyz.shape[0]
        cnt = np.zeros(n, np.int32)
        ijk = np.zeros((n, 27, 3), np.int32)
        m = self._params.search_method if method is None else method
        self._ck(self._L.ndtb_neighbors(self._h, _p(xyz), n, m, _p(cnt), _p(ijk)))
        return cnt, ijk

    def evalDerivatives(self, T, p, compute_hessian=True):
        Tc = _colmajor(T)
        p = np.ascontiguousarray(p, dtype=np.float64)
        s = C.c_double()
        g, H = np.zeros(6), np.zeros((6, 6))
        self._ck(self._L.ndtb_eval_derivatives(self._h, _p(Tc), _p(p), int(compute_hessian), C.byref(s), _p(g), _p(H)))
        return s.value, g, H

    def computeHessian(self, T, p):
        Tc = _colmajor(T)
        p = np.ascontiguousarray(p, dtype=np.float64)
        H = np.zeros((6, 6))
        self._ck(self._L.ndtb_compute_hessian(self._h, _p(Tc), _p(p), _p(H)))
        return H

    # ---- filter_node's front end on the device (reference src/filter_node.cpp:236-262) ----
    def filterScan(self, raw, r_min=1.0, r_max=60.0, leaf=0.5, mean_k=30, stddev_mul=1.0):
        """NaN removal, range gate, pcl::VoxelGrid, pcl::StatisticalOutlierRemoval; the result becomes the source.
        Returns (points after the gate, after the downsample, after the outlier removal)."""
        raw = _cloud(raw)
        fp = FilterParams(r_min, r_max, leaf, mean_k, stddev_mul)
        counts = np.zeros(3, dtype=np.uint64)
        self._ck(self._L.ndtb_filter_scan(self._h, _p(raw), raw.shape[0], raw.strides[0], C.byref(fp), _p(counts)))
        self._n_source = int(counts[2])
        return tuple(int(c) for c in counts)

    def getSource(self):
        out = np.zeros((max(1, self._n_source), 4), np.float32)
        n = C.c_size_t()
        self._ck(self._L.ndtb_get_source(self._h, _p(out), out.shape[0], 16, C.byref(n)))
        return out[:n.value].copy()

    # ---- odom_node's localmap on the device (reference src/odom_node.cpp:218-233, 289-290, 329-356) ----
    def localmapClear(self):
        self._ck(self._L.ndtb_localmap_clear(self._h))

    def localmapAppendSource(self, T, leaf=0.0, also_tmp=True):
        """localmap (+ tmp_map) += voxel-downsampled (leaf > 0) T * source.  Returns the number of points appended."""
        n = C.c_size_t()
        Tc = _colmajor(T)
        self._ck(self._L.ndtb_localmap_append_source(self._h, _p(Tc), float(leaf), int(also_tmp), C.byref(n)))
        return n.value

    def localmapSnapshot(self):
        self._ck(self._L.ndtb_localmap_snapshot(self._h))

    def localmapSetTarget(self, force_full=False):
        inc = C.c_int()
        self._ck(self._L.ndtb_localmap_set_target(self._h, int(force_full), C.byref(inc)))
        return bool(inc.value)

    def localmapRoll(self):
        self._ck(self._L.ndtb_localmap_roll(self._h))

    def localmapSizes(self):
        a = np.zeros(3, dtype=np.uint64)
        self._ck(self._L.ndtb_localmap_sizes(self._h, _p(a)))
        return tuple(int(x) for x in a)

    def localmapDownload(self, which=0):
        """which: 0 localmap, 1 tmp_map, 2 the cloud appended last.  Returns (n, 4) float32 x y z intensity."""
        cap = max(1, self.localmapSizes()[0] if which == 0 else self.localmapSizes()[1] if which == 1 else self._n_source)
        out = np.zeros((cap, 4), np.float32)
        n = C.c_size_t()
        self._ck(self._L.ndtb_localmap_download(self._h, int(which), _p(out), cap, 16, C.byref(n)))
        return out[:n.value].copy()

    # ---- the keyframe-count window on the device (reference src/odom_node.cpp:737-802 ExtractSurroundKeyframes) ----
    def windowAddSource(self, T, leaf=1.0):
        """tf_cloud = VoxelGrid(leaf)(T * source) kept in HBM.  Returns (id, points)."""
        i, n = C.c_int(), C.c_size_t()
        Tc = _colmajor(T)
        self._ck(self._L.ndtb_window_add_source(self._h, _p(Tc), float(leaf), C.byref(i), C.byref(n)))
        return i.value, n.value

    def windowDrop(self, id):
        self._ck(self._L.ndtb_window_drop(self._h, int(id)))

    def windowClear(self):
        self._ck(self._L.ndtb_window_clear(self._h))

    def windowSetTarget(self, ids, leaf=0.5):
        """pc_target_ = VoxelGrid(leaf)(concatenation of the listed clouds, in that order); setInputTarget.  Returns its size."""
        a = np.ascontiguousarray(ids, dtype=np.int32)
        n = C.c_size_t()
        self._ck(self._L.ndtb_window_set_target(self._h, _p(a), a.shape[0], float(leaf), C.byref(n)))
        self._n_target = n.value
        return n.value

    def getTarget(self, cap):
        out = np.zeros((max(1, int(cap)), 4), np.float32)
        n = C.c_size_t()
        self._ck(self._L.ndtb_get_target(self._h, _p(out), out.shape[0], 16, C.byref(n)))
        return out[:n.value].copy()

    def lastTiming(self):
        b, a = C.c_double(), C.c_double()
        self._L.ndtb_last_timing(self._h, C.byref(b), C.byref(a))
        return dict(build_ms=b.value, align_ms=a.value)


class BatchNDT:
    """Batched registration on one device (north_star item 4): targets and sources resident in HBM,
    any number of {target, source, guess} items per run."""

    def __init__(self, device=0):
        self._L = load_library()
        self._h = C.c_void_p()
        if self._L.ndtb_batch_create(device, C.byref(self._h)):
            raise NdtError(self._L.ndtb_batch_last_error(None).decode())
        self.params = Params()
        self._L.ndtb_default_params(C.byref(self.params))

    def __del__(self):
        try:
            if self._h:
                self._L.ndtb_batch_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise NdtError(self._L.ndtb_batch_last_error(self._h).decode())

    def set_params(self, **kw):
        for k, v in kw.items():
            if not hasattr(self.params, k):
                raise KeyError(k)
            setattr(self.params, k, v)
        self._ck(self._L.ndtb_batch_set_params(self._h, C.byref(self.params)))

    def clear(self):
        self._ck(self._L.ndtb_batch_clear(self._h))

    def set_stream(self, cuda_stream_ptr):
        """Run on the caller's stream (e.g. torch.cuda.current_stream().cuda_stream)."""
        self._ck(self._L.ndtb_batch_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def add_target(self, cloud):
        cloud = _cloud(cloud)
        i = C.c_int()
        self._ck(self._L.ndtb_batch_add_target(self._h, _p(cloud), cloud.shape[0], cloud.strides[0], 0, C.byref(i)))
        return i.value

    def add_source(self, cloud):
        cloud = _cloud(cloud)
        i = C.c_int()
        self._ck(self._L.ndtb_batch_add_source(self._h, _p(cloud), cloud.shape[0], cloud.strides[0], 0, C.byref(i)))
        return i.value

    def add_target_ptr(self, ptr, n, stride=16, on_device=True):
        """on_device: False / 0 host pointer (async upload), True / 1 device pointer (copied), 2 device pointer borrowed."""
        i = C.c_int()
        self._ck(self._L.ndtb_batch_add_target(self._h, C.c_void_p(ptr), n, stride, int(on_device), C.byref(i)))
        return i.value

    def add_source_ptr(self, ptr, n, stride=16, on_device=True):
        i = C.c_int()
        self._ck(self._L.ndtb_batch_add_source(self._h, C.c_void_p(ptr), n, stride, int(on_device), C.byref(i)))
        return i.value

    @staticmethod
    def pointer_table(ptrs, ns):
        """(void*[n], size_t[n]) for add_targets_ptr / add_sources_ptr; build it once for clouds that stay where they are."""
        n = len(ptrs)
        return (C.c_void_p * n)(*ptrs), (C.c_size_t * n)(*ns), n

    def add_targets_ptr(self, table, stride=16, on_device=True):
        """Many clouds in one C-ABI call; table = BatchNDT.pointer_table(pointers, sizes).  Returns the first id."""
        i = C.c_int()
        self._ck(self._L.ndtb_batch_add_targets(self._h, table[2], table[0], table[1], stride, int(on_device), C.byref(i)))
        return i.value

    def add_sources_ptr(self, table, stride=16, on_device=True):
        i = C.c_int()
        self._ck(self._L.ndtb_batch_add_sources(self._h, table[2], table[0], table[1], stride, int(on_device), C.byref(i)))
        return i.value

    def wait_uploads(self):
        self._ck(self._L.ndtb_batch_wait_uploads(self._h))

    def run_raw(self, item_target, item_source, guesses):
        """Runs the registrations; returns the ndtb_result records as a numpy structured array
        (fields T [16, column-major], iterations, converged, score, trans_probability, n_evals)."""
        it = np.ascontiguousarray(item_target, dtype=np.int32)
        isrc = np.ascontiguousarray(item_source, dtype=np.int32)
        n = it.shape[0]
        G = np.ascontiguousarray(np.asarray(guesses, dtype=np.float32).reshape(n, 4, 4).transpose(0, 2, 1)).reshape(n, 16)
        res = np.zeros(n, dtype=RESULT_DTYPE)
        self._ck(self._L.ndtb_batch_run(self._h, n, _p(it), _p(isrc), _p(G), _p(res)))
        return res

    def run(self, item_target, item_source, guesses):
        res = self.run_raw(item_target, item_source, guesses)
        T = res["T"].reshape(-1, 4, 4).transpose(0, 2, 1)
        return [dict(T=T[i].copy(), iterations=int(res["iterations"][i]), converged=bool(res["converged"][i]),
                     score=float(res["score"][i]), trans_probability=float(res["trans_probability"][i]),
                     n_evals=int(res["n_evals"][i])) for i in range(res.shape[0])]

    def fitness(self, n_items, max_range=float(np.finfo(np.float64).max)):
        """getFitnessScore of every registration of the last run (loop-closure gate, pgo_node.cpp:460-476)."""
        out = np.zeros(n_items, dtype=np.float64)
        self._ck(self._L.ndtb_batch_fitness(self._h, float(max_range), _p(out)))
        return out

    # ---- loop-closure submaps on the device (pgo_node.cpp:530-554) ----
    def add_keyframe(self, cloud):
        """Stores a keyframe cloud (sensor frame; column 3 = intensity when present) in HBM; returns its id."""
        cloud = _cloud(cloud)
        i = C.c_int()
        self._ck(self._L.ndtb_batch_add_keyframe(self._h, _p(cloud), cloud.shape[0], cloud.strides[0], C.byref(i)))
        return i.value

    def clear_keyframes(self):
        self._ck(self._L.ndtb_batch_clear_keyframes(self._h))

    def num_keyframes(self):
        return self._L.ndtb_batch_num_keyframes(self._h)

    def add_submap(self, root, half_window, poses, leaf, as_target=True):
        """LoopFindNearKeyframesCloud: keyframes root - half_window .. root + half_window through their poses ((n, 4, 4)
        float32, one per keyframe id), stacked, VoxelGrid(leaf).  Returns (target or source id, points)."""
        P = np.ascontiguousarray(np.asarray(poses, dtype=np.float32).reshape(-1, 4, 4).transpose(0, 2, 1)).reshape(-1, 16)
        i, n = C.c_int(), C.c_size_t()
        self._ck(self._L.ndtb_batch_add_submap(self._h, int(bool(as_target)), int(root), int(half_window), _p(P), P.shape[0], float(leaf),
                                               C.byref(i), C.byref(n)))
        return i.value, n.value

    def add_submaps(self, roots, half_windows, poses, leaf, as_target):
        """Many submaps in one call, assembled concurrently on the batch's build lanes.  roots / half_windows / as_target:
        sequences of equal length.  Returns (ids, points) as int arrays."""
        P = np.ascontiguousarray(np.asarray(poses, dtype=np.float32).reshape(-1, 4, 4).transpose(0, 2, 1)).reshape(-1, 16)
        r = np.ascontiguousarray(roots, dtype=np.int32)
        hw = np.ascontiguousarray(half_windows, dtype=np.int32)
        at = np.ascontiguousarray([1 if a else 0 for a in as_target], dtype=np.int32)
        ids = np.zeros(r.shape[0], dtype=np.int32)
        n = np.zeros(r.shape[0], dtype=np.uint64)
        self._ck(self._L.ndtb_batch_add_submaps(self._h, r.shape[0], _p(at), _p(r), _p(hw), _p(P), P.shape[0], float(leaf), _p(ids), _p(n)))
        return ids, n.astype(np.int64)

    def download(self, id, n, is_target=True):
        """(n, 4) float32 x y z intensity of a batch target / source."""
        out = np.zeros((n, 4), dtype=np.float32)
        m = C.c_size_t()
        self._ck(self._L.ndtb_batch_download(self._h, int(bool(is_target)), int(id), _p(out), n, 16, C.byref(m)))
        return out[:m.value]

    def diag(self, n_items):
        """Per item of the last run: replayed (hybrid mode ran it again with the strict kernel), borderline, n_rise, last_norm."""
        rp, bl, nr = (np.zeros(n_items, np.int32) for _ in range(3))
        ln = np.zeros(n_items, np.float64)
        self._ck(self._L.ndtb_batch_get_diag(self._h, _p(rp), _p(bl), _p(nr), _p(ln)))
        return dict(replayed=rp, borderline=bl, n_rise=nr, last_norm=ln)

    def step_history(self, n_items, max_steps=32):
        """|delta_p| of every Newton solve of every item's final pass, (n_items, max_steps) float32 (0 beyond the last iteration)."""
        out = np.zeros((n_items, max_steps), np.float32)
        self._ck(self._L.ndtb_batch_get_step_history(self._h, _p(out), max_steps))
        return out

    def stats(self):
        s = BatchStats()
        self._ck(self._L.ndtb_batch_get_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in BatchStats._fields_}


class MultiNDT:
    """One batch job cut over several GPUs of one box inside the library (ndtb_multi_*): one host thread and one
    ndtb_batch per listed device, items dealt in contiguous blocks by target id, results gathered in item order."""

    def __init__(self, devices):
        self._L = load_library()
        self._h = C.c_void_p()
        self.devices = [int(d) for d in devices]
        dv = (C.c_int * len(self.devices))(*self.devices)
        if self._L.ndtb_multi_create(len(self.devices), dv, C.byref(self._h)):
            raise NdtError(self._L.ndtb_multi_last_error(None).decode())
        self.params = Params()
        self._L.ndtb_default_params(C.byref(self.params))

    def __del__(self):
        try:
            if self._h:
                self._L.ndtb_multi_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise NdtError(self._L.ndtb_multi_last_error(self._h).decode())

    def set_params(self, **kw):
        for k, v in kw.items():
            if not hasattr(self.params, k):
                raise KeyError(k)
            setattr(self.params, k, v)
        self._ck(self._L.ndtb_multi_set_params(self._h, C.byref(self.params)))

    def run_raw(self, targets, sources, item_target, item_source, guesses):
        """targets / sources: lists of (n, 4) float32 host arrays (kept alive by the caller for the call)."""
        tg = [_cloud(t) for t in targets]
        sc = [_cloud(s) for s in sources]
        stride = tg[0].strides[0] if tg else 16
        assert all(a.strides[0] == stride for a in tg + sc)
        tp, tn, _ = BatchNDT.pointer_table([a.ctypes.data for a in tg], [a.shape[0] for a in tg])
        sp, sn, _ = BatchNDT.pointer_table([a.ctypes.data for a in sc], [a.shape[0] for a in sc])
        it = np.ascontiguousarray(item_target, dtype=np.int32)
        isrc = np.ascontiguousarray(item_source, dtype=np.int32)
        n = it.shape[0]
        G = np.ascontiguousarray(np.asarray(guesses, dtype=np.float32).reshape(n, 4, 4).transpose(0, 2, 1)).reshape(n, 16)
        res = np.zeros(n, dtype=RESULT_DTYPE)
        self._ck(self._L.ndtb_multi_run(self._h, len(tg), tp, tn, len(sc), sp, sn, stride, n, _p(it), _p(isrc), _p(G), _p(res)))
        return res

    def stats(self):
        d = len(self.devices)
        items, targets = np.zeros(d, np.int32), np.zeros(d, np.int32)
        dev_ms, wall_ms = np.zeros(d), np.zeros(d)
        self._ck(self._L.ndtb_multi_get_stats(self._h, _p(items), _p(targets), _p(dev_ms), _p(wall_ms)))
        return dict(items=items, targets=targets, device_ms=dev_ms, wall_ms=wall_ms)
