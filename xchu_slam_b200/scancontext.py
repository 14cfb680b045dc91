"""Host-side mirror of the reference's SCManager (xchu_mapping/include/scancontext/Scancontext.h:56-121, Scancontext.cpp = SC)
over the device implementation (include/ndt_b200/sc_c_api.h, csrc/scancontext.cu): same member names, argument meaning and
results, plus batched entry points (many scans / pairs / queries per call) that are the point of running it on the GPU.
SURVEY.md section 8(f) rank 4.  No CPU fallback: the CUDA library must be present.

The descriptors, ring keys and sector keys live in HBM; the host keeps only the reference's tree bookkeeping
(`tree_making_period_conter`, the size of the key set the kd-tree was last built over: SC:350-363) -- the search itself is an
exact k-nearest scan of the ring keys on the device, which is what the reference's kd-tree (eps 0) returns.
"""
import ctypes as C

import numpy as np

from .ndt import NdtError, load_library

SYMBOLS = ["ndtb_sc_default_params", "ndtb_sc_create", "ndtb_sc_destroy", "ndtb_sc_last_error", "ndtb_sc_set_dist_thres",
           "ndtb_sc_add_scans", "ndtb_sc_size", "ndtb_sc_clear", "ndtb_sc_get", "ndtb_sc_distance", "ndtb_sc_detect"]


class SCParams(C.Structure):
    _fields_ = [("num_ring", C.c_int), ("num_sector", C.c_int), ("max_radius", C.c_double), ("lidar_height", C.c_double),
                ("search_ratio", C.c_double), ("num_exclude_recent", C.c_int), ("num_candidates", C.c_int),
                ("tree_making_period", C.c_int), ("dist_thres", C.c_double)]


_bound = False


def _lib():
    global _bound
    L = load_library()
    if not _bound:
        vp, sz = C.c_void_p, C.c_size_t
        L.ndtb_sc_default_params.argtypes = [C.POINTER(SCParams)]
        L.ndtb_sc_default_params.restype = None
        L.ndtb_sc_create.argtypes = [C.c_int, C.POINTER(SCParams), C.POINTER(vp)]
        L.ndtb_sc_destroy.argtypes = [vp]
        L.ndtb_sc_last_error.argtypes = [vp]
        L.ndtb_sc_last_error.restype = C.c_char_p
        L.ndtb_sc_set_dist_thres.argtypes = [vp, C.c_double]
        L.ndtb_sc_add_scans.argtypes = [vp, C.c_int, vp, vp, sz, C.POINTER(C.c_int)]
        L.ndtb_sc_size.argtypes = [vp]
        L.ndtb_sc_clear.argtypes = [vp]
        L.ndtb_sc_get.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.ndtb_sc_distance.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.ndtb_sc_detect.argtypes = [vp, C.c_int, vp, vp, vp, vp, vp, vp]
        _bound = True
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class SCManager:
    """SCManager of the reference, descriptors resident on one device."""

    def __init__(self, device=0, **params):
        self._L = _lib()
        self.params = SCParams()
        self._L.ndtb_sc_default_params(C.byref(self.params))
        for k, v in params.items():
            if not hasattr(self.params, k):
                raise KeyError(k)
            setattr(self.params, k, v)
        self._h = C.c_void_p()
        if self._L.ndtb_sc_create(device, C.byref(self.params), C.byref(self._h)):
            raise NdtError(self._L.ndtb_sc_last_error(None).decode())
        self.tree_making_period_conter = 0      # sic (Scancontext.h:83)
        self._tree_size = 0                     # keys in the tree built last: polarcontext_invkeys_to_search_.size()

    def __del__(self):
        try:
            if self._h:
                self._L.ndtb_sc_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise NdtError(self._L.ndtb_sc_last_error(self._h).decode())

    # ---- reference members ----
    def setSCdistThres(self, thres):
        self.params.dist_thres = float(thres)
        self._ck(self._L.ndtb_sc_set_dist_thres(self._h, float(thres)))

    def makeAndSaveScancontextAndKeys(self, scan_down):
        """SC:249-260 for one scan ((n, >=3) float32, sensor frame)."""
        return self.add_scans([scan_down])

    def detectLoopClosureID(self):
        """SC:331-423 for the descriptor added last: (loop id or -1, yaw difference in radians as a float)."""
        n = self.size()
        if n < self.params.num_exclude_recent + 1:                                           # SC:343-347
            return -1, 0.0
        if self.tree_making_period_conter % self.params.tree_making_period == 0:             # SC:350-361
            self._tree_size = n - self.params.num_exclude_recent
        self.tree_making_period_conter += 1
        r = self.detect([n - 1], [self._tree_size])
        return int(r["loop_id"][0]), float(r["yaw_diff_rad"][0])

    def getConstRefRecentSCD(self):
        return self.get(self.size() - 1)["desc"]

    def distanceBtnScanContext(self, id1, id2):
        """SC:118-148 between two stored descriptors: (distance, argmin column shift)."""
        d, s = self.distance([id1], [id2])
        return float(d[0]), int(s[0])

    # ---- batched entry points ----
    def add_scans(self, scans):
        """Descriptors + keys of many scans in one launch; returns the id of the first."""
        arrs = [np.ascontiguousarray(s, dtype=np.float32) for s in scans]
        for a in arrs:
            if a.ndim != 2 or a.shape[1] < 3:
                raise ValueError("scan must be (n, >=3) float32")
        if not arrs:
            return self.size()
        stride = arrs[0].strides[0] if arrs[0].shape[0] else 4 * arrs[0].shape[1]
        for a in arrs:
            if a.shape[0] and a.strides[0] != stride:
                raise ValueError("all scans of a call must share one record stride")
        ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
        ns = (C.c_size_t * len(arrs))(*[a.shape[0] for a in arrs])
        first = C.c_int()
        self._ck(self._L.ndtb_sc_add_scans(self._h, len(arrs), ptrs, ns, stride, C.byref(first)))
        return first.value

    def size(self):
        return self._L.ndtb_sc_size(self._h)

    def clear(self):
        self._ck(self._L.ndtb_sc_clear(self._h))
        self.tree_making_period_conter = 0
        self._tree_size = 0

    def get(self, i):
        R, S = self.params.num_ring, self.params.num_sector
        d, rk, sk, ik = np.zeros(R * S), np.zeros(R), np.zeros(S), np.zeros(R, np.float32)
        self._ck(self._L.ndtb_sc_get(self._h, int(i), _p(d), _p(rk), _p(sk), _p(ik)))
        return dict(desc=np.ascontiguousarray(d.reshape(S, R).T), ring=rk, sector=sk, invkey=ik)

    def distance(self, id1, id2):
        a, b = np.ascontiguousarray(id1, dtype=np.int32), np.ascontiguousarray(id2, dtype=np.int32)
        dist, shift = np.zeros(a.shape[0]), np.zeros(a.shape[0], np.int32)
        self._ck(self._L.ndtb_sc_distance(self._h, a.shape[0], _p(a), _p(b), _p(dist), _p(shift)))
        return dist, shift

    def detect(self, query_ids, tree_sizes):
        q, t = np.ascontiguousarray(query_ids, dtype=np.int32), np.ascontiguousarray(tree_sizes, dtype=np.int32)
        n = q.shape[0]
        out = dict(loop_id=np.zeros(n, np.int32), yaw_diff_rad=np.zeros(n, np.float32), min_dist=np.zeros(n), nn_idx=np.zeros(n, np.int32))
        self._ck(self._L.ndtb_sc_detect(self._h, n, _p(q), _p(t), _p(out["loop_id"]), _p(out["yaw_diff_rad"]), _p(out["min_dist"]),
                                        _p(out["nn_idx"])))
        return out
