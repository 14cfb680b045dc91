"""ctypes bindings for the ScanContext checker:

  OracleSC -- oracle/sc_oracle.cpp (libsc_oracle.so), the CPU restatement of the reference's
              include/scancontext/Scancontext.cpp;
  RefSC    -- oracle/_ref/libsc_ref.so, the reference's OWN Scancontext.cpp compiled unmodified over the stand-in
              Eigen / PCL headers (oracle/ref_sc_shim.cpp), built only where /root/reference exists.

TEST INFRASTRUCTURE ONLY.  Both expose the same methods; descriptors are (num_ring, num_sector) float64 arrays.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libsc_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libsc_ref.so")
REFERENCE_SRC = "/root/reference/xchu_mapping/include/scancontext/Scancontext.cpp"
R, S = 20, 60     # Scancontext.h:85-86

_lib = _ref = None
vp, sz = C.c_void_p, C.c_size_t


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-C", _HERE, "-s", "libsc_oracle.so"])
        L = C.CDLL(_LIB)
        L.sco_create.restype = vp
        L.sco_destroy.argtypes = [vp]
        L.sco_set_dist_thres.argtypes = [vp, C.c_double]
        L.sco_xy2theta.argtypes = [C.c_float, C.c_float]
        L.sco_xy2theta.restype = C.c_float
        L.sco_make.argtypes = [vp, vp, sz, sz, vp]
        L.sco_keys.argtypes = [vp, vp, vp, vp]
        L.sco_dist_direct.argtypes = [vp, vp, vp]
        L.sco_dist_direct.restype = C.c_double
        L.sco_fast_align.argtypes = [vp, vp, vp]
        L.sco_distance.argtypes = [vp, vp, vp, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.sco_add.argtypes = [vp, vp, sz, sz]
        L.sco_size.argtypes = [vp]
        L.sco_size.restype = sz
        L.sco_get.argtypes = [vp, sz, vp, vp, vp, vp]
        L.sco_detect.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        _lib = L
    return _lib


def ref_available():
    if os.path.exists(REFERENCE_SRC):
        subprocess.check_call(["make", "-C", _HERE, "-s", "_ref/libsc_ref.so"])
    return os.path.exists(_REF)


def ref():
    global _ref
    if _ref is None:
        if not ref_available():
            raise RuntimeError("oracle/_ref/libsc_ref.so is absent and /root/reference is not here to build it")
        L = C.CDLL(_REF)
        L.ref_sc_create.restype = vp
        L.ref_sc_destroy.argtypes = [vp]
        L.ref_sc_constants.argtypes = [vp] + [vp] * 9
        L.ref_sc_make.argtypes = [vp, vp, sz, sz, vp]
        L.ref_sc_keys.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp]
        L.ref_sc_dist_direct.argtypes = [vp, vp, vp, C.c_int, C.c_int]
        L.ref_sc_dist_direct.restype = C.c_double
        L.ref_sc_fast_align.argtypes = [vp, vp, vp, C.c_int]
        L.ref_sc_distance.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]
        L.ref_sc_add.argtypes = [vp, vp, sz, sz]
        L.ref_sc_detect.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_float)]
        L.ref_sc_size.argtypes = [vp]
        L.ref_sc_size.restype = sz
        L.ref_sc_get.argtypes = [vp, sz, vp, vp, vp, vp]
        L.ref_sc_xy2theta.argtypes = [C.c_float, C.c_float]
        L.ref_sc_xy2theta.restype = C.c_float
        _ref = L
    return _ref


def _cloud(pts):
    a = np.ascontiguousarray(pts, dtype=np.float32)
    assert a.ndim == 2 and a.shape[1] >= 3
    return a


def _cm(desc):     # (R, S) row-major view -> column-major buffer as Eigen stores MatrixXd
    return np.ascontiguousarray(np.asarray(desc, dtype=np.float64).T).reshape(-1)


def _from_cm(buf):
    return np.ascontiguousarray(buf.reshape(S, R).T)


class _Base:
    def get(self, i):
        d, rk, sk, ik = np.zeros(R * S), np.zeros(R), np.zeros(S), np.zeros(R, np.float32)
        self._get(i, d, rk, sk, ik)
        return dict(desc=_from_cm(d), ring=rk, sector=sk, invkey=ik)


class OracleSC(_Base):
    def __init__(self):
        self.L = lib()
        self.h = vp(self.L.sco_create())

    def __del__(self):
        try:
            self.L.sco_destroy(self.h)
        except Exception:
            pass

    def set_dist_thres(self, t):
        self.L.sco_set_dist_thres(self.h, float(t))

    def xy2theta(self, x, y):
        return self.L.sco_xy2theta(float(x), float(y))

    def make(self, pts):
        a = _cloud(pts)
        d = np.zeros(R * S)
        self.L.sco_make(self.h, _p(a), a.shape[0], a.shape[1], _p(d))
        return _from_cm(d)

    def keys(self, desc):
        rk, sk = np.zeros(R), np.zeros(S)
        d = _cm(desc)
        self.L.sco_keys(self.h, _p(d), _p(rk), _p(sk))
        return rk, sk

    def dist_direct(self, d1, d2):
        a, b = _cm(d1), _cm(d2)
        return self.L.sco_dist_direct(self.h, _p(a), _p(b))

    def fast_align(self, k1, k2):
        a, b = np.ascontiguousarray(k1, dtype=np.float64), np.ascontiguousarray(k2, dtype=np.float64)
        return self.L.sco_fast_align(self.h, _p(a), _p(b))

    def distance(self, d1, d2):
        a, b = _cm(d1), _cm(d2)
        dist, shift = C.c_double(), C.c_int()
        self.L.sco_distance(self.h, _p(a), _p(b), C.byref(dist), C.byref(shift))
        return dist.value, shift.value

    def add(self, pts):
        a = _cloud(pts)
        self.L.sco_add(self.h, _p(a), a.shape[0], a.shape[1])

    def size(self):
        return self.L.sco_size(self.h)

    def _get(self, i, d, rk, sk, ik):
        self.L.sco_get(self.h, i, _p(d), _p(rk), _p(sk), _p(ik))

    def detect(self, full=False):
        lid, yaw, md, nn, ts = C.c_int(), C.c_float(), C.c_double(), C.c_int(), C.c_int()
        self.L.sco_detect(self.h, C.byref(lid), C.byref(yaw), C.byref(md), C.byref(nn), C.byref(ts))
        if full:
            return dict(loop_id=lid.value, yaw=yaw.value, min_dist=md.value, nn_idx=nn.value, tree_size=ts.value)
        return lid.value, yaw.value


class RefSC(_Base):
    def __init__(self):
        self.L = ref()
        self.h = vp(self.L.ref_sc_create())

    def __del__(self):
        try:
            self.L.ref_sc_destroy(self.h)
        except Exception:
            pass

    def constants(self):
        ring, sector, ex, cand, per = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        mr, lh, sr, dt = C.c_double(), C.c_double(), C.c_double(), C.c_double()
        self.L.ref_sc_constants(self.h, C.byref(ring), C.byref(sector), C.byref(mr), C.byref(lh), C.byref(sr), C.byref(ex), C.byref(cand),
                                C.byref(dt), C.byref(per))
        return dict(num_ring=ring.value, num_sector=sector.value, max_radius=mr.value, lidar_height=lh.value, search_ratio=sr.value,
                    exclude_recent=ex.value, candidates=cand.value, dist_thres=dt.value, tree_period=per.value)

    def xy2theta(self, x, y):
        return self.L.ref_sc_xy2theta(float(x), float(y))

    def make(self, pts):
        a = _cloud(pts)
        d = np.zeros(R * S)
        self.L.ref_sc_make(self.h, _p(a), a.shape[0], a.shape[1], _p(d))
        return _from_cm(d)

    def keys(self, desc):
        rk, sk = np.zeros(R), np.zeros(S)
        d = _cm(desc)
        self.L.ref_sc_keys(self.h, _p(d), R, S, _p(rk), _p(sk))
        return rk, sk

    def dist_direct(self, d1, d2):
        a, b = _cm(d1), _cm(d2)
        return self.L.ref_sc_dist_direct(self.h, _p(a), _p(b), R, S)

    def fast_align(self, k1, k2):
        a, b = np.ascontiguousarray(k1, dtype=np.float64), np.ascontiguousarray(k2, dtype=np.float64)
        return self.L.ref_sc_fast_align(self.h, _p(a), _p(b), S)

    def distance(self, d1, d2):
        a, b = _cm(d1), _cm(d2)
        dist, shift = C.c_double(), C.c_int()
        self.L.ref_sc_distance(self.h, _p(a), _p(b), R, S, C.byref(dist), C.byref(shift))
        return dist.value, shift.value

    def add(self, pts):
        a = _cloud(pts)
        self.L.ref_sc_add(self.h, _p(a), a.shape[0], a.shape[1])

    def size(self):
        return self.L.ref_sc_size(self.h)

    def _get(self, i, d, rk, sk, ik):
        self.L.ref_sc_get(self.h, i, _p(d), _p(rk), _p(sk), _p(ik))

    def detect(self):
        lid, yaw = C.c_int(), C.c_float()
        self.L.ref_sc_detect(self.h, C.byref(lid), C.byref(yaw))
        return lid.value, yaw.value
