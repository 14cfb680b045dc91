"""Experiment: where one odom_node-style scan spends its time through the C ABI (host wall clock per call)."""
import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xchu_slam_b200 import ndt, synth
src, tgt, gt, guess = synth.make_case(n_source=20000, n_target=30000, seed=synth.BASE_SEED, keyframe_leaf=1.0)
def xyzi(a):    # pcl::PointXYZI layout, 32 B / point
    out = np.zeros((a.shape[0], 8), np.float32); out[:, :3] = a[:, :3]; out[:, 3] = 1; return out
src32, tgt32 = xyzi(src), xyzi(tgt)
extra = xyzi(synth.make_case(n_source=4000, n_target=30000, seed=synth.BASE_SEED + 9, keyframe_leaf=1.0)[0])
g = ndt.NormalDistributionsTransform()
g.setResolution(1.0); g.setTransformationEpsilon(0.01); g.setMaximumIterations(30)
L, h = g._L, g._h
r = ndt.Result(); G = np.ascontiguousarray(guess.T).reshape(-1).copy()
out = np.zeros((20000, 8), np.float32); fit = C.c_double(); inc = C.c_int()
T = {}
def tm(name, f):
    t0 = time.perf_counter(); f(); T.setdefault(name, []).append((time.perf_counter() - t0) * 1e3)
for rep in range(60):
    grown = np.concatenate([tgt32, extra[: 3000 + 10 * rep]])
    tm("set_target full 30k (stride 32)", lambda: L.ndtb_set_target(h, tgt32.ctypes.data, tgt32.shape[0], 32))
    tm("set_target_auto +3k (incremental)", lambda: L.ndtb_set_target_auto(h, grown.ctypes.data, grown.shape[0], 32, C.byref(inc)))
    tm("set_target full 33k (stride 32)", lambda: L.ndtb_set_target(h, grown.ctypes.data, grown.shape[0], 32))
    tm("set_source 20k (stride 32)", lambda: L.ndtb_set_source(h, src32.ctypes.data, 20000, 32))
    tm("set_source 20k (stride 16)", lambda: L.ndtb_set_source(h, src.ctypes.data, 20000, 16))
    tm("align", lambda: L.ndtb_align(h, G.ctypes.data, C.byref(r)))
    tm("get_output (stride 32)", lambda: L.ndtb_get_output(h, out.ctypes.data, 32))
    tm("fitness (first after new target)", lambda: L.ndtb_fitness(h, C.c_double(1e300), None, C.byref(fit)))
    tm("fitness (again)", lambda: L.ndtb_fitness(h, C.c_double(1e300), None, C.byref(fit)))
print("iterations", r.iterations, "n_evals", r.n_evals, "incremental", inc.value)
for k, v in T.items():
    v = np.array(v[10:]); print("%-40s median %.3f ms  p10 %.3f  p90 %.3f" % (k, np.median(v), np.percentile(v, 10), np.percentile(v, 90)))
print(g.lastTiming())
