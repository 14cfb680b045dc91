"""Experiment: stage timings of the e2e upload / compute pipeline with 1 and 2 batches in flight."""
import sys, os, time, threading
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xchu_slam_b200 import ndt, synth
nsub = int(sys.argv[1]) if len(sys.argv) > 1 else 64
w = synth.batch_workload(n_submaps=nsub, scans_per_submap=8, hyps_per_scan=8, cache_dir="/tmp/ndt_b200_bench_cache")
tg = [torch.from_numpy(t).pin_memory() for t in w["targets"]]
sc = [torch.from_numpy(s).pin_memory() for s in w["sources"]]
it, isrc, G = w["item_target"], w["item_source"], w["guesses"]
def mk():
    b = ndt.BatchNDT(); b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30); return b
bs = [mk(), mk()]
ul, cl = threading.Lock(), threading.Lock()
T0 = time.perf_counter()
log = []
def step(b, name, pipe):
    t = [time.perf_counter()]
    if pipe: ul.acquire()
    t.append(time.perf_counter())
    b.clear()
    for x in tg: b.add_target_ptr(x.data_ptr(), x.shape[0], 16, 0)
    for x in sc: b.add_source_ptr(x.data_ptr(), x.shape[0], 16, 0)
    t.append(time.perf_counter())
    if pipe:
        b.wait_uploads(); ul.release()
    t.append(time.perf_counter())
    if pipe: cl.acquire()
    t.append(time.perf_counter())
    b.run_raw(it, isrc, G)
    t.append(time.perf_counter())
    if pipe: cl.release()
    log.append((name, [round((x - T0) * 1e3, 1) for x in t], b.stats()["total_ms"]))
for b in bs: step(b, "warm", False); step(b, "warm", False)
log.clear(); T0 = time.perf_counter()
for k in range(3): step(bs[0], "single%d" % k, False)
for l in log: print(l)
log.clear(); T0 = time.perf_counter()
def worker(j):
    for k in range(4): step(bs[j], "t%d_s%d" % (j, k), True)
th = [threading.Thread(target=worker, args=(j,)) for j in range(2)]
[t.start() for t in th]; [t.join() for t in th]
print("pipelined total", (time.perf_counter() - T0) * 1e3 / 8, "ms/step")
for l in sorted(log, key=lambda x: x[1][0]): print(l)
