#!/usr/bin/env python
"""bench.py -- NDT aligns/sec (20k-pt scan vs 1M-pt map) on N B200s; % HBM roofline.

Default workload = BASELINE configs[4] (C5): ONE job of S distinct 1M-pt submaps x (scans x hypotheses) independent aligns of
20k-pt scans (res 1.0 m, DIRECT7, step 0.1, eps 0.01, <= 30 iterations) -- 64 x 8 x 8 = 4096 registrations -- INCLUDING the
voxel-table build of every submap.  At N GPUs the job is SHARDED (shard.plan_shards: items ordered by submap, contiguous
blocks; a rank uploads and builds only its own submaps) and the per-item results come back through one NCCL all_gather per
step inside the timed region: STRONG scaling, value = 4096 x steps / max-over-ranks device time.  The same line carries, under
"weak", the figure for N copies of the whole job (per-GPU work fixed), which is what round 1 reported.

  value : inputs resident in HBM when the timed region starts (device pointers handed to the C ABI)
  e2e   : same job through the C ABI with HOST (pinned) buffers; H2D of every cloud and D2H of the results inside the timed region
  value_hybrid / value_factorised_only : the same resident job with strict_math -2 / 0 instead of the default (term-by-term kernel)
  --impl reference : the reference algorithm's CPU restatement (oracle/, OpenMP on all host cores) on a bounded sample
  --workload c1|c3|c4 : single-align latency lines (build / align / end to end) with the derivative kernel's roofline where ONE
          registration owns the voxel table; --workload c2 : scans/s of the odometry sequence, both refresh strategies
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "NDT aligns/sec (20k-pt scan vs 1M-pt map)"
UNIT = "aligns/s"
CACHE = os.environ.get("NDT_BENCH_CACHE", "/tmp/ndt_b200_bench_cache")
DTYPE = "f32 terms, f64 accumulate"
NDT_PARAMS = dict(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c5", choices=["c5", "c1", "c2", "c3", "c4"])
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong: the 4096-align job is cut over the GPUs (BASELINE config C5); weak: every GPU runs the whole job")
    ap.add_argument("--submaps", type=int, default=int(os.environ.get("NDT_BENCH_SUBMAPS", 64)))
    ap.add_argument("--scans", type=int, default=8, help="distinct scans per submap")
    ap.add_argument("--hyps", type=int, default=8, help="initial-guess hypotheses per scan")
    ap.add_argument("--n-target", type=int, default=1000000)
    ap.add_argument("--n-source", type=int, default=20000)
    ap.add_argument("--cpu-aligns", type=int, default=1024, help="aligns in the bounded CPU sample (~10 ms each on 16 cores)")
    ap.add_argument("--cpu-passes", type=int, default=4, help="the CPU sample is timed in this many passes: median / min / max")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cpp-pipeline", action="store_true", help="skip the examples/batch_pipeline.cpp arm of the e2e number")
    ap.add_argument("--no-variants", action="store_true", help="skip value_strict / value_factorised_only / the other scaling mode")
    ap.add_argument("--strict-math", type=int, default=int(os.environ.get("NDTB_BENCH_STRICT", -1)),
                    help="-1 library default (term-by-term kernel: the reference arithmetic by construction), -2 hybrid (factorised kernel + strict replay of flagged registrations), 0 factorised kernel only, 1 term-by-term kernel by name")
    ap.add_argument("--inflight", type=int, default=1,
                    help="batches in flight per GPU for `value` (host threads, one BatchNDT and stream each)")
    ap.add_argument("--inflight-e2e", type=int, default=2,
                    help="batches in flight per GPU for `e2e`: the H2D of one step overlaps the compute of another")
    ap.add_argument("--c2-scans", type=int, default=300)
    return ap.parse_args()


def workload_config(a, n_gpus, scaling):
    n = a.submaps * a.scans * a.hyps
    per = "the job is sharded over the GPUs, %d aligns per step in total" % n if scaling == "strong" else "%d aligns per GPU per step" % n
    return {"workload": "C5 batched registration: %d submaps x %d scans x %d hypotheses = %d aligns (%s), %d-pt scan vs %d-pt submap, "
                        "res 1.0 m, DIRECT7, step 0.1, eps 0.01, max_iter 30, voxel-table build of every submap included"
                        % (a.submaps, a.scans, a.hyps, n, per, a.n_source, a.n_target),
            "aligns_per_step": n if scaling == "strong" else n * n_gpus, "n_source": a.n_source, "n_target": a.n_target,
            "resolution": 1.0, "search": "DIRECT7",
            "parallelism": ("one job sharded by submap over %d GPU(s), no collective inside a registration, one all_gather of the results per step"
                            if scaling == "strong" else "independent copies of the job on %d GPU(s), no collective") % n_gpus,
            "l2_policy": "inputs_exceed_l2 (%.2f GB of clouds per job vs 126 MB L2)"
                         % ((a.submaps * a.n_target + a.submaps * a.scans * a.n_source) * 16 / 1e9)}


def get_workload(a):
    from xchu_slam_b200 import synth
    return synth.batch_workload(n_submaps=a.submaps, scans_per_submap=a.scans, hyps_per_scan=a.hyps, n_target=a.n_target,
                                n_source=a.n_source, cache_dir=CACHE)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for k, nme in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def load_ncu_constants():
    """Per-launch figures of the derivative kernel taken from ONE ncu --set full capture (profiles/roofline_traffic.json): DRAM
    bytes per launch and warp instructions per 32 source points.  They are properties of the kernel binary and the workload
    shape, not of this run; the times they are combined with are measured live."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
    except Exception:
        return {}


# ---------------------------------------------------------------------------------------------------------------------
# CPU legs (the oracle: test infrastructure, executed here only as the timed baseline and as the checker of the GPU results)
# ---------------------------------------------------------------------------------------------------------------------
def cpu_sample(a, w, n_aligns, passes=1, with_one_thread=True):
    """Oracle (CPU restatement of ndt_omp) on a bounded sample of the workload, BASELINE.md section 2 protocol: built on this
    machine with -O3 -march=native -fopenmp -ffp-contract=off, OpenMP on ALL host cores, `passes` passes over disjoint parts
    of the sample (median / min / max of the per-pass rate); each submap's grid built once, serially, as the reference does,
    and charged at the workload's one build per scans*hyps aligns; the M-point kd-tree pcl::Registration::align builds for a
    new target timed beside it; plus the as-shipped ONE-thread figure (odom_node.cpp:74) on a small sub-sample.
    Returns (cpu_baseline dict, {item: result}) -- the per-item results are what the GPU batch is checked against."""
    from oracle import oracle
    nthreads = len(os.sched_getaffinity(0))   # all host cores, whatever OMP_NUM_THREADS torchrun exported
    o = oracle.OracleNDT(native=True)
    o.set_params(search_method=oracle.DIRECT7, num_threads=nthreads, **NDT_PARAMS)
    per_submap = a.scans * a.hyps
    it = np.asarray(w["item_target"])
    n_aligns = min(n_aligns, len(it))
    subs = []          # (submap, [items]) in sample order
    n = 0
    for sub in range(len(w["targets"])):
        if n >= n_aligns:
            break
        ids = np.flatnonzero(it == sub)[:min(per_submap, n_aligns - n)]
        if len(ids):
            subs.append((sub, ids))
            n += len(ids)
    passes = max(1, min(passes, len(subs)))
    per_pass = [subs[k::passes] for k in range(passes)]
    results, rates, t_build_all, t_align_all, t_kd_all, evals, n_built = {}, [], 0.0, 0.0, 0.0, 0, 0
    wall0 = time.perf_counter()
    for group in per_pass:
        t_build, t_align, n_g = 0.0, 0.0, 0
        for sub, ids in group:
            t0 = time.perf_counter()
            o.set_target(w["targets"][sub])
            t_build += time.perf_counter() - t0
            if n_built < 2:      # the raw-target kd-tree (pcl::Registration::initCompute), timed on the first submaps only
                t0 = time.perf_counter()
                o.build_nn()
                t_kd_all += time.perf_counter() - t0
            n_built += 1
            for i in ids:
                o.set_source(w["sources"][w["item_source"][i]])
                t0 = time.perf_counter()
                r = o.align(w["guesses"][i])
                t_align += time.perf_counter() - t0
                evals += len(r["trace"])
                results[int(i)] = dict(T=r["T"], iterations=r["iterations"], converged=r["converged"])
                n_g += 1
        # a partly used submap still pays its whole build: charge builds at the workload's 1 build : per_submap aligns
        rates.append(n_g / (t_align + (t_build / len(group)) * (n_g / per_submap)))
        t_build_all += t_build; t_align_all += t_align
    wall = time.perf_counter() - wall0
    kd = t_kd_all / min(2, n_built)
    total = t_align_all + (t_build_all / n_built) * (n / per_submap)
    out = {"value": float(np.median(rates)), "unit": UNIT, "cores": nthreads, "kind": "port",
           "passes": {"n": passes, "median": float(np.median(rates)), "min": float(np.min(rates)), "max": float(np.max(rates)),
                      "pooled": n / total},
           "with_target_kdtree": n / (total + kd * (n / per_submap)),
           "build_s": t_build_all / n_built, "kdtree_build_s": kd, "align_ms": 1e3 * t_align_all / n,
           "sample": "%d aligns vs %d submaps in %d passes (%.3f s serial grid build per submap, one build per %d aligns as in the workload; "
                     "%.1f ms/align, %.1f evaluations/align; %.1f s of CPU wall time); oracle = dependency-free restatement of ndt_omp, "
                     "g++ %s built on this host, %d OpenMP threads; `with_target_kdtree` also charges the %.3f s kd-tree over the "
                     "1M target points that pcl::Registration::align builds per new target"
                     % (n, n_built, passes, t_build_all / n_built, per_submap, 1e3 * t_align_all / n, evals / n, wall,
                        oracle.NATIVE_FLAGS if o.native else "-O2 -fopenmp -ffp-contract=off (portable build: the native compile failed)",
                        nthreads, kd)}
    if with_one_thread:
        o1 = oracle.OracleNDT(native=True)
        o1.set_params(search_method=oracle.DIRECT7, num_threads=1, **NDT_PARAMS)
        sub, ids = subs[0]
        o1.set_target(w["targets"][sub])
        t0 = time.perf_counter()
        for i in ids[:8]:
            o1.set_source(w["sources"][w["item_source"][i]])
            o1.align(w["guesses"][i])
        t1 = (time.perf_counter() - t0) / len(ids[:8])
        out["one_thread_as_shipped"] = {"value": 1.0 / (t1 + (t_build_all / n_built) / per_submap), "align_ms": 1e3 * t1, "aligns": int(len(ids[:8])),
                                        "note": "odom_node configures ndt_omp with ONE thread (src/odom_node.cpp:74)"}
    return out, results


MODE_NAMES = {-1: "term-by-term kernel (library default: every float operation of the reference in its order)",
              -2: "hybrid: factorised kernel, then strict replay of flagged registrations inside ndtb_batch_run",
              0: "factorised kernel only", 1: "term-by-term kernel"}


def parity_report(cpu_results, gpu_res, mode):
    """GPU batch vs CPU oracle on the same registrations: north_star's gate, registration by registration."""
    dts, n_eq, n_in, worst = [], 0, 0, None
    for i, r in cpu_results.items():
        Tg = np.asarray(gpu_res["T"][i], dtype=np.float32).reshape(4, 4).T
        dt = float(np.max(np.abs(Tg[:3, 3] - r["T"][:3, 3]))); dr = float(np.max(np.abs(Tg[:3, :3] - r["T"][:3, :3])))
        eq = int(gpu_res["iterations"][i]) == r["iterations"] and bool(gpu_res["converged"][i]) == r["converged"]
        ok = eq and dt < 1e-4 and dr < 1e-5
        n_eq += int(eq); n_in += int(ok)
        dts.append(dt)
        if not ok and (worst is None or dt > worst["translation_diff_m"]):
            worst = {"item": int(i), "iterations_gpu": int(gpu_res["iterations"][i]), "iterations_cpu": r["iterations"],
                     "translation_diff_m": dt, "rotation_diff": dr}
    n = len(cpu_results)
    return {"registrations_compared": n, "iterations_and_convergence_equal": n_eq, "also_final_transform_within_1e-4m_1e-5rad": n_in,
            "all_within_gate": n_in == n, "worst_outside_gate": worst,
            "translation_diff_m": {"median": float(np.median(dts)), "p99": float(np.percentile(dts, 99)), "max": float(np.max(dts))},
            "gpu_mode": mode,
            "note": "GPU batch (the timed path) vs the CPU oracle on the same items of the last timed step; gate = equal iteration "
                    "counts and convergence flags, final transform within 1e-4 m / 1e-5 rad (north_star); the bench exits non-zero "
                    "when any registration is outside it"}


def run_reference(a):
    """Reference arm: the reference algorithm on the host cores.  `value` times the oracle port (the faster of the two CPU
    builds, so the conservative baseline); `reference_build` beside it times oracle/_ref, the reference's own ndt_omp sources
    compiled over stand-in Eigen/PCL headers (the real ones are absent from this image) -- DESIGN.md sections 6 and 7."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from xchu_slam_b200 import synth
    try:
        w = get_workload(a)
    except MemoryError:
        w = synth.batch_workload(n_submaps=1, scans_per_submap=a.scans, hyps_per_scan=a.hyps, n_target=a.n_target, n_source=a.n_source)
    per_step = max(64, a.cpu_aligns // max(1, a.steps))
    for _ in range(max(0, min(a.warmup, 1))):
        cpu_sample(a, w, 2, with_one_thread=False)
    t0 = time.perf_counter()
    vals = [cpu_sample(a, w, per_step, passes=1, with_one_thread=(k == 0))[0] for k in range(a.steps)]
    dt = time.perf_counter() - t0
    rates = [x["value"] for x in vals]
    v = float(np.median(rates))
    cb = dict(vals[0]); cb["value"] = v
    cb["passes"] = {"n": len(rates), "median": v, "min": float(np.min(rates)), "max": float(np.max(rates))}
    rb = None
    try:
        from oracle import ref
        if ref.available():      # the reference's own NDT sources over the stand-in Eigen/PCL headers, same items, a smaller sample
            nthreads = len(os.sched_getaffinity(0))
            r = ref.RefNDT()
            r.set_params(search_method=2, num_threads=nthreads, **NDT_PARAMS)
            t0 = time.perf_counter(); r.set_target(w["targets"][0]); tb = time.perf_counter() - t0
            ids = np.flatnonzero(np.asarray(w["item_target"]) == 0)[:24]
            t0 = time.perf_counter()
            for i in ids:
                r.set_source(w["sources"][w["item_source"][i]])
                r.align(w["guesses"][i])
            ta = time.perf_counter() - t0
            per = a.scans * a.hyps
            rb = {"value": len(ids) / (ta + tb * len(ids) / per), "unit": UNIT, "cores": nthreads, "aligns": int(len(ids)), "build_s": tb,
                  "align_ms": 1e3 * ta / len(ids),
                  "note": "oracle/_ref: the reference's ndt_omp sources compiled against stand-in Eigen/PCL headers whose eager, "
                          "unvectorised evaluation is slower than Eigen's; reported beside the (faster) port, which stays the baseline"}
    except Exception as ex:      # the reference build is optional evidence, never a reason to fail the arm
        rb = {"unavailable": str(ex)[:200]}
    line = {"impl": "reference", "reference_build": rb, "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": DTYPE,
            "data": "synthetic", "config": workload_config(a, a.gpus, a.scaling), "cpu_baseline": cb,
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------
# C5 on the GPUs
# ---------------------------------------------------------------------------------------------------------------------
class Job:
    """A set of registrations and the clouds they need, held twice: pinned on the host (e2e) and resident in HBM (value)."""

    def __init__(self, w, idx, dev):
        import torch
        from xchu_slam_b200 import ndt, shard
        self.idx = np.asarray(idx)
        tids, sids, lt, ls = shard.local_problem(self.idx, w["item_target"], w["item_source"])
        self.n_items = len(self.idx)
        self.it, self.isrc = lt, ls
        self.G = np.ascontiguousarray(np.asarray(w["guesses"], dtype=np.float32)[self.idx])
        self.tg_host = [torch.from_numpy(w["targets"][t]).pin_memory() for t in tids]
        self.sc_host = [torch.from_numpy(w["sources"][s]).pin_memory() for s in sids]
        self.tg_dev = [t.to(dev) for t in self.tg_host]
        self.sc_dev = [s.to(dev) for s in self.sc_host]
        PT = ndt.BatchNDT.pointer_table
        # the clouds stay where they are between steps: their pointer tables are built once, a step hands them over in two calls
        self.tab_tg_host = PT([t.data_ptr() for t in self.tg_host], [t.shape[0] for t in self.tg_host])
        self.tab_sc_host = PT([s.data_ptr() for s in self.sc_host], [s.shape[0] for s in self.sc_host])
        self.tab_tg_dev = PT([t.data_ptr() for t in self.tg_dev], [t.shape[0] for t in self.tg_dev])
        self.tab_sc_dev = PT([s.data_ptr() for s in self.sc_dev], [s.shape[0] for s in self.sc_dev])
        self.h2d_bytes = sum(t.numel() * 4 for t in self.tg_host) + sum(s.numel() * 4 for s in self.sc_host) + self.n_items * 64
        self.d2h_bytes = self.n_items * C.sizeof(ndt.Result)


def run_c5(a):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # host threads building voxel tables per rank (the library's default is 12): share the host's cores between the ranks
    if "NDTB_BUILD_LANES" not in os.environ:
        os.environ["NDTB_BUILD_LANES"] = str(max(4, min(12, (os.cpu_count() or 16) // max(1, world))))

    # ---- rank 0: the workload (generated once, cached) and the CPU sample, BEFORE the process group exists: the other ranks wait
    # in the rendezvous (a TCP store, no GPU spinning in an NCCL barrier, no host cores stolen from the OpenMP threads) ----------
    w, cb, cpu_results = None, None, None
    if rank == 0:
        w = get_workload(a)
        if not a.no_cpu_baseline:
            cb, cpu_results = cpu_sample(a, w, a.cpu_aligns, passes=a.cpu_passes)
    dist = None
    if world > 1:
        import datetime
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(minutes=30))
    if w is None:
        w = get_workload(a)       # cached by rank 0 before it joined the group

    def barrier():
        if dist is not None:
            dist.barrier()

    from xchu_slam_b200 import ndt, shard
    n_total = len(w["item_target"])
    plan = shard.plan_shards(w["item_target"], world)
    width = max(len(p) for p in plan)
    jobs = {"strong": Job(w, plan[rank], dev)}
    jobs["weak"] = jobs["strong"] if world == 1 else None      # built on demand: every rank then holds the whole job
    torch.cuda.synchronize()

    def make_batches(n, strict):
        out = []
        for _ in range(n):
            bk = ndt.BatchNDT(device=local)
            bk.set_params(search_method=ndt.DIRECT7, strict_math=strict, **NDT_PARAMS)
            out.append(bk)
        # batch 0 runs on torch's current stream, so torch's events bracket the library's work directly
        out[0].set_stream(torch.cuda.current_stream().cuda_stream)
        return out

    batches = make_batches(max(1, a.inflight, a.inflight_e2e), a.strict_math)
    upload_lock, compute_lock = threading.Lock(), threading.Lock()
    gather_buf = torch.zeros((width, shard.RESULT_WIDTH), dtype=torch.float64, device=dev)
    gather_out = torch.zeros((world * width, shard.RESULT_WIDTH), dtype=torch.float64, device=dev) if world > 1 else None

    def pack(res):
        out = np.zeros((len(res), shard.RESULT_WIDTH))
        out[:, :16] = res["T"]; out[:, 16] = res["iterations"]; out[:, 17] = res["converged"]; out[:, 18] = res["score"]
        out[:, 19] = res["trans_probability"]; out[:, 20] = res["n_evals"]
        return out

    pin_in = torch.zeros((width, shard.RESULT_WIDTH), dtype=torch.float64).pin_memory() if world > 1 else None
    pin_out = torch.zeros((world * width, shard.RESULT_WIDTH), dtype=torch.float64).pin_memory() if world > 1 else None

    def gather(res):
        """The job's only exchange: every rank's result rows to every rank (21 doubles per registration), over NCCL; staged
        through page-locked buffers, one stream synchronisation at the end."""
        if len(res):
            pin_in.numpy()[:len(res)] = pack(res)
        gather_buf.copy_(pin_in, non_blocking=True)
        dist.all_gather_into_tensor(gather_out, gather_buf)
        pin_out.copy_(gather_out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        full = np.zeros((n_total, shard.RESULT_WIDTH))
        g = pin_out.numpy()
        for r in range(world):
            if len(plan[r]):
                full[plan[r]] = g[r * width:r * width + len(plan[r])]
        return full

    def step(job, resident, b, pipelined=False, do_gather=False):
        """pipelined (several batches in flight): a two-stage pipeline -- one batch uploads while another computes."""
        if job.n_items == 0:
            res = np.zeros(0, dtype=ndt.RESULT_DTYPE)
            if do_gather:
                with compute_lock:
                    return gather(res), None
            return res, None
        if pipelined:
            with upload_lock:
                b.clear()
                b.add_targets_ptr(job.tab_tg_host, 16, 0)
                b.add_sources_ptr(job.tab_sc_host, 16, 0)
                b.wait_uploads()
            with compute_lock:
                res = b.run_raw(job.it, job.isrc, job.G)
                st = b.stats()
                return (gather(res) if do_gather else res), st
        b.clear()
        if resident:
            b.add_targets_ptr(job.tab_tg_dev, 16, 2)      # resident in HBM: borrowed in place
            b.add_sources_ptr(job.tab_sc_dev, 16, 2)
        else:
            b.add_targets_ptr(job.tab_tg_host, 16, 0)     # pinned host memory: H2D inside the call
            b.add_sources_ptr(job.tab_sc_host, 16, 0)
        res = b.run_raw(job.it, job.isrc, job.G)  # returns after the device finished and the result records are on the host
        st = b.stats()
        return (gather(res) if do_gather else res), st

    KEYS = ("deriv_ms", "build_ms", "total_ms", "kernel_launches", "deriv_launches", "points_evaluated", "pairs_evaluated", "replayed",
            "replay_ms", "replay_deriv_ms")

    def timed(job, bset, resident, steps, inflight, do_gather):
        """EXACTLY `steps` steps.  inflight > 1: that many host threads, each driving its own BatchNDT (own streams) and
        taking steps round-robin; the GPU overlaps one step's uploads / tail iterations with another step's compute."""
        barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        agg = {k: 0.0 for k in KEYS}
        out = {}
        lock = threading.Lock()

        def worker(j):
            torch.cuda.set_device(local)
            for k in range(j, steps, inflight):
                res, st = step(job, resident, bset[j], pipelined=(inflight > 1 and not resident), do_gather=do_gather)
                with lock:
                    out[k] = res
                    if st:
                        for key in KEYS:
                            agg[key] += st[key]
        if inflight <= 1:
            worker(0)
        else:
            th = [threading.Thread(target=worker, args=(j,)) for j in range(inflight)]
            for t in th:
                t.start()
            for t in th:
                t.join()
        e1.record()          # every run_raw / gather has returned: all of the steps' device work precedes this record
        torch.cuda.synchronize()
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            tt = torch.tensor([ms] + [agg[k] for k in KEYS], device=dev, dtype=torch.float64)
            mx = tt.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            dist.all_reduce(tt, op=dist.ReduceOp.SUM)
            ms = float(mx[0].item())
            agg = {k: float(tt[1 + i].item()) for i, k in enumerate(KEYS)}      # summed over ranks
            agg["deriv_ms_max_rank"] = float(mx[1].item())
        else:
            agg["deriv_ms_max_rank"] = agg["deriv_ms"]
        return ms, agg, out[steps - 1]

    def measure(job, bset, do_gather, steps, warm, with_e2e=True):
        for _ in range(warm):
            for bk in bset[:max(1, a.inflight)]:
                step(job, True, bk, do_gather=do_gather)
        ms, agg, res = timed(job, bset, True, steps, max(1, a.inflight), do_gather)
        r = {"ms": ms, "agg": agg, "res": res}
        if with_e2e:
            for _ in range(min(warm, 1)):
                for bk in bset[:max(1, a.inflight_e2e)]:
                    step(job, False, bk, do_gather=do_gather)
            r["ms_e2e"], r["agg_e2e"], _ = timed(job, bset, False, steps, max(1, a.inflight_e2e), do_gather)
        return r

    def as_fields(res):
        if isinstance(res, np.ndarray) and res.dtype.names is None:      # gathered rows -> the fields the parity check reads
            return {"T": res[:, :16], "iterations": res[:, 16].astype(int), "converged": res[:, 17].astype(int)}
        return res

    def sample_parity(res, job, mode_name):
        """The registrations of the last timed step that the CPU oracle also ran, against it."""
        res = as_fields(res)
        have = set(int(i) for i in (job.idx if (main_mode == "weak" or world == 1) else range(n_total)))
        cmp_items = {i: r for i, r in cpu_results.items() if i in have}
        if main_mode == "weak" or world == 1:      # results are in the job's local order
            pos = {int(i): k for k, i in enumerate(job.idx)}
            order = sorted(cmp_items)
            res = {"T": np.asarray(res["T"])[[pos[i] for i in order]], "iterations": np.asarray(res["iterations"])[[pos[i] for i in order]],
                   "converged": np.asarray(res["converged"])[[pos[i] for i in order]]}
            cmp_items = {k: cmp_items[i] for k, i in enumerate(order)}
        return parity_report(cmp_items, res, mode_name)

    main_mode = a.scaling
    if main_mode == "weak" and jobs["weak"] is None:
        jobs["weak"] = Job(w, np.arange(n_total), dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    M = measure(jobs[main_mode], batches, main_mode == "strong" and world > 1, a.steps, a.warmup)
    clocks = sampler.stop() if rank == 0 else None
    n_per_step = n_total if main_mode == "strong" else n_total * world

    extra = {}
    if not a.no_variants:
        vsteps = max(2, min(3, a.steps))
        for name, strict in (("value_hybrid", -2), ("value_factorised_only", 0)):
            if strict == a.strict_math:
                continue
            bs = make_batches(1, strict)
            r = measure(jobs[main_mode], bs, main_mode == "strong" and world > 1, vsteps, 1, with_e2e=False)
            extra[name] = {"value": n_per_step * vsteps / (r["ms"] / 1e3), "ms_per_step": r["ms"] / vsteps, "steps": vsteps,
                           "strict_math": strict,
                           "note": ("opt-in hybrid mode: factorised kernel, then a strict replay of the flagged registrations inside the same call; "
                                    "24571 of 24576 registrations of six synthetic worlds inside the gate (profiles/r02/r02_c11_hybrid_seeds.txt)" if strict == -2 else
                                    "factorised kernel only, no strict replay: round 1's headline path; NOT parity-clean "
                                    "(~1 % of registrations end outside 1e-4 m / 1e-5 rad)")}
            if rank == 0 and cb is not None:
                extra[name]["parity_vs_cpu_sample"] = sample_parity(r["res"], jobs[main_mode], MODE_NAMES[strict])
            del bs
        if world > 1:
            other = "weak" if main_mode == "strong" else "strong"
            if jobs[other] is None:
                jobs[other] = Job(w, np.arange(n_total), dev)
            r = measure(jobs[other], batches, other == "strong", a.steps, 1)
            n_other = n_total if other == "strong" else n_total * world
            extra[other] = {"value": n_other * a.steps / (r["ms"] / 1e3), "ms_per_step": r["ms"] / a.steps,
                            "e2e": n_other * a.steps / (r["ms_e2e"] / 1e3), "e2e_ms_per_step": r["ms_e2e"] / a.steps, "aligns_per_step": n_other,
                            "note": "the job cut over the GPUs, results gathered over NCCL" if other == "strong" else
                                    "every GPU runs its own copy of the whole job (per-GPU work fixed)"}

    # ---- the same job through examples/batch_pipeline.cpp: the end-to-end pipeline with no interpreter in the loop, 16-byte and
    # packed 12-byte host records (rank 0, one GPU: it is a second process on this rank's device, run while this one is idle) ----
    if rank == 0 and world == 1 and not a.no_cpp_pipeline:
        extra["e2e_cpp"] = cpp_pipeline(a, w, jobs[main_mode], local)

    rc = 0
    if rank == 0:
        ms, agg = M["ms"], M["agg"]
        value = n_per_step * a.steps / (ms / 1e3)
        e2e = n_per_step * a.steps / (M["ms_e2e"] / 1e3)
        line = c5_line(a, world, main_mode, value, e2e, M, agg, clocks, jobs[main_mode], n_per_step)
        line.update(extra)
        res = as_fields(M["res"])
        conv = int(np.count_nonzero(np.asarray(res["converged"])))
        line["converged_frac"] = conv / max(1, len(res["converged"]))
        line["mean_iterations"] = float(np.mean(res["iterations"]))
        if cb is not None:
            par = sample_parity(M["res"], jobs[main_mode], MODE_NAMES[a.strict_math])
            line["parity_vs_cpu_sample"] = par
            line["cpu_baseline"] = cb
            if not par["all_within_gate"] and a.strict_math != 0:
                rc = 1
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rc:
        sys.stderr.write("bench.py: GPU results outside the parity gate (see parity_vs_cpu_sample)\n")
        sys.exit(rc)


def cpp_pipeline(a, w, job, device):
    import subprocess
    import tempfile
    from xchu_slam_b200 import build, jobfile
    out = {}
    try:
        exe = build.build_batch_pipeline()
        tids, sids = sorted(set(int(t) for t in np.asarray(w["item_target"])[job.idx])), sorted(set(int(s) for s in np.asarray(w["item_source"])[job.idx]))
        for stride in (16, 12):
            with tempfile.TemporaryDirectory(prefix="ndtb_job_") as d:
                path = jobfile.write_job(os.path.join(d, "job.bin"), [w["targets"][t] for t in tids], [w["sources"][s] for s in sids],
                                         job.it, job.isrc, np.asarray(w["guesses"], dtype=np.float32)[job.idx], stride=stride)
                r = subprocess.run([exe, path, "--steps", str(a.steps), "--warmup", str(max(2, a.warmup)), "--inflight", str(max(1, a.inflight_e2e)),
                                    "--strict-math", str(a.strict_math), "--device", str(device)], capture_output=True, text=True, timeout=900)
            if r.returncode != 0:
                out["stride_%d" % stride] = {"error": r.stderr.strip()[-300:]}
                continue
            out["stride_%d" % stride] = json.loads(r.stdout.strip().splitlines()[-1])
        out["note"] = ("examples/batch_pipeline.cpp over the C ABI: page-locked host clouds, two batches in flight, host clock around whole steps "
                       "(uploads, table builds, registrations, results back); stride_12 = packed x y z records, 3/4 of the bytes")
    except Exception as e:      # the C++ arm is an explanation of the e2e number, not the number: never fail the line for it
        out["error"] = repr(e)[:300]
    return out


def roofline_block(agg, ms_total, clocks, K=7, sm_count=148, shared_tables=True, strict=False):
    """B_pt = 16 + 8*K + 48*h bytes per evaluated point (points and (point, voxel) pairs COUNTED by the kernel; the
    term-by-term kernel actually fetches a 64-byte record per pair, the algorithmic figure stays 8d's), over the kernel's own time (CUDA events
    around every derivative launch of the first pass).  What BINDS the kernel is stated from the ncu capture: with 64 registrations
    sharing each submap's table the gather is served by L2 (DRAM traffic ~1 % of the algorithmic bytes), and the limiter is
    instruction issue / the L1 (shared-memory) pipe -- `frac_issue` is the share of the SMs' issue slots the kernel's warp
    instructions fill at the SM clock measured during the run."""
    peaks, ncu = load_peaks(), load_ncu_constants()
    if strict:
        ncu = ncu.get("strict", {})
    rec = 48     # mean 3 x f64 + symmetric icov 6 x f32 per valid neighbour (the strict kernel's 64-byte record carries the full 3 x 3)
    peak = float(peaks.get("hbm_gbs", 6650.0))
    alg_bytes = agg["points_evaluated"] * (16 + 8 * K) + rec * agg["pairs_evaluated"]
    deriv_ms = agg["deriv_ms"]
    achieved = alg_bytes / (deriv_ms / 1e3) / 1e9 if deriv_ms > 0 else 0.0
    launches = max(1, int(agg["deriv_launches"]))
    wi32 = ncu.get("warp_instructions_per_32_points")
    sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0
    frac_issue = None
    if wi32 and deriv_ms > 0:
        frac_issue = (agg["points_evaluated"] / 32.0 * wi32) / (sm_count * 4 * (deriv_ms / 1e3) * sm_mhz * 1e6)
    if not shared_tables:
        bound = "hbm"
    elif strict:
        bound = ("instruction issue + the FP64 accumulate path (f32->f64 converts on the XU pipe, shared-memory accumulators); "
                 "L2-resident gather: HBM is NOT the binding roof here")
    else:
        bound = "issue slots / L1 look-ups (L2-resident gather; HBM is NOT the binding roof here)"
    return {"bound": bound,
            "kernel": "k_derivatives<DIRECT7> (term-by-term)" if strict else "k_derivatives_fast<DIRECT7>", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak if peak else None, "traffic": ncu.get("dram_bytes_per_launch"),
            "frac_issue": frac_issue, "warp_instructions_per_32_points": wi32, "sm_mhz_used": sm_mhz,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s",
            "algorithmic_bytes_per_launch": alg_bytes / launches, "launches": launches,
            "kernel_ms_per_launch": deriv_ms / launches,
            "mean_neighbours_per_point": agg["pairs_evaluated"] / max(1, agg["points_evaluated"]),
            "kernel_share_of_step": agg["deriv_ms_max_rank"] / ms_total if ms_total else None,
            "note": "achieved = algorithmic bytes (16 + 8*7 + %d*neighbours per evaluated point, counted by the kernel) / kernel time; "
                    "`traffic` = DRAM bytes of one launch from the ncu capture named in profiles/roofline_traffic.json. The gather is "
                    "served by L2 (64 registrations share a table), so the fraction of the HBM roof says how close the kernel runs to "
                    "what HBM could feed, not what binds it: that is `bound` / `frac_issue`; the HBM-bound case is the single-align "
                    "line of --workload c3" % rec}


def c5_line(a, world, mode, value, e2e, M, agg, clocks, job, n_per_step):
    ms, steps = M["ms"], a.steps
    rf = roofline_block(agg, ms, clocks, strict=a.strict_math in (-1, 1))
    ae = M["agg_e2e"]
    return {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": a.warmup,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": mode, "vs_baseline": None,
            "dtype": DTYPE, "data": "synthetic", "config": workload_config(a, world, mode),
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": job.h2d_bytes, "d2h_bytes_per_step": job.d2h_bytes,
                    "ms_per_step": M["ms_e2e"] / steps, "batches_in_flight": max(1, a.inflight_e2e),
                    "bytes_are": "per rank (this rank's shard of the job)" if mode == "strong" and world > 1 else "per rank"},
            "gpu_launches": int(agg["kernel_launches"]),
            "kernel_mode": MODE_NAMES[a.strict_math],
            "roofline": rf,
            "breakdown_ms_per_step": {"voxel_build": agg["build_ms"] / steps / world, "derivatives": agg["deriv_ms"] / steps / world,
                                      "device_total": agg["total_ms"] / steps / world, "strict_replay": agg["replay_ms"] / steps / world,
                                      "strict_replay_derivatives": agg["replay_deriv_ms"] / steps / world,
                                      "registrations_replayed_per_step": agg["replayed"] / steps,
                                      "e2e_device_total": ae["total_ms"] / steps / world,
                                      "note": "mean over ranks of the per-rank device times (CUDA events inside the library)"},
            "clocks": clocks}


# ---------------------------------------------------------------------------------------------------------------------
# single-align workloads (C1 / C3 / C4) and the odometry sequence (C2): latency lines, one GPU
# ---------------------------------------------------------------------------------------------------------------------
SINGLE = {"c1": dict(n_source=20000, n_target=300000, res=1.0, seed_off=0, kw={}, text="C1: 20k-pt scan vs 300k-pt localmap, res 1.0 m"),
          "c3": dict(n_source=20000, n_target=5000000, res=0.5, seed_off=3, kw=dict(keyframe_leaf=0.25), text="C3: 20k-pt scan vs 5M-pt localmap, res 0.5 m"),
          "c4": dict(n_source=120000, n_target=1000000, res=2.0, seed_off=4, kw=dict(dense=True), text="C4: undownsampled 120k-pt scan vs 1M-pt map, res 2.0 m")}


def run_single(a):
    import torch
    from xchu_slam_b200 import ndt, synth
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    cfg = SINGLE[a.workload]
    src, tgt, gt, guess = synth.make_case(n_source=cfg["n_source"], n_target=cfg["n_target"], seed=synth.BASE_SEED + cfg["seed_off"], **cfg["kw"])
    methods = [("DIRECT7", ndt.DIRECT7)] + ([("DIRECT1", ndt.DIRECT1), ("KDTREE", ndt.KDTREE)] if a.workload == "c3" else [])
    out = {}
    sampler = ClockSampler(0)
    sampler.start()
    for name, m in methods:
        g = ndt.NormalDistributionsTransform()
        g.setResolution(cfg["res"]); g.setStepSize(0.1); g.setTransformationEpsilon(0.01); g.setMaximumIterations(30)
        g.setNeighborhoodSearchMethod(m)
        build_ms, align_ms, e2e_ms = [], [], []
        for k in range(a.warmup + a.steps):
            t0 = time.perf_counter()
            g.setInputTarget(tgt)
            t1 = time.perf_counter()
            g.setInputSource(src)
            g.align(guess)
            t2 = time.perf_counter()
            if k >= a.warmup:
                build_ms.append(1e3 * (t1 - t0)); e2e_ms.append(1e3 * (t2 - t0))
        for k in range(a.warmup + 4 * a.steps):      # align with the target prebuilt (device time of the whole Newton loop)
            g.align(guess)
            if k >= a.warmup:
                align_ms.append(g.lastTiming()["align_ms"])
        r = g.result()
        # the derivative kernel alone, ONE registration per launch (nothing shares the voxel table): a one-item batch, both kernels
        roof = {}
        for kname, strict in (("factorised", 0), ("term_by_term", 1)):
            b = ndt.BatchNDT()
            b.set_params(resolution=cfg["res"], step_size=0.1, trans_eps=0.01, max_iter=30, search_method=m, strict_math=strict)
            b.add_target(tgt); b.add_source(src)
            for _ in range(3):
                b.run([0], [0], [guess])
            st = b.stats()
            st["deriv_ms_max_rank"] = st["deriv_ms"]
            K = {ndt.DIRECT7: 7, ndt.DIRECT1: 1, ndt.KDTREE: 27}[m]
            rb = roofline_block(st, st["total_ms"], None, K=K, shared_tables=False)
            roof[kname] = {k: rb[k] for k in ("achieved", "peak", "frac", "kernel_ms_per_launch", "launches", "algorithmic_bytes_per_launch", "mean_neighbours_per_point")}
        out[name] = {"target_build_ms": {"median": float(np.median(build_ms)), "min": float(np.min(build_ms))},
                     "align_prebuilt_target_ms": {"median": float(np.median(align_ms)), "min": float(np.min(align_ms))},
                     "end_to_end_ms": {"median": float(np.median(e2e_ms)), "min": float(np.min(e2e_ms))},
                     "iterations": r["iterations"], "converged": r["converged"], "derivative_kernel_single_registration": roof}
    clocks = sampler.stop()
    first = out[methods[0][0]]
    line = {"metric": "single align latency (%s)" % cfg["text"], "value": first["end_to_end_ms"]["median"], "unit": "ms", "n_gpus": 1, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": first["end_to_end_ms"]["median"], "higher_is_better": False, "scaling": "n/a", "vs_baseline": None,
            "dtype": DTYPE, "data": "synthetic",
            "config": {"workload": cfg["text"] + ", step 0.1, eps 0.01, max_iter 30; handle path (term-by-term kernel, one CUDA graph per align); "
                                                 "value = setInputTarget + setInputSource + align from host memory"},
            "methods": out, "clocks": clocks,
            "roofline": {"bound": "latency (one 20k-pt registration is ~4 MB per evaluation: launch / dependency latency, not bandwidth)",
                         "note": "per-method HBM figures under methods.*.derivative_kernel_single_registration"}}
    print(json.dumps(line), flush=True)


def run_c2(a):
    from xchu_slam_b200 import build, synth
    build.build_cuda()
    exe = build.build_examples()
    tmp = "/tmp/ndt_c2"
    os.makedirs(tmp, exist_ok=True)
    seq = synth.write_sequence(os.path.join(tmp, "seq%d.bin" % a.c2_scans), n_scans=a.c2_scans, n_points=20000)

    def run(*flags):
        outp = os.path.join(tmp, "bench_c2.txt")
        r = subprocess.run([exe, seq, outp] + list(flags), capture_output=True, text=True)
        if r.returncode != 0:
            raise SystemExit("odom_sequence failed: " + r.stderr[-400:])
        rows = np.array([l.split() for l in open(outp) if not l.startswith("#")], dtype=np.float64)
        s = [l for l in open(outp) if l.startswith("#")][0].split()
        return rows, dict(zip(s[1::2], s[2::2]))
    run("--res", "1.0", "--max-scans", "20", "--device-localmap")
    modes = {}
    for name, flags in (("distance_threshold_device_localmap", ["--res", "1.0", "--device-localmap"]),
                        ("distance_threshold_device_localmap_full_rebuild", ["--res", "1.0", "--device-localmap", "--no-incremental"]),
                        ("distance_threshold_host_localmap", ["--res", "1.0"]),
                        ("keyframe_count_window_device", ["--res", "2.0", "--refresh", "surround", "--device-localmap"]),
                        ("keyframe_count_window_host", ["--res", "2.0", "--refresh", "surround"])):
        rows, s = run(*flags)
        err = np.linalg.norm(rows[:, 4:6] - rows[:, 14:16], axis=1)
        modes[name] = {"scans_per_s": float(s["scans_per_s"]), "align_ms_mean": float(s["align_ms_mean"]), "target_ms_mean": float(s["target_ms_mean"]),
                       "fitness_ms_mean": float(s["fitness_ms_mean"]), "surround_ms_mean": float(s.get("surround_ms_mean", 0)),
                       "target_sets": int(s["target_sets"]), "incremental": int(s["incremental"]), "surround_sets": int(s.get("surround_sets", 0)),
                       "converged_frac": float(np.mean(rows[:, 2])), "xy_drift_max_m": float(err.max())}
    v = modes["distance_threshold_device_localmap"]["scans_per_s"]
    line = {"metric": "odometry scans/s (C2: sequential scan-to-localmap NDT, 20k-pt scans)", "value": v, "unit": "scans/s", "n_gpus": 1,
            "steps": a.c2_scans, "warmup": 20, "ms_per_step": 1e3 / v, "higher_is_better": True, "scaling": "n/a", "vs_baseline": None, "dtype": DTYPE,
            "data": "synthetic", "config": {"workload": "C2: %d-scan synthetic sequence, examples/odom_sequence (the mirror of odom_node's loop over the "
                                                        "drop-in header), DIRECT7, eps 0.01, <= 30 iterations; one step = one scan" % a.c2_scans},
            "modes": modes}
    print(json.dumps(line), flush=True)


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "c5":
        run_c5(a)
    elif a.workload == "c2":
        run_c2(a)
    else:
        run_single(a)


if __name__ == "__main__":
    main()
