"""Host-side sharding of independent registrations (xchu_slam_b200/shard.py): plan properties on one
process, and the N>1 path with world_size 2 over gloo on CPU.  The local work in the gloo test is
done by the CPU oracle standing in for the per-rank BatchNDT (a test may use the oracle; the product
runner is shard.gpu_runner)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_plan_is_a_balanced_partition_grouped_by_target():
    from xchu_slam_b200 import shard
    rng = np.random.default_rng(0)
    it = rng.integers(0, 9, 101)
    for world in (1, 2, 3, 4, 8):
        plan = shard.plan_shards(it, world)
        allidx = np.concatenate(plan)
        assert sorted(allidx.tolist()) == list(range(101))                 # a partition
        sizes = [len(p) for p in plan]
        assert max(sizes) - min(sizes) <= 1                                # balanced
        # contiguous in target order: a target is split over at most two neighbouring ranks per cut
        tmax = -1
        for p in plan:
            if len(p):
                assert it[p].min() >= tmax
                tmax = it[p].max()
        n_build = sum(len(np.unique(it[p])) for p in plan if len(p))
        assert n_build <= len(np.unique(it)) + world - 1                   # table builds: at most one extra per cut


def test_plan_deals_whole_submaps_round_robin_when_that_is_balanced():
    """The C5 shape: 64 submaps x 64 registrations.  Every submap's table is built on exactly one rank, the ranks hold equal
    item counts, and neighbouring submaps (correlated difficulty) sit on different ranks."""
    from xchu_slam_b200 import shard
    it = np.repeat(np.arange(64), 64)
    for world in (2, 4, 8):
        plan = shard.plan_shards(it, world)
        assert sorted(np.concatenate(plan).tolist()) == list(range(4096))
        assert all(len(p) == 4096 // world for p in plan)
        owners = [set(np.unique(it[p]).tolist()) for p in plan]
        assert sum(len(o) for o in owners) == 64                            # no table built twice
        for r, o in enumerate(owners):
            assert o == set(range(r, 64, world))
        for p in plan:
            assert np.all(np.diff(it[p]) >= 0)                              # items of a rank stay grouped by target


def test_plan_more_ranks_than_items():
    from xchu_slam_b200 import shard
    plan = shard.plan_shards([0, 0, 1], 8)
    assert sum(len(p) for p in plan) == 3 and all(len(p) <= 1 for p in plan)


def test_local_problem_remaps_ids():
    from xchu_slam_b200 import shard
    it = np.array([5, 5, 7, 9, 9, 9])
    isrc = np.array([3, 4, 3, 8, 8, 2])
    tids, sids, lt, ls = shard.local_problem(np.array([2, 3, 5]), it, isrc)
    assert tids.tolist() == [7, 9] and sids.tolist() == [2, 3, 8]
    assert tids[lt].tolist() == [7, 9, 9] and sids[ls].tolist() == [3, 8, 2]


def test_pack_unpack_round_trip():
    from xchu_slam_b200 import shard
    r = [dict(T=np.arange(16, dtype=np.float32).reshape(4, 4), iterations=7, converged=True, score=1.5, trans_probability=0.25, n_evals=8)]
    u = shard.unpack_results(shard.pack_results(r))[0]
    assert np.array_equal(u["T"], r[0]["T"]) and u["iterations"] == 7 and u["converged"] and u["score"] == 1.5 and u["n_evals"] == 8


def _oracle_runner(targets, sources, it, isrc, guesses):
    from oracle import oracle
    out = []
    for k in range(len(it)):
        o = oracle.OracleNDT()
        o.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, num_threads=1)
        o.set_target(targets[it[k]])
        o.set_source(sources[isrc[k]])
        r = o.align(guesses[k])
        r["n_evals"] = len(r["trace"])
        out.append(r)
    return out


def _problem():
    gold = np.load(os.path.join(ROOT, "tests", "golden", "ndt_tiny.npz"))
    from xchu_slam_b200 import synth
    tgt, src, gt = gold["target"], gold["source"], gold["gt"]
    targets = [tgt, tgt[: len(tgt) * 3 // 4], tgt[len(tgt) // 8:]]
    sources = [src, src[::2].copy()]
    it = np.array([0, 1, 2, 0, 1, 2, 0], np.int32)
    isrc = np.array([0, 1, 0, 1, 0, 1, 0], np.int32)
    guesses = np.stack([synth.perturbed_guess(gt, 300 + i) for i in range(len(it))])
    return targets, sources, it, isrc, guesses


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    from xchu_slam_b200 import shard
    dist.init_process_group("gloo", rank=rank, world_size=world)
    targets, sources, it, isrc, guesses = _problem()
    res = shard.run_sharded(_oracle_runner, targets, sources, it, isrc, guesses)
    q.put((rank, shard.pack_results(res)))
    dist.barrier()
    dist.destroy_process_group()


def test_run_sharded_world2_gloo_matches_single_rank():
    import torch.multiprocessing as mp
    from xchu_slam_b200 import shard
    targets, sources, it, isrc, guesses = _problem()
    single = shard.pack_results(shard.run_sharded(_oracle_runner, targets, sources, it, isrc, guesses))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for r in range(2):                       # every rank holds the full, identically ordered result list
        assert np.array_equal(got[r], single)
    assert np.all(single[:, 16] >= 1)


# ---- the same job on real devices (marked gpu; the two-device tests skip on a one-GPU box) --------------------------------
def _c5_small():
    from xchu_slam_b200 import synth
    return synth.batch_workload(n_submaps=4, scans_per_submap=4, hyps_per_scan=4, n_target=200000, n_source=8000)


def _single_batch(w, device=0):
    from xchu_slam_b200 import ndt
    b = ndt.BatchNDT(device=device)
    b.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2)
    for t in w["targets"]:
        b.add_target(t)
    for s in w["sources"]:
        b.add_source(s)
    return b.run_raw(w["item_target"], w["item_source"], w["guesses"])


def _same_results(a, b):
    """SURVEY.md section 4, multi-GPU row: identical per-item results regardless of sharding.  Iteration counts, convergence
    flags and evaluation counts EQUAL; transforms equal to float rounding (the FP64 partial sums of an evaluation are grouped
    by the block size, which follows the number of points a device has in flight: ~1e-16 per evaluation)."""
    assert np.array_equal(a["iterations"], b["iterations"]) and np.array_equal(a["converged"], b["converged"])
    assert np.array_equal(a["n_evals"], b["n_evals"])
    assert np.max(np.abs(a["T"] - b["T"])) < 1e-6
    assert np.max(np.abs(a["score"] - b["score"]) / np.maximum(1.0, np.abs(b["score"]))) < 1e-9
    return float(np.mean(np.all(a["T"] == b["T"], axis=1)))


@pytest.mark.gpu
def test_multi_entry_point_one_device_listed_twice_matches_single_batch():
    """ndtb_multi_run (the library-side sharding a C++ caller such as pgo_node gets) with device 0 listed twice: two
    ndtb_batches on two host threads, each with half of the items, must return what one batch returns."""
    from xchu_slam_b200 import ndt
    w = _c5_small()
    single = _single_batch(w)
    m = ndt.MultiNDT([0, 0])
    m.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2)
    res = m.run_raw(w["targets"], w["sources"], w["item_target"], w["item_source"], w["guesses"])
    frac_bitwise = _same_results(res, single)
    assert frac_bitwise > 0.9
    st = m.stats()
    assert st["items"].tolist() == [32, 32] and st["targets"].tolist() == [2, 2]     # each half builds only its own submaps
    # one listed device = the single batch, bit for bit (same block decomposition)
    m1 = ndt.MultiNDT([0])
    m1.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2)
    r1 = m1.run_raw(w["targets"], w["sources"], w["item_target"], w["item_source"], w["guesses"])
    assert np.array_equal(r1["T"], single["T"]) and np.array_equal(r1["iterations"], single["iterations"])


@pytest.mark.gpu
def test_multi_entry_point_two_devices():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    from xchu_slam_b200 import ndt
    w = _c5_small()
    single = _single_batch(w)
    m = ndt.MultiNDT([0, 1])
    m.set_params(resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2)
    res = m.run_raw(w["targets"], w["sources"], w["item_target"], w["item_source"], w["guesses"])
    _same_results(res, single)


def _nccl_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    from xchu_slam_b200 import shard
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    w = _c5_small()
    runner = shard.gpu_runner(rank, resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2)
    res = shard.run_sharded(runner, w["targets"], w["sources"], w["item_target"], w["item_source"], w["guesses"],
                            device=torch.device("cuda", rank))
    q.put((rank, shard.pack_results(res)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_run_sharded_world2_nccl_matches_one_gpu():
    """shard.run_sharded with the product runner on two GPUs: NCCL all_gather of the 21-double result rows; every rank ends
    with the full list, equal to the one-GPU run."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    import torch.multiprocessing as mp
    from xchu_slam_b200 import shard
    w = _c5_small()
    one = shard.pack_results(shard.run_sharded(shard.gpu_runner(0, resolution=1.0, step_size=0.1, trans_eps=0.01, max_iter=30, search_method=2),
                                               w["targets"], w["sources"], w["item_target"], w["item_source"], w["guesses"]))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(got[0], got[1])
    assert np.array_equal(got[0][:, 16:18], one[:, 16:18]) and np.array_equal(got[0][:, 20], one[:, 20])     # iterations, converged, n_evals
    assert np.max(np.abs(got[0][:, :16] - one[:, :16])) < 1e-6
