"""Sharding of independent registrations over the GPUs of one box (north_star item 4).

A single registration does not shard (20k points, a 43-double reduction, <= 32 strictly sequential
Newton steps); independent registrations do, perfectly.  One process per GPU (torch.distributed);
whole submaps are dealt round-robin (or, when that would unbalance the ranks, items sorted by target id are cut
into contiguous blocks) so that each distinct submap's voxel table is built on as few GPUs as possible; every rank runs its block through its own
BatchNDT; there is NO collective inside a registration, only one final gather of the per-item
results (20 doubles each).  The same code runs over gloo on CPU tensors (tests) and nccl on CUDA
tensors (bench / production).
"""
import numpy as np

RESULT_WIDTH = 21   # T[16], iterations, converged, score, trans_probability, n_evals


def plan_shards(item_target, world_size):
    """Returns a list (one entry per rank) of index arrays into the item list.

    Whole submaps are dealt round-robin in target-id order (distinct target j -> rank j mod world_size) when that leaves the
    ranks' item counts within max(1, 2 %) of each other: every voxel table is then built on exactly one GPU, and neighbouring
    submaps -- windows over the same stretch of the drive, whose registrations are hard or easy together -- land on different
    ranks (measured on the C5 job at 8 ranks: heaviest rank 1.08x the mean, against 1.39x for contiguous blocks of submaps).
    Otherwise (few or very unequal submaps) the items are ordered by target id (stable) and cut into world_size contiguous
    blocks whose sizes differ by at most one, so that a submap is split over at most two neighbouring ranks per cut."""
    item_target = np.asarray(item_target)
    order = np.argsort(item_target, kind="stable")
    n = order.shape[0]
    if n and world_size > 1:
        distinct, inv = np.unique(item_target, return_inverse=True)
        if len(distinct) >= world_size:
            rank_of_item = inv % world_size
            sizes = np.bincount(rank_of_item, minlength=world_size)
            if sizes.max() - sizes.min() <= max(1, int(0.02 * n / world_size)):
                return [order[rank_of_item[order] == r] for r in range(world_size)]
    base, extra = divmod(n, world_size)
    out, pos = [], 0
    for r in range(world_size):
        cnt = base + (1 if r < extra else 0)
        out.append(order[pos:pos + cnt])
        pos += cnt
    return out


def local_problem(idx, item_target, item_source):
    """Restrict a batch to the items `idx`: returns (target_ids, source_ids, local_item_target,
    local_item_source) where the local item arrays index into target_ids / source_ids."""
    it = np.asarray(item_target)[idx]
    isrc = np.asarray(item_source)[idx]
    tids, lt = np.unique(it, return_inverse=True)
    sids, ls = np.unique(isrc, return_inverse=True)
    return tids, sids, lt.astype(np.int32), ls.astype(np.int32)


def pack_results(results):
    out = np.zeros((len(results), RESULT_WIDTH), np.float64)
    for i, r in enumerate(results):
        out[i, :16] = np.asarray(r["T"], dtype=np.float64).reshape(-1)
        out[i, 16:] = [r["iterations"], int(r["converged"]), r["score"], r["trans_probability"], r.get("n_evals", 0)]
    return out


def unpack_results(arr):
    return [dict(T=row[:16].reshape(4, 4).astype(np.float32), iterations=int(row[16]), converged=bool(row[17]),
                 score=float(row[18]), trans_probability=float(row[19]), n_evals=int(row[20])) for row in arr]


def gpu_runner(device, **params):
    """runner for run_sharded on a CUDA device: a BatchNDT with the given ndtb_params fields."""
    from . import ndt
    b = ndt.BatchNDT(device=device)
    b.set_params(**params)

    def run(targets, sources, it, isrc, guesses):
        b.clear()
        for t in targets:
            b.add_target(t)
        for s in sources:
            b.add_source(s)
        return b.run(it, isrc, guesses)
    return run


def run_sharded(runner, targets, sources, item_target, item_source, guesses, group=None, device="cpu"):
    """Runs the whole batch across the ranks of `group` (default world) and returns, on EVERY rank,
    the list of per-item results in the original item order.  `runner(targets, sources, item_target,
    item_source, guesses) -> [result dict]` does the local work (gpu_runner on a GPU box).
    Without an initialised process group this is the single-rank case."""
    import torch
    import torch.distributed as dist
    multi = dist.is_available() and dist.is_initialized()
    world = dist.get_world_size(group) if multi else 1
    rank = dist.get_rank(group) if multi else 0
    n = len(item_target)
    plan = plan_shards(item_target, world)
    idx = plan[rank]
    local = np.zeros((0, RESULT_WIDTH))
    if len(idx):
        tids, sids, lt, ls = local_problem(idx, item_target, item_source)
        res = runner([targets[t] for t in tids], [sources[s] for s in sids], lt, ls, np.asarray(guesses)[idx])
        local = pack_results(res)
    if not multi:
        full = np.zeros((n, RESULT_WIDTH))
        full[idx] = local
        return unpack_results(full)
    # the only communication of the whole job: one all_gather of (padded) result rows
    width = max(len(p) for p in plan)
    buf = torch.zeros((width, RESULT_WIDTH), dtype=torch.float64, device=device)
    if len(idx):
        buf[:len(idx)] = torch.from_numpy(local).to(device)
    gathered = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(gathered, buf, group=group)
    full = np.zeros((n, RESULT_WIDTH))
    for r in range(world):
        if len(plan[r]):
            full[plan[r]] = gathered[r][:len(plan[r])].cpu().numpy()
    return unpack_results(full)
