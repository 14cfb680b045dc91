"""Loop-closure refinement on the device: the host-side mirror of PGO::ICPRefine and
PGO::LoopFindNearKeyframesCloud (reference xchu_mapping/src/pgo_node.cpp:404-476, 530-554), SURVEY.md
section 8(f) rank 3.

What the reference does per loop candidate (loop_kf_idx, curr_kf_idx), serially on the CPU:
  source = LoopFindNearKeyframesCloud(key = curr, submap_size = 0,  root_idx = loop)
  target = LoopFindNearKeyframesCloud(key = loop, submap_size = 25, root_idx = loop)
  pcl::IterativeClosestPoint(source -> target); accept when converged and getFitnessScore() <= 0.3.
LoopFindNearKeyframesCloud stacks the keyframes root_idx - submap_size .. root_idx + submap_size, each moved into
the map frame by its own pose (TransformCloud2Map, pgo_node.cpp:568-590), and VoxelGrid-downsamples the stack
at `filter_size`.  Its `key` argument is never read (pgo_node.cpp:537: `keyNear = root_idx + i`), so the
reference's "current" cloud is in fact the loop keyframe itself; `source_mode="reference"` reproduces that,
`"current"` registers the current keyframe placed by its own pose (what the argument list suggests was meant),
`"current_at_loop_pose"` places the current keyframe with the loop keyframe's pose (the SC-A-LOAM original).

Here every candidate of a call is assembled in HBM from the resident keyframe store (ndtb_batch_add_keyframe /
ndtb_batch_add_submap), all candidates are registered concurrently by the batched NDT (ndtb_batch_run) from the
identity guess, and gated on the batched getFitnessScore (ndtb_batch_fitness).  PCL's ICP itself is outside
/root/reference and is replaced, not restated: the registration results are NDT's (parity: oracle NDT on the same
clouds), the clouds and the fitness gate are the reference's.
"""
import numpy as np

from .ndt import BatchNDT


def get_transformation(x, y, z, roll, pitch, yaw):
    """pcl::getTransformation(float x, ..., float yaw) -> Affine3f (PCL common/eigen.h), float arithmetic throughout; sinf / cosf
    taken as (float) sin / cos of the float argument widened to double (the build-wide convention)."""
    f = np.float32
    x, y, z, roll, pitch, yaw = f(x), f(y), f(z), f(roll), f(pitch), f(yaw)
    cosf = lambda a: f(np.cos(np.float64(a)))
    sinf = lambda a: f(np.sin(np.float64(a)))
    A, B, Cc, D, E, F = cosf(yaw), sinf(yaw), cosf(pitch), sinf(pitch), cosf(roll), sinf(roll)
    DE, DF = f(D * E), f(D * F)
    T = np.zeros((4, 4), dtype=np.float32)
    T[0] = [f(A * Cc), f(f(A * DF) - f(B * E)), f(f(B * F) + f(A * DE)), x]
    T[1] = [f(B * Cc), f(f(A * E) + f(B * DF)), f(f(B * DE) - f(A * F)), y]
    T[2] = [f(-D), f(Cc * F), f(Cc * E), z]
    T[3, 3] = 1.0
    return T


class LoopRefiner:
    """Keyframe store + batched loop-candidate refinement on one device.

    filter_size: leaf of downSizeFilterICP (pgo_node.cpp:75); history_keyframe_search_num: half window of the target
    submap (pgo_node.cpp:422); loop_fitness_score_threshold: pgo_node.cpp:442.  ndt_params go to ndtb_batch_set_params."""

    def __init__(self, device=0, filter_size=0.5, history_keyframe_search_num=25, loop_fitness_score_threshold=0.3,
                 source_mode="reference", **ndt_params):
        if source_mode not in ("reference", "current", "current_at_loop_pose"):
            raise ValueError("source_mode")
        self.batch = BatchNDT(device)
        if ndt_params:
            self.batch.set_params(**ndt_params)
        self.filter_size = float(filter_size)
        self.half_window = int(history_keyframe_search_num)
        self.threshold = float(loop_fitness_score_threshold)
        self.source_mode = source_mode
        self.poses = []      # 4x4 float32 per keyframe: pcl::getTransformation(keyframePosesUpdated[i])

    def add_keyframe(self, cloud, pose6):
        """keyframeLaserClouds.push_back / keyframePosesUpdated.push_back (pgo_node.cpp:216-219).  cloud: (n, >=3) float32 in
        the sensor frame (column 3 = intensity); pose6: x y z roll pitch yaw."""
        kid = self.batch.add_keyframe(cloud)
        self.poses.append(get_transformation(*pose6))
        assert kid == len(self.poses) - 1
        return kid

    def update_poses(self, poses6):
        """keyframePosesUpdated after an iSAM2 update (pgo_node.cpp:504-527): one pose6 per stored keyframe."""
        if len(poses6) != len(self.poses):
            raise ValueError("one pose per keyframe")
        self.poses = [get_transformation(*p) for p in poses6]

    def num_keyframes(self):
        return len(self.poses)

    def _pose_array(self, override=None):
        P = np.stack(self.poses).astype(np.float32) if self.poses else np.zeros((0, 4, 4), np.float32)
        if override:
            P = P.copy()
            for k, T in override.items():
                P[k] = T
        return P

    def refine(self, loop_pairs, download_clouds=False, timing=None):
        """loop_pairs: [(loop_kf_idx, curr_kf_idx), ...].  One batched run; per pair a dict(loop, curr, accepted, converged,
        fitness, T (4x4 float32, the correction in the map frame), iterations, n_source, n_target [, source, target]).
        timing: a dict that receives the host wall time (s) of the three stages (assemble, register, fitness)."""
        import time
        t0 = time.perf_counter()
        b = self.batch
        b.clear()
        n_kf = len(self.poses)
        P = self._pose_array()
        # the distinct submaps of the call: targets keyed by the loop keyframe, sources by (keyframe, whose pose places it)
        tkeys, skeys, pair_keys = [], [], []
        for loop_idx, curr_idx in loop_pairs:
            loop_idx, curr_idx = int(loop_idx), int(curr_idx)
            if self.source_mode == "reference":
                skey = (loop_idx, loop_idx)
            elif self.source_mode == "current":
                skey = (curr_idx, curr_idx)
            else:
                skey = (curr_idx, loop_idx)
            if loop_idx not in tkeys:
                tkeys.append(loop_idx)
            if skey not in skeys:
                skeys.append(skey)
            pair_keys.append((loop_idx, curr_idx, skey))
        # everything placed by the stored poses goes to the device in ONE call (assembled concurrently on the build lanes) ...
        roots, hws, ast, slot = list(tkeys), [self.half_window] * len(tkeys), [True] * len(tkeys), {}
        for skey in skeys:
            kf, pose_of = skey
            in_range = 0 <= kf < n_kf and 0 <= pose_of < n_kf
            if in_range and kf != pose_of:
                continue
            slot[skey] = len(roots)
            roots.append(kf if in_range else -1)        # no such keyframe: empty cloud
            hws.append(0)
            ast.append(False)
        ids, npts = b.add_submaps(roots, hws, P, self.filter_size, ast) if roots else ([], [])
        tgt_cache = {k: (int(ids[i]), int(npts[i])) for i, k in enumerate(tkeys)}
        src_cache = {k: (int(ids[i]), int(npts[i])) for k, i in slot.items()}
        # ... a keyframe placed by ANOTHER keyframe's pose (current_at_loop_pose) needs its own pose table
        for skey in skeys:
            if skey not in src_cache:
                kf, pose_of = skey
                src_cache[skey] = b.add_submap(kf, 0, self._pose_array({kf: P[pose_of]}), self.filter_size, as_target=False)
        item_t, item_s, meta = [], [], []
        for loop_idx, curr_idx, skey in pair_keys:
            (tid, tn), (sid, sn) = tgt_cache[loop_idx], src_cache[skey]
            meta.append(dict(loop=loop_idx, curr=curr_idx, n_source=sn, n_target=tn, target_id=tid, source_id=sid))
            item_t.append(tid)
            item_s.append(sid)
        # a candidate whose clouds are empty cannot be registered (the reference's ICP would refuse the empty input)
        live = [i for i, m in enumerate(meta) if m["n_source"] > 0 and m["n_target"] > 0]
        out = [dict(m, accepted=False, converged=False, fitness=float(np.finfo(np.float64).max), T=np.eye(4, dtype=np.float32),
                    iterations=0) for m in meta]
        t1 = t2 = t3 = time.perf_counter()
        if live:
            guesses = np.tile(np.eye(4, dtype=np.float32), (len(live), 1, 1))
            res = b.run([item_t[i] for i in live], [item_s[i] for i in live], guesses)
            t2 = time.perf_counter()
            fit = b.fitness(len(live))
            t3 = time.perf_counter()
            for j, i in enumerate(live):
                r = res[j]
                out[i].update(converged=r["converged"], fitness=float(fit[j]), T=r["T"], iterations=r["iterations"],
                              accepted=bool(r["converged"] and fit[j] <= self.threshold))      # pgo_node.cpp:443
        if timing is not None:
            timing.update(assemble_s=t1 - t0, register_s=t2 - t1, fitness_s=t3 - t2)
        if download_clouds:
            for o in out:
                o["source"] = b.download(o["source_id"], o["n_source"], is_target=False)
                o["target"] = b.download(o["target_id"], o["n_target"], is_target=True)
        return out
