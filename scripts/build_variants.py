"""Kernel variants for scripts/gpurun_variants.sh: each entry is compiled into variants/lib_<name>.so (git-ignored, travels to the
GPU box), then the in-tree library is rebuilt with the default macros (the variant builds share its object paths).
    python scripts/build_variants.py [name ...]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from xchu_slam_b200 import build as b  # noqa: E402

VARIANTS = {
    # strict (term-by-term) kernel: FP64 accumulators in shared memory, 128 registers, two blocks per SM (ncu r02: 241 registers,
    # one block = 8 warps per SM, 59 % of cycles with no eligible warp)
    "strict_smem": ("NDTB_STRICT_ACC_SMEM=1", "NDTB_MINBLOCKS=2"),
    "strict_mb2": ("NDTB_MINBLOCKS=2",),
    # 64-B aligned voxel records, one 256-bit + one 128-bit load per (point, voxel) pair (ndt_internal.h, ndt_fast.cu)
    "rec64": ("NDTB_FAST_REC64=1",),
    # the same with the next point's seven cell codes fetched one iteration ahead
    "rec64_pfp2": ("NDTB_FAST_REC64=1", "NDTB_FAST_PFP=2"),
    # three resident blocks per SM (more registers per thread, fewer warps)
    "minblocks3": ("NDTB_FAST_MINBLOCKS=3",),
    # FP64 accumulators in shared memory free ~56 registers: the next point's record AND cell codes prefetched without spills
    # (ptxas: 124 registers at 4 blocks/SM, 96 at 5 blocks/SM, 0 spill bytes); combinations not in the r01 sweeps
    # unpredicated record loads on a clamped index: ~64 fewer static instructions in the DIRECT7 kernel (no 12-register copy per
    # cell); 72 B of spills at 128 registers alone, none with shared-memory accumulators
    "uncond": ("NDTB_FAST_UNCOND=1",),
    "uncond_rec64": ("NDTB_FAST_UNCOND=1", "NDTB_FAST_REC64=1"),
    "uncond_smem": ("NDTB_FAST_UNCOND=1", "NDTB_FAST_ACC_SMEM=1"),
    "uncond_smem_pfp2_rec64": ("NDTB_FAST_UNCOND=1", "NDTB_FAST_ACC_SMEM=1", "NDTB_FAST_PFP=2", "NDTB_FAST_REC64=1"),
    "smem_pfp2": ("NDTB_FAST_ACC_SMEM=1", "NDTB_FAST_PFP=2"),
    "smem_pfp2_rec64": ("NDTB_FAST_ACC_SMEM=1", "NDTB_FAST_PFP=2", "NDTB_FAST_REC64=1"),
    "smem_pfp2_mb5": ("NDTB_FAST_ACC_SMEM=1", "NDTB_FAST_PFP=2", "NDTB_FAST_MINBLOCKS=5"),
    "smem_pfp2_rec64_mb5": ("NDTB_FAST_ACC_SMEM=1", "NDTB_FAST_PFP=2", "NDTB_FAST_REC64=1", "NDTB_FAST_MINBLOCKS=5"),
}


def main():
    names = sys.argv[1:] or list(VARIANTS)
    out = os.path.join(ROOT, "variants")
    os.makedirs(out, exist_ok=True)
    try:
        for n in names:
            b.build_cuda(defines=VARIANTS[n], lib=os.path.join(out, "lib_%s.so" % n))
            print("built variants/lib_%s.so  (%s)" % (n, " ".join(VARIANTS[n])))
    finally:
        b.build_cuda(force=True)
        print("in-tree library rebuilt with the default macros")


if __name__ == "__main__":
    main()
