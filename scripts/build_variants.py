"""Kernel variants for scripts/gpurun_variants.sh (the factorised-kernel variants of rounds 1-2 -- 64-byte records, unpredicated
loads, shared-memory accumulators, deeper prefetch -- were all measured slower than the default and removed from the sources:
profiles/r02/r02_c3_variants.txt): each entry is compiled into variants/lib_<name>.so (git-ignored, travels to the
GPU box), then the in-tree library is rebuilt with the default macros (the variant builds share its object paths).
    python scripts/build_variants.py [name ...]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from xchu_slam_b200 import build as b  # noqa: E402

VARIANTS = {
    # strict (term-by-term) kernel: the round-2 default is shared-memory FP64 accumulators + neighbour loop not unrolled + two
    # blocks per SM (profiles/r02/r02_c4_strict_variants.txt, r02_c5_strict_variants.txt); these rebuild the earlier settings
    # leading accumulators (score, gradient, first Hessian rows) in registers, the rest in shared memory: the final binary's
    # strict kernel is bound by its shared-memory traffic (L1/TEX 93 %, profiles/r02/r02_final_deriv_strict_full.txt)
    # strict kernel at three 128-thread blocks per SM (168 registers): most of the 43 accumulators fit in registers without spills
    "strict_t128_b3_r25": ("NDTB_THREADS=128", "NDTB_MINBLOCKS=3", "NDTB_STRICT_REG_ACCS=25"),
    "strict_t128_b3_r31": ("NDTB_THREADS=128", "NDTB_MINBLOCKS=3", "NDTB_STRICT_REG_ACCS=31"),
    "strict_t128_b3_r37": ("NDTB_THREADS=128", "NDTB_MINBLOCKS=3", "NDTB_STRICT_REG_ACCS=37"),
    "strict_t192_b2_r31": ("NDTB_THREADS=192", "NDTB_MINBLOCKS=2", "NDTB_STRICT_REG_ACCS=31"),
    # factorised kernel with the scalar core of every pair (w, q, exp, f) in the reference's float operations
    "fast_exactcore": ("NDTB_FAST_EXACT_CORE=1",),
    "strict_reg7": ("NDTB_STRICT_REG_ACCS=7",),
    "strict_reg13": ("NDTB_STRICT_REG_ACCS=13",),
    "strict_reg19": ("NDTB_STRICT_REG_ACCS=19",),
    "strict_r01": ("NDTB_NOUNROLL=0", "NDTB_STRICT_ACC_SMEM=0", "NDTB_MINBLOCKS=1"),
    "strict_nounroll_regs": ("NDTB_NOUNROLL=1", "NDTB_STRICT_ACC_SMEM=0", "NDTB_MINBLOCKS=1"),
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
