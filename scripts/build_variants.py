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
    # strict kernel: the default is three 128-thread blocks per SM with 25 of the 43 accumulators in registers; every other split,
    # block shape (64 x 6, 96 x 4, 192 x 2), the per-point vectors in shared memory, a hand-written double exponential, the
    # neighbour loop unrolled by two, the next source record prefetched, the cell codes kept in registers behind a bit mask and the
    # dense-index box in registers were all measured slower (profiles/r02/r02_c12 / c13 / c15 / c16_strict_variants.txt).  These rebuild earlier settings:
    "strict_r02": ("NDTB_THREADS=256", "NDTB_MINBLOCKS=2", "NDTB_STRICT_REG_ACCS=7"),
    "strict_t128_b3_r19": ("NDTB_STRICT_REG_ACCS=19",),
    "strict_t128_b3_r31": ("NDTB_STRICT_REG_ACCS=31",),
    "strict_r01": ("NDTB_THREADS=256", "NDTB_NOUNROLL=0", "NDTB_STRICT_ACC_SMEM=0", "NDTB_MINBLOCKS=1"),
    "strict_nounroll_regs": ("NDTB_THREADS=256", "NDTB_NOUNROLL=1", "NDTB_STRICT_ACC_SMEM=0", "NDTB_MINBLOCKS=1"),
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
