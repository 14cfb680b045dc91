"""Builds the in-tree native libraries: libndt_b200.so (CUDA, sm_100a) and _lidar_synth.so.

nvcc cross-compiles without a GPU.  -fmad=false is part of the numerics contract (DESIGN.md
section 5): float expressions are evaluated as written, like the reference's x86-64 build.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libndt_b200.so")
SOURCES = ["voxel_build.cu", "ndt_core.cu", "ndt_fast.cu", "ndt_api.cu", "scancontext.cu", "no_alloc.cu"]
# ndt_fast.cu carries explicit round-to-nearest intrinsics where bits matter and wants FMA contraction elsewhere
NO_FMAD_FALSE = {"ndt_fast.cu"}
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC", "-ccbin", "g++"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_cuda(force=False, verbose=False, defines=(), lib=None):
    """defines: extra -D macros (kernel tuning variants); lib: output path override."""
    global LIB
    if lib:
        LIB_SAVE, LIB = LIB, lib
        try:
            return _build_cuda(True, verbose, defines)
        finally:
            LIB = LIB_SAVE
    return _build_cuda(force, verbose, defines)


def _build_cuda(force=False, verbose=False, defines=()):
    hdrs = [os.path.join(CSRC, "ndt_internal.h"), os.path.join(CSRC, "ndt_device.cuh"), os.path.join(HERE, "..", "include", "ndt_b200", "ndt_c_api.h"),
            os.path.join(HERE, "..", "include", "ndt_b200", "sc_c_api.h")]
    objs = []
    procs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(CSRC, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            flags = [f for f in NVCC_FLAGS if not (f == "-fmad=false" and src in NO_FMAD_FALSE)]
            cmd = ["nvcc"] + flags + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError("nvcc failed on " + src)
        if verbose:
            sys.stderr.write(out)
    if force or procs or _stale(LIB, objs):
        tmp = LIB + ".tmp%d" % os.getpid()     # linked beside the target, then renamed: a reader never sees a half-written library
        subprocess.check_call(["nvcc", "-shared", "-o", tmp] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                                       "-Xcompiler", "-fPIC", "-ccbin", "g++", "-lcudart"])
        os.replace(tmp, LIB)
    return LIB


EXAMPLES = os.path.join(HERE, "..", "examples")


def build_examples(force=False):
    """Host C++ above the C ABI: the headless mirror of odom_node's loop over include/pclomp/ndt_omp.h."""
    src = os.path.join(EXAMPLES, "odom_sequence.cpp")
    exe = os.path.join(EXAMPLES, "odom_sequence")
    deps = [src, LIB, os.path.join(HERE, "..", "include", "pclomp", "ndt_omp.h"), os.path.join(HERE, "..", "include", "pclomp", "pcl_compat.h"),
            os.path.join(HERE, "..", "include", "ndt_b200", "ndt_c_api.h")]
    if force or _stale(exe, deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", src, "-o", exe, "-L" + HERE, "-lndt_b200",
                               "-Wl,-rpath,$ORIGIN/../xchu_slam_b200"])
    build_sc_example(force)
    build_batch_pipeline(force)
    return exe


def build_batch_pipeline(force=False):
    """examples/batch_pipeline.cpp: the batch entry points driven from C++ (two-stage upload / compute pipeline)."""
    src = os.path.join(EXAMPLES, "batch_pipeline.cpp")
    exe = os.path.join(EXAMPLES, "batch_pipeline")
    deps = [src, LIB, os.path.join(HERE, "..", "include", "ndt_b200", "ndt_c_api.h")]
    if force or _stale(exe, deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-pthread", src, "-o", exe, "-L" + HERE, "-lndt_b200",
                               "-Wl,-rpath,$ORIGIN/../xchu_slam_b200"])
    return exe


def build_sc_example(force=False):
    """examples/sc_sequence.cpp over include/scancontext/Scancontext.h (drop-in for the reference's SCManager)."""
    inc = os.path.join(HERE, "..", "include")
    src = os.path.join(EXAMPLES, "sc_sequence.cpp")
    exe = os.path.join(EXAMPLES, "sc_sequence")
    deps = [src, LIB, os.path.join(inc, "scancontext", "Scancontext.h"), os.path.join(inc, "pclomp", "pcl_compat.h"),
            os.path.join(inc, "ndt_b200", "sc_c_api.h")]
    if force or _stale(exe, deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I" + inc, src, "-o", exe, "-L" + HERE, "-lndt_b200",
                               "-Wl,-rpath,$ORIGIN/../xchu_slam_b200"])
    return exe


def build_all(force=False, verbose=False):
    from . import synth
    synth.build(force)
    lib = build_cuda(force, verbose)
    build_examples(force)
    return lib


if __name__ == "__main__":
    sys.path.insert(0, os.path.dirname(HERE))
    build_cuda(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
