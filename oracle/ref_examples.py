"""TEST INFRASTRUCTURE: the example callers (examples/odom_sequence.cpp, examples/sc_sequence.cpp) compiled against the REFERENCE'S OWN
headers over the stand-in Eigen / PCL headers of oracle/ref_stubs/ -> examples/*_ref: the CPU arm of config C2 and the trajectory
/ loop-detection oracles of the drop-in headers.  Built only where /root/reference exists; the binaries travel to the GPU box.
Lives under oracle/ so that the product package has no path into the oracle tree."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
EXAMPLES = os.path.join(HERE, "..", "examples")


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


REFERENCE_INCLUDE = "/root/reference/xchu_mapping/include"


def build_example_reference(force=False):
    """TEST INFRASTRUCTURE: the same odom_sequence.cpp compiled against the REFERENCE'S OWN pclomp headers (over the
    stand-in Eigen/PCL headers of oracle/ref_stubs/) -> examples/odom_sequence_ref, the CPU arm of config C2 and the
    trajectory oracle of the drop-in header.  Built only where /root/reference exists; the binary travels to the GPU box.
    Returns the path, or None when it is neither buildable nor prebuilt."""
    src = os.path.join(EXAMPLES, "odom_sequence.cpp")
    exe = os.path.join(EXAMPLES, "odom_sequence_ref")
    stubs = os.path.join(HERE, "ref_stubs")
    if os.path.exists(os.path.join(REFERENCE_INCLUDE, "pclomp", "ndt_omp_impl.hpp")):
        deps = [src] + [os.path.join(d, f) for d, _, fs in os.walk(stubs) for f in fs] + [os.path.join(HERE, "thirdparty_math.h")]
        if force or _stale(exe, deps):
            subprocess.check_call(["g++", "-std=c++17", "-O2", "-fopenmp", "-ffp-contract=off", "-w", "-DODOM_USE_REFERENCE", "-I" + stubs,
                                   "-I" + REFERENCE_INCLUDE, src, "-o", exe])
    build_sc_example_reference(force)
    return exe if os.path.exists(exe) else None


def build_sc_example_reference(force=False):
    """TEST INFRASTRUCTURE: sc_sequence.cpp compiled over the REFERENCE'S OWN scancontext/Scancontext.{h,cpp} (stand-in headers of
    oracle/ref_stubs/) -> examples/sc_sequence_ref.  Built only where /root/reference exists; the binary travels to the GPU box."""
    src = os.path.join(EXAMPLES, "sc_sequence.cpp")
    exe = os.path.join(EXAMPLES, "sc_sequence_ref")
    stubs = os.path.join(HERE, "ref_stubs")
    if os.path.exists(os.path.join(REFERENCE_INCLUDE, "scancontext", "Scancontext.cpp")):
        deps = [src] + [os.path.join(d, f) for d, _, fs in os.walk(stubs) for f in fs]
        if force or _stale(exe, deps):
            subprocess.check_call(["g++", "-std=c++17", "-O2", "-fopenmp", "-ffp-contract=off", "-w", "-DSC_USE_REFERENCE", "-I" + stubs,
                                   "-I" + REFERENCE_INCLUDE, src, "-o", exe])
    return exe if os.path.exists(exe) else None




def build_shim_pcl_mode(force=False):
    """TEST INFRASTRUCTURE: tests/cpp/shim_pcl_mode.cpp -- include/pclomp/ndt_omp.h in its PCL build (derived from
    pcl::Registration), compiled over the stand-in PCL / Eigen headers of oracle/ref_stubs (whose kd-tree stand-in wraps the
    reference's nanoflann, hence the reference include path) and linked to libndt_b200.so -> examples/shim_pcl_mode.
    Built only where /root/reference exists; the binary travels to the GPU box."""
    root = os.path.join(HERE, "..")
    src = os.path.join(root, "tests", "cpp", "shim_pcl_mode.cpp")
    exe = os.path.join(EXAMPLES, "shim_pcl_mode")
    stubs = os.path.join(HERE, "ref_stubs")
    libdir = os.path.join(root, "xchu_slam_b200")
    hdr = os.path.join(root, "include", "pclomp", "ndt_omp.h")
    if os.path.exists(os.path.join(REFERENCE_INCLUDE, "scancontext", "nanoflann.hpp")):
        deps = [src, hdr, os.path.join(root, "include", "ndt_b200", "ndt_c_api.h")]
        if force or _stale(exe, deps):
            subprocess.check_call(["g++", "-std=c++17", "-O2", "-w", "-I" + os.path.join(root, "include"), "-I" + stubs,
                                   "-I" + REFERENCE_INCLUDE, src, "-o", exe, "-L" + libdir, "-lndt_b200",
                                   "-Wl,-rpath,$ORIGIN/../xchu_slam_b200"])
    return exe if os.path.exists(exe) else None
