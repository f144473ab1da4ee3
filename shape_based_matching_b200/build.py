"""In-tree build of the native pieces (no JIT cache: the .so files travel with the repo snapshot).

    python -m shape_based_matching_b200.build          # CUDA library + host C++ check + oracle
"""
import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
LIB = os.path.join(PKG, "libsbm_b200.so")
NVCC = os.environ.get("SBM_NVCC", "/usr/local/cuda/bin/nvcc")
# the environment may export CXX=/opt/gcc/bin/g++ (no libgomp): always use the system compiler
HOST_CXX = "/usr/bin/g++"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-ccbin", HOST_CXX, "-shared", "-Xcompiler", "-fPIC", "-Xptxas", "-v",
]


def _run(cmd, **kw):
    print("+", " ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, **kw)
    if r.returncode != 0:
        print(r.stdout)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    return r.stdout


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_cuda(force=False):
    srcs = [os.path.join(PKG, "csrc", f) for f in ("sbm_capi.cu", "sbm_kernels.cuh", "sbm_nccl.hpp")] + [
        os.path.join(ROOT, "include", "sbm_capi.h")
    ]
    if force or _stale(LIB, srcs):
        out = _run([NVCC] + NVCC_FLAGS + ["-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(PKG, "csrc"),
                                           srcs[0], "-o", LIB])
        with open(os.path.join(PKG, "csrc", "ptxas_info.txt"), "w") as f:
            f.write(out)
    return LIB


HOST_LIB = os.path.join(PKG, "libline2Dup_b200.so")
DRIVER = os.path.join(ROOT, "tests", "cpp", "dropin_driver")


def build_host(force=False):
    """The drop-in C++ API (include/line2Dup.h) over the C ABI, and the C++ test driver that uses it."""
    inc = os.path.join(ROOT, "include")
    src = os.path.join(PKG, "csrc", "line2Dup.cpp")
    src_c = os.path.join(PKG, "csrc", "line2Dup_c.cpp")
    deps = [src, src_c, os.path.join(inc, "line2Dup.h"), os.path.join(inc, "sbm_capi.h"), os.path.join(inc, "sbm", "cv_compat.hpp"),
            os.path.join(inc, "sbm", "imgproc_compat.hpp"), LIB]
    if force or _stale(HOST_LIB, deps):
        _run([HOST_CXX, "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-I" + inc, src, src_c, "-o", HOST_LIB, "-L" + PKG, "-lsbm_b200", "-lz",
              "-Wl,-rpath,$ORIGIN"])
    drv = os.path.join(ROOT, "tests", "cpp", "dropin_driver.cpp")
    if os.path.exists(drv) and (force or _stale(DRIVER, [drv, HOST_LIB])):
        _run([HOST_CXX, "-std=c++17", "-O2", "-Wall", "-I" + inc, drv, "-o", DRIVER, "-L" + PKG, "-lline2Dup_b200", "-lsbm_b200", "-lz",
              "-Wl,-rpath,$ORIGIN/../../shape_based_matching_b200"])
    gdrv = os.path.join(ROOT, "tests", "cpp", "group_driver.cpp")
    gbin = os.path.join(ROOT, "tests", "cpp", "group_driver")
    if os.path.exists(gdrv) and (force or _stale(gbin, [gdrv, HOST_LIB])):
        _run([HOST_CXX, "-std=c++17", "-O2", "-Wall", "-I" + inc, gdrv, "-o", gbin, "-L" + PKG, "-lline2Dup_b200", "-lsbm_b200", "-lz", "-pthread",
              "-Wl,-rpath,$ORIGIN/../../shape_based_matching_b200"])
    return HOST_LIB


def build_oracle(force=False):
    """Build the CHECKER (test infrastructure): the C restatement and, where the reference tree exists,
    the reference's own code (oracle/_ref).  Building it is not using it."""
    odir = os.path.join(ROOT, "oracle")
    port_src = os.path.join(odir, "sbm_oracle.cpp")
    port = os.path.join(odir, "libsbm_oracle.so")
    if os.path.exists(port_src) and (force or _stale(port, [port_src])):
        _run(["make", "-C", odir, "port"])
    ref = os.path.join(odir, "_ref", "libsbm_ref.so")
    if os.path.exists("/root/reference/line2Dup.cpp"):
        deps = [os.path.join(odir, "ref_driver.cpp"), os.path.join(odir, "shim", "opencv2", "imgproc.hpp"),
                os.path.join(ROOT, "include", "sbm", "cv_compat.hpp")]
        if force or _stale(ref, deps):
            _run(["make", "-C", odir, "ref"])


def build_all(force=False):
    build_cuda(force)
    build_host(force)
    build_oracle(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    print("ok:", LIB)
