"""Builds liborbx.so (the CUDA kernels + C ABI of include/orbx.h) in-tree with nvcc for sm_100a."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "liborbx.so")
SOURCES = ["extractor.cu", "k_pyramid.cu", "k_fast.cu", "k_quadtree.cu", "k_blur.cu", "k_describe.cu", "k_match.cu",
           "k_search.cu", "k_bow.cu", "k_color.cu", "k_undistort.cu", "k_stereo.cu", "matcher.cu", "vocabulary.cu", "noalloc.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--fmad=false",
              "-Xcompiler", "-fPIC,-O2,-Wall,-ffp-contract=off", "-Xptxas", "-v", "-cudart", "static"]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def build(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh", ".inc"))]
    deps.append(os.path.join(HERE, "..", "include", "orbx.h"))
    if not force and os.path.exists(SO) and all(os.path.getmtime(d) <= os.path.getmtime(SO) for d in deps):
        return SO
    objs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    procs = []
    for s in srcs:
        o = os.path.join(HERE, "build", os.path.basename(s) + ".o")
        objs.append(o)
        if not force and os.path.exists(o) and all(os.path.getmtime(d) <= os.path.getmtime(o) for d in deps if not d.endswith(".cu") or d == s):
            continue
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", s, "-o", o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        else:
            with open(os.path.join(HERE, "build", os.path.basename(s) + ".ptxas.log"), "w") as f:
                f.write(out)
        if p.returncode:
            raise RuntimeError(f"nvcc failed on {s}")
    cmd = [_nvcc(), "-shared", "-o", SO] + objs + ["-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a"]
    subprocess.check_call(cmd)
    return SO


def build_variant(name: str, defines: list) -> str:
    """Development aid: liborbx_<name>.so compiled with extra -D flags (select it with ORBX_LIB=<path>)."""
    out = os.path.join(HERE, f"liborbx_{name}.so")
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    flags = [f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")]
    subprocess.check_call([_nvcc()] + flags + ["-D" + d for d in defines] + ["-shared", "-o", out] + srcs)
    return out


CLASSES_SO = os.path.join(HERE, "liborbx_classes.so")
HARNESS = os.path.join(HERE, "..", "tests", "cpp", "frame_harness")


def build_host(force: bool = False) -> str:
    """The C++ class layer (ORB_SLAM2::ORBextractor over the C ABI) and the Frame-shaped harness of the tests."""
    inc = os.path.join(HERE, "..", "include")
    src = os.path.join(HERE, "host", "ORBextractor.cc")
    hdrs = [os.path.join(inc, f) for f in os.listdir(inc)]
    cxx = os.environ.get("CXX", "g++")
    flags = ["-std=c++17", "-O2", "-fPIC", "-Wall", "-Wextra", "-ffp-contract=off", "-I", inc]
    if force or not os.path.exists(CLASSES_SO) or any(os.path.getmtime(d) > os.path.getmtime(CLASSES_SO) for d in hdrs + [src, SO]):
        subprocess.check_call([cxx] + flags + ["-shared", src, "-o", CLASSES_SO, "-L", HERE, "-lorbx", "-Wl,-rpath,$ORIGIN"])
    hsrc = os.path.join(HERE, "..", "tests", "cpp", "frame_harness.cpp")
    if os.path.exists(hsrc) and (force or not os.path.exists(HARNESS) or
                                 any(os.path.getmtime(d) > os.path.getmtime(HARNESS) for d in hdrs + [hsrc, CLASSES_SO])):
        subprocess.check_call([cxx] + flags + [hsrc, "-o", HARNESS, "-L", HERE, "-lorbx_classes", "-lorbx", "-pthread",
                                               "-Wl,-rpath,$ORIGIN/../../orb_slam_cuda_b200"])
    # the same harness with the class layer's ORBX_HAVE_OPENCV branches switched on: the test-only <opencv2/...> of
    # oracle/ref_matcher/shim stands in for OpenCV (its cv::Mat expressions evaluate like OpenCV's CV_32F paths), and
    # host/ORBextractor.cc is compiled along because cv::Mat then has that header's layout
    shim = os.path.join(HERE, "..", "oracle", "ref_matcher", "shim")
    hcv = HARNESS + "_cv"
    if os.path.isdir(shim) and os.path.exists(hsrc) and (force or not os.path.exists(hcv) or any(
            os.path.getmtime(d) > os.path.getmtime(hcv) for d in hdrs + [hsrc, src, SO, os.path.join(shim, "opencv2", "core", "core.hpp")])):
        subprocess.check_call([cxx, "-std=c++17", "-O2", "-w", "-ffp-contract=off", "-pthread", "-I", shim, "-I", inc, hsrc, src, "-o", hcv,
                               "-L", HERE, "-lorbx", "-Wl,-rpath,$ORIGIN/../../orb_slam_cuda_b200"])
    return CLASSES_SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
