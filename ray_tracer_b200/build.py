"""Builds ray_tracer_b200/libb200pt.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

-fmad=false: every float operator in the kernels is one IEEE-754 operation, so results are
reproducible against the CPU oracle (csrc/pt_math.h); box tests use explicit FMA intrinsics.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb200pt.so")
SOURCES = ["api.cu", "megakernel.cu", "wavefront.cu", "bvh2_build.cu", "wide_build.cu", "comm.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
              "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC,-O2,-fno-fast-math,-ffp-contract=off", "-shared",
              "-cudart", "static"]


def nvcc_path():
    for p in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if p and os.path.exists(p):
            return p
    raise RuntimeError("nvcc not found")


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "b200pt.h"), __file__]
    return any(os.path.getmtime(d) > t for d in deps)


HOST_DIR = os.path.join(HERE, "host")
HOST_LIB = os.path.join(HERE, "libb200pt_host.so")
HOST_SOURCES = ["scene_builder.cpp", "host_api.cpp"]


def build_host(force=False, verbose=False):
    """Host-side scene builder (plain C++, no CUDA): libb200pt_host.so."""
    srcs = [os.path.join(HOST_DIR, s) for s in HOST_SOURCES]
    deps = srcs + [os.path.join(HOST_DIR, "scene_builder.h"), os.path.join(CSRC, "pt_layout.h")]
    if not force and os.path.exists(HOST_LIB) and all(os.path.getmtime(d) <= os.path.getmtime(HOST_LIB) for d in deps):
        return HOST_LIB
    cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-shared", "-o", HOST_LIB] + srcs
    if verbose:
        print(" ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    return HOST_LIB


HEADLESS_BIN = os.path.join(HERE, "b200pt_headless")
HEADLESS_SOURCES = ["headless_main.cpp", "headless_engine.cpp"]


def build_headless(force=False, verbose=False):
    """b200pt_headless: the C++ host program above the C ABI (host/headless_engine.h), linked against the two
    libraries next to it (rpath $ORIGIN). Needs libb200pt.so and libb200pt_host.so to exist."""
    srcs = [os.path.join(HOST_DIR, s) for s in HEADLESS_SOURCES]
    deps = srcs + [os.path.join(HOST_DIR, "headless_engine.h"), os.path.join(HOST_DIR, "scene_builder.h"),
                   os.path.join(HERE, "..", "include", "b200pt.h"), LIB, HOST_LIB]
    if not force and os.path.exists(HEADLESS_BIN) and all(os.path.getmtime(d) <= os.path.getmtime(HEADLESS_BIN) for d in deps):
        return HEADLESS_BIN
    cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-o", HEADLESS_BIN] + srcs + \
          ["-L" + HERE, "-lb200pt", "-lb200pt_host", "-Wl,-rpath,$ORIGIN", "-pthread"]
    if verbose:
        print(" ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    return HEADLESS_BIN


def build(force=False, verbose=False, extra=(), out=None):
    """out: alternative output path (A/B variants with -D flags in `extra`; loaded with B200PT_LIB=<path>)."""
    build_host(force=force, verbose=verbose)
    if out is None and not force and not needs_build():
        return LIB
    env = dict(os.environ)
    env.pop("CXX", None)    # the image's $CXX points at a g++ without its spec files
    cmd = [nvcc_path()] + NVCC_FLAGS + list(extra) + ["-ccbin", "/usr/bin/g++", "-o", out or LIB] + [os.path.join(CSRC, s) for s in SOURCES] + ["-ldl"]
    if verbose:
        print(" ".join(cmd))
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout + r.stderr)
    return out or LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True, extra=["-Xptxas", "-v"] if "--ptxas" in sys.argv else ()))
    print(build_headless(force="--force" in sys.argv, verbose=True))
