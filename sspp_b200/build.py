"""In-tree builds: nvcc for the CUDA engine (sm_100a only), g++ for the host C++ mirror."""
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(ROOT)
CSRC = os.path.join(ROOT, "csrc")
HOST = os.path.join(ROOT, "host")
LIB_GPU = os.path.join(CSRC, "libsspp_gpu.so")
LIB_HOST = os.path.join(HOST, "libsspp_host.so")

# -fmad=false / -ffp-contract=off: no multiply-add is ever fused, on device or host (bit parity
# with the reference's x86-64 build, CMakeLists.txt:6-8).  -lineinfo so ncu maps SASS to source.
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off", "-shared",
]
GXX = "/usr/bin/g++"  # the image's $CXX links libstdc++ statically, which breaks dlopen-ed C++ libraries


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def build_gpu(force=False, verbose=False):
    # every header of the one translation unit triggers a rebuild (a hand-kept list once missed dedup_kernels.cuh)
    srcs = [os.path.join(CSRC, "sspp_gpu.cu")] + sorted(
        os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))
    srcs.append(os.path.join(REPO, "include", "sspp_gpu.h"))
    if not force and _newer(LIB_GPU, srcs):
        return LIB_GPU
    extra = os.environ.get("SSPP_NVCC_EXTRA", "").split()  # e.g. -DEDGE_MIN_BLOCKS=2 for tuning runs
    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_GPU, srcs[0]]
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB_GPU


def build_host(force=False):
    src_dir = os.path.join(HOST, "src")
    if not os.path.isdir(src_dir):
        return None
    srcs = sorted(os.path.join(src_dir, f) for f in os.listdir(src_dir) if f.endswith(".cpp"))
    hdr_dir = os.path.join(HOST, "include", "sspp")
    deps = srcs + [os.path.join(hdr_dir, f) for f in os.listdir(hdr_dir)] + [os.path.join(REPO, "include", "sspp_gpu.h")]
    if not force and _newer(LIB_HOST, deps):
        return LIB_HOST
    cmd = [GXX, "-std=c++17", "-O2", "-frounding-math", "-ffp-contract=off", "-fPIC", "-shared", "-o", LIB_HOST,
           "-I", os.path.join(HOST, "include"), "-I", os.path.join(REPO, "include")] + srcs + [
               "-L", CSRC, "-lsspp_gpu", "-Wl,-rpath," + CSRC, "-Wl,-rpath,$ORIGIN/../csrc"]
    subprocess.run(cmd, check=True)
    return LIB_HOST


def build_all(force=False, verbose=False):
    build_gpu(force, verbose)
    build_host(force)
