"""Build recipe for the sm_100a shared library (in-tree, so the .so travels with the repo snapshot)."""
import os
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "librmb_b200.so")
SOURCES = ["act_engine.cu"]
HEADERS = ["ptx.cuh", "gemm_tc.cuh", "gemm_plan.h", "tma_host.h", "attention_tc.cuh", "elementwise.cuh", "conv_halo.cuh", "diffusion_kernels.cuh", "diffusion_engine.cuh", "imgenc_engine.cuh", "gemm_gn.cuh", "pointcloud.cuh", "resize.cuh"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    deps.append(os.path.join(PKG_DIR, "..", "include", "rmb_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=True):
    """Compile csrc/*.cu for sm_100a into librmb_b200.so (cross-compiles without a GPU)."""
    if not force and not needs_build():
        return LIB_PATH
    cmd = [
        _nvcc(), "-shared", "-Xcompiler", "-fPIC", "-O3", "-std=c++17", "-lineinfo",
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-Xptxas", "-v" if verbose else "-O3",
        "-o", LIB_PATH,
    ] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building librmb_b200.so")
    if verbose:
        log = os.path.join(PKG_DIR, "build_ptxas.log")
        with open(log, "w") as f:
            f.write(res.stdout + res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
