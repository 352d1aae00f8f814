"""Build recipe for the sm_100a shared library (in-tree, so the .so travels with the repo snapshot)."""
import os
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "librmb_b200.so")             # product default: IEEE fp16 operands
LIB_PATH_BF16 = os.path.join(PKG_DIR, "librmb_b200_bf16.so")   # same sources with -DRMB_ACT_BF16 (RMB_LIB_DTYPE=bf16)
SOURCES = ["act_engine.cu"]
HEADERS = ["ptx.cuh", "gemm_tc.cuh", "gemm_plan.h", "tma_host.h", "attention_tc.cuh", "elementwise.cuh", "conv_halo.cuh", "diffusion_kernels.cuh", "diffusion_engine.cuh", "imgenc_engine.cuh", "gemm_gn.cuh", "pointcloud.cuh", "resize.cuh", "rollout_glue.cuh", "launch.h", "stem_pool.cuh", "gemm_ln.cuh", "attention_v2.cuh", "attention_v3.cuh", "attention_v4.cuh", "ik.cuh"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build(lib_path=LIB_PATH):
    if not os.path.exists(lib_path):
        return True
    t = os.path.getmtime(lib_path)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    deps.append(os.path.join(PKG_DIR, "..", "include", "rmb_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _cmd(lib_path, bf16, verbose):
    return [
        _nvcc(), "-shared", "-Xcompiler", "-fPIC", "-O3", "-std=c++17", "-lineinfo",
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-Xptxas", "-v" if verbose else "-O3",
    ] + (["-DRMB_ACT_BF16"] if bf16 else []) + ["-o", lib_path + ".tmp"] + [os.path.join(CSRC, s) for s in SOURCES]


def build(force=False, verbose=True, bf16_too=True, stage_only=False):
    """Compile csrc/*.cu for sm_100a into librmb_b200.so and (bf16_too) librmb_b200_bf16.so, side by side
    (cross-compiles without a GPU; the two nvcc runs go in parallel). Each library is written to <name>.tmp and moved
    into place atomically; stage_only leaves the .tmp files for a later commit_staged()."""
    jobs = []
    for lib_path, bf16 in ((LIB_PATH, False), (LIB_PATH_BF16, True)):
        if bf16 and not bf16_too:
            continue
        if force or needs_build(lib_path):
            jobs.append((lib_path, bf16, subprocess.Popen(_cmd(lib_path, bf16, verbose), stdout=subprocess.PIPE,
                                                           stderr=subprocess.STDOUT, text=True)))
    for lib_path, bf16, proc in jobs:
        out, _ = proc.communicate()
        if proc.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError(f"nvcc failed building {os.path.basename(lib_path)}")
        if verbose:
            with open(os.path.join(PKG_DIR, "build_ptxas_bf16.log" if bf16 else "build_ptxas.log"), "w") as f:
                f.write(out)
    if not stage_only:
        commit_staged()
    return LIB_PATH


def commit_staged():
    for lib_path in (LIB_PATH, LIB_PATH_BF16):
        if os.path.exists(lib_path + ".tmp"):
            os.replace(lib_path + ".tmp", lib_path)


if __name__ == "__main__":
    if "--commit" in sys.argv:
        commit_staged()
    else:
        print(build(force="--force" in sys.argv, bf16_too="--no-bf16" not in sys.argv, stage_only="--stage" in sys.argv))
