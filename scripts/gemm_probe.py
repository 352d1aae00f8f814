"""Time rmb_gemm on the ACT encoder GEMM shapes (CUDA events, L2 flushed by the size of the operands themselves)."""
import argparse
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from robomanipbaselines_b200._lib import load, check, act_torch_dtype

ap = argparse.ArgumentParser()
ap.add_argument("--shapes", default="154112x3200x512x1,154112x1024x512x0,154112x512x512x6,154112x512x3200x6,154112x512x512x0")
ap.add_argument("--iters", type=int, default=10)
a = ap.parse_args()
lib = load()
ACT = act_torch_dtype()
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
for s in a.shapes.split(","):
    M, N, K, flags = (int(v) for v in s.split("x"))
    A = torch.randn(M, K, device=dev).to(ACT)
    W = (torch.randn(N, K, device=dev) / K ** 0.5).to(ACT)
    bias = torch.randn(N, device=dev)
    res = torch.randn(M, N, device=dev) if flags & 4 else torch.randn(M, N, device=dev).to(ACT)
    out = torch.empty(M, N, device=dev, dtype=torch.float32 if flags & 2 else ACT)
    p = lambda t: t.data_ptr()
    if not flags & 4:
        res = None
    check(lib.rmb_gemm(p(A), p(W), p(bias), (p(res) if res is not None else None), p(out), M, N, K, flags, st), "gemm")
    torch.cuda.synchronize()
    worst = 0.0
    for r0 in (0, max(0, M // 2 - 1000), max(0, M - 2048)):       # head, middle and tail row blocks against torch
        sl = slice(r0, min(M, r0 + 2048))
        ref = A[sl].float() @ W.float().t() + bias
        if res is not None:
            ref = ref + res[sl].float()
        if flags & 1:
            ref = ref.relu()
        worst = max(worst, float((out[sl].float() - ref).abs().max() / ref.abs().max().clamp_min(1.0)))
    print(f"  max rel err vs torch: {worst:.3e}", flush=True)
    for _ in range(3):
        check(lib.rmb_gemm(p(A), p(W), p(bias), (p(res) if res is not None else None), p(out), M, N, K, flags, st), "gemm")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(a.iters):
        check(lib.rmb_gemm(p(A), p(W), p(bias), (p(res) if res is not None else None), p(out), M, N, K, flags, st), "gemm")
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    print(f"{M}x{N}x{K} flags={flags}: {ms:.4f} ms  {2.0 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)
