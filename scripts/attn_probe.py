"""Time rmb_attention at the ACT encoder shape for several key counts (fixed per-CTA cost vs per-key-block cost)."""
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from robomanipbaselines_b200._lib import load, check, act_torch_dtype

lib = load()
ACT = act_torch_dtype()
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
B, H, D, Lq = 256, 8, 512, 602
once = "--once" in sys.argv          # one shape, few launches: the command to run under ncu
variants = [int(a) for a in sys.argv[1:] if a.isdigit()] or [lib.rmb_set_attention_variant(0)]
if once:
    B = 64
for variant, Lk in [(v, l) for v in variants for l in ((602,) if once else (128, 256, 384, 512, 602, 640, 896))]:
    assert lib.rmb_set_attention_variant(variant) == variant
    QK = torch.randn(B * max(Lq, Lk), 2 * D, device=dev).to(ACT)
    V = torch.randn(B * Lk, D, device=dev).to(ACT)
    out = torch.zeros(B * Lq, D, device=dev, dtype=ACT)
    p = lambda t: t.data_ptr()
    run = lambda: check(lib.rmb_attention(p(QK), 2 * D, 0, 0, p(QK), 2 * D, D, p(V), D, 0, p(out), D, B, H, Lq, Lk, st), "attention")
    for _ in range(3):
        run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(1 if once else 10):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / (1 if once else 10)
    print(f"v{variant} Lq {Lq} Lk {Lk}: {ms:.4f} ms  {4.0 * B * H * Lq * Lk * 64 / ms / 1e9:.1f} TFLOP/s", flush=True)
