"""Timeline of attention_v3_kernel (CTA 0, stream 0): builds librmb_b200_trace.so with -DRMB_ATTN_TRACE (never the product
library), runs rmb_attention at the ACT encoder shape and prints per-phase clock statistics.

    python scripts/attn_trace.py --build      (CPU container: compile only)
    python scripts/attn_trace.py [B]          (GPU box)
"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robomanipbaselines_b200 import build as _build  # noqa: E402

# --tag NAME picks librmb_b200_<NAME>.so; --flags "-DA=1 -DB" are extra nvcc flags of --build; --notrace builds without the log
TAG = sys.argv[sys.argv.index("--tag") + 1] if "--tag" in sys.argv else "trace"
TRACE_LIB = os.path.join(_build.PKG_DIR, f"librmb_b200_{TAG}.so")

if "--build" in sys.argv:
    cmd = _build._cmd(TRACE_LIB, False, False)
    extra = sys.argv[sys.argv.index("--flags") + 1].split() if "--flags" in sys.argv else []
    for fl in ([] if "--notrace" in sys.argv else ["-DRMB_ATTN_TRACE"]) + extra:
        cmd.insert(1, fl)
    subprocess.check_call(cmd)
    os.replace(TRACE_LIB + ".tmp", TRACE_LIB)
    print("built", TRACE_LIB)
    sys.exit(0)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import robomanipbaselines_b200._lib as L  # noqa: E402

L.LIB_PATH = TRACE_LIB
lib = L.load()
TRACED = hasattr(lib, "rmb_attention_trace_dump")
if TRACED:
    lib.rmb_attention_trace_dump.restype = ctypes.c_int
    lib.rmb_attention_trace_dump.argtypes = [ctypes.c_void_p, ctypes.c_int]
ACT = L.act_torch_dtype()
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
args = [a for i, a in enumerate(sys.argv[1:], 1) if a.isdigit() and sys.argv[i - 1] not in ("--tag", "--flags")]
B, H, D, Lq, Lk = (int(args[0]) if args else 256), 8, 512, 602, 602
QK = torch.randn(B * Lq, 2 * D, device=dev).to(ACT)
V = torch.randn(B * Lk, D, device=dev).to(ACT)
out = torch.zeros(B * Lq, D, device=dev, dtype=ACT)
p = lambda t: t.data_ptr()
run = lambda: L.check(lib.rmb_attention(p(QK), 2 * D, 0, 0, p(QK), 2 * D, D, p(V), D, 0, p(out), D, B, H, Lq, Lk, st), "attention")
buf = np.zeros(2 * 16384, dtype=np.int64)
for _ in range(3):
    run()
if not TRACED:                                     # plain timing of an experimental build
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    ref = torch.softmax((QK[:Lq, :64].float() @ QK[:Lk, D:D + 64].float().T) * 0.125, dim=1) @ V[:Lk, :64].float()
    err = (out[:Lq, :64].float() - ref).abs().max().item()
    print(f"[{TAG}] B {B} Lq {Lq} Lk {Lk}: {ms:.4f} ms  {4.0 * B * H * Lq * Lk * 64 / ms / 1e9:.1f} TFLOP/s; head 0 of episode 0 max-abs err vs fp32 {err:.3e}")
    sys.exit(0)
lib.rmb_attention_trace_dump(buf.ctypes.data, 2 * 16384)      # drop the warm-up launches
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); run(); e1.record()
lib.rmb_attention_trace_dump(buf.ctypes.data, 2 * 16384)
print(f"launch {e0.elapsed_time(e1):.4f} ms (with tracing)")
words = buf[buf != 0]
ev = np.stack([words >> 48, words & 0xffffffffffff], axis=1)
soft = ev[ev[:, 0] < 100]
iss = ev[ev[:, 0] >= 100]
t0 = ev[:, 1].min()
print(f"span {ev[:, 1].max() - t0} clocks; softmax events {len(soft)}, issuer events {len(iss)}")


def stats(name, seq, names):
    d = {}
    for (a, ta), (b, tb) in zip(seq[:-1], seq[1:]):
        d.setdefault((int(a), int(b)), []).append(int(tb - ta))
    print(name)
    tot = sum(sum(v) for v in d.values())
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        v = np.array(v)
        print(f"  {names.get(k[0], k[0])} -> {names.get(k[1], k[1])}: n {len(v):5d}  mean {v.mean():8.1f}  p50 {np.median(v):8.1f}  max {v.max():8d}  share {v.sum() / tot:6.1%}")


SN = {1: "S ready", 2: "S in regs", 3: "max done", 4: "exp01 done", 5: "PV(n-1) done", 6: "exp23 done", 8: "O written", 7: "P ready"}
IN = {101: "S consumed", 102: "S(n+1) issued", 103: "P ready", 104: "V landed", 105: "PV issued"}
stats("softmax warp 0 (stream 0, CTA 0):", soft, SN)
stats("issuer (stream 0, CTA 0):", iss, IN)
k = min(60, len(ev))
order = np.argsort(ev[:, 1], kind="stable")
print("first events (clock - t0, id):", [(int(ev[i, 1] - t0), int(ev[i, 0])) for i in order[:k]])
mid = order[len(order) // 2: len(order) // 2 + 40]
print("mid-run events:", [(int(ev[i, 1] - t0), int(ev[i, 0])) for i in mid])
