"""GPU parity tests of the individual sm_100a kernels, called through the C ABI (ctypes), against plain PyTorch fp32
references of the same op computed from the same bf16-rounded inputs."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from robomanipbaselines_b200 import _lib
from robomanipbaselines_b200._lib import check, cur_stream, ptr

pytestmark = pytest.mark.gpu


def _act_dtype():
    return _lib.act_torch_dtype()


def _bf16(t):
    """Round to the library's 16-bit operand type (fp16 by default, bf16 with -DRMB_ACT_BF16)."""
    return t.to(_act_dtype())


def _report(name, got, ref):
    err = (got.float() - ref.float()).abs().max().item()
    scale = ref.float().abs().max().item()
    print(f"[{name}] max-abs err {err:.4e} (ref max {scale:.3e})")
    return err, scale


@pytest.mark.parametrize("M,N,K,flags", [
    (128, 64, 64, 0), (128, 128, 64, 0), (128, 256, 128, 0), (300, 512, 512, 1), (1000, 3200, 512, 1),
    (1204, 512, 3200, 2 | 4), (77, 1024, 512, 2), (4096, 512, 512, 0), (20000, 256, 576, 1),
    # CTA-pair launches (a tile per SM): even / odd row-tile counts, ragged last column tile (N = 3200, 1088), residuals
    (19000, 256, 128, 0), (4700, 3200, 512, 1), (38001, 64, 64, 2), (9000, 1088, 192, 2 | 4), (5000, 1024, 256, 0),
    # bit 3: no residual -> 16-bit outputs go through the TMA-store epilogue (single CTA and pair, ragged M and N)
    (300, 512, 512, 8 | 1), (129, 128, 64, 8), (1000, 3200, 512, 8 | 1), (4700, 3200, 512, 8 | 1), (19000, 256, 128, 8),
    (38001, 64, 64, 8), (77, 1024, 512, 8), (9000, 1088, 192, 8), (5000, 1024, 256, 8 | 2),
])
def test_gemm(lib, cuda, M, N, K, flags):
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K)
    A = _bf16(torch.randn(M, K, generator=g)).to(cuda)
    W = _bf16(torch.randn(N, K, generator=g) / K ** 0.5).to(cuda)
    bias = torch.randn(N, generator=g).to(cuda)
    res_f32 = bool(flags & 4)
    res = torch.randn(M, N, generator=g).to(cuda)
    if not res_f32:
        res = _bf16(res)
    out = torch.empty(M, N, device=cuda, dtype=torch.float32 if flags & 2 else _act_dtype())
    use_res = not (flags & 8)
    out.fill_(-77.0)
    check(lib.rmb_gemm(ptr(A), ptr(W), ptr(bias), ptr(res) if use_res else None, ptr(out), M, N, K, flags & 7, cur_stream()), "gemm")
    torch.cuda.synchronize()
    ref = A.float() @ W.float().t() + bias + (res.float() if use_res else 0.0)
    if flags & 1:
        ref = ref.relu()
    err, scale = _report(f"gemm {M}x{N}x{K} f{flags}", out, ref)
    assert err <= 2e-2 * max(scale, 1.0)


@pytest.mark.parametrize("M,K,with_pos", [(602 * 40, 512, False), (602 * 33 + 77, 3200, True), (300, 512, True), (128, 64, False)])
def test_gemm_residual_layernorm(lib, cuda, M, K, with_pos):
    """The fused out-projection / FFN-down + residual + LayerNorm launch vs fp32 torch on the same 16-bit operands:
    fp32 residual stream within 1e-4 (the accuracy of the stand-alone LayerNorm kernel), 16-bit copies within one
    rounding step; ragged last row tile, odd number of row tiles (one CTA of the last pair idle), tiny M."""
    g = torch.Generator(device="cpu").manual_seed(M + K)
    L = 602
    A = _bf16(torch.randn(M, K, generator=g)).to(cuda)
    W = _bf16(torch.randn(512, K, generator=g) / K ** 0.5).to(cuda)
    bias = torch.randn(512, generator=g).to(cuda)
    res = torch.randn(M, 512, generator=g).to(cuda)
    gamma = (1 + 0.1 * torch.randn(512, generator=g)).to(cuda)
    beta = (0.1 * torch.randn(512, generator=g)).to(cuda)
    pos = torch.randn(L, 512, generator=g).to(cuda)
    out = torch.full((M, 512), -77.0, device=cuda)
    y_act = torch.full((M, 512), -7.0, device=cuda, dtype=_act_dtype())
    ypos = torch.full((M, 512), -7.0, device=cuda, dtype=_act_dtype()) if with_pos else None
    check(lib.rmb_gemm_residual_layernorm(ptr(A), ptr(W), ptr(bias), ptr(res), ptr(gamma), ptr(beta), ptr(pos) if with_pos else None,
                                          L, ptr(out), ptr(y_act), ptr(ypos), M, K, cur_stream()), "gemm_ln")
    torch.cuda.synchronize()
    v = A.double() @ W.double().t() + bias.double() + res.double()
    ref = F.layer_norm(v, (512,), gamma.double(), beta.double(), 1e-5)
    err = (out.double() - ref).abs().max().item()
    print(f"[gemm+res+LN {M}x512x{K}] fp32 stream max-abs err {err:.3e}")
    assert err <= 1e-4
    step = 2.0 ** -8 if _act_dtype() == torch.bfloat16 else 2.0 ** -11
    assert (y_act.double() - ref).abs().max().item() <= step * ref.abs().max().item() + 1e-4
    if with_pos:
        refp = ref + pos.double()[torch.arange(M, device=cuda) % L]
        assert (ypos.double() - refp).abs().max().item() <= step * refp.abs().max().item() + 1e-4


def _conv_ref(x_nhwc, w_oihw, bias, res_nhwc, stride, relu):
    y = F.conv2d(x_nhwc.float().permute(0, 3, 1, 2), w_oihw.float(), bias, stride=stride, padding=w_oihw.shape[-1] // 2)
    y = y.permute(0, 2, 3, 1)
    if res_nhwc is not None:
        y = y + res_nhwc.float()
    return y.relu() if relu else y


@pytest.mark.parametrize("B,H,W,Cin,Cout,k,stride,relu,use_res", [
    (2, 16, 16, 64, 64, 3, 1, 1, 0), (3, 30, 40, 64, 128, 3, 2, 1, 0), (3, 30, 40, 64, 128, 1, 2, 0, 0),
    (2, 15, 20, 512, 512, 3, 1, 1, 1), (5, 15, 20, 512, 512, 1, 1, 0, 0), (2, 60, 80, 128, 128, 3, 1, 1, 1),
    (2, 120, 160, 64, 64, 3, 1, 1, 1), (2, 60, 80, 128, 256, 3, 2, 1, 0), (3, 30, 40, 256, 512, 3, 2, 1, 0),
    # enough row tiles for CTA pairs: 3x3 stride-1 convs with 128 output channels take the halo3 stages (gemm_tc.cuh)
    (8, 60, 80, 128, 128, 3, 1, 1, 1), (9, 60, 80, 128, 128, 3, 1, 1, 0), (40, 28, 36, 128, 128, 3, 1, 0, 1),
])
def test_conv(lib, cuda, B, H, W, Cin, Cout, k, stride, relu, use_res):
    g = torch.Generator(device="cpu").manual_seed(B + H + W + Cin + Cout + k + stride)
    x = _bf16(torch.randn(B, H, W, Cin, generator=g)).to(cuda)
    w = _bf16(torch.randn(Cout, Cin, k, k, generator=g) / (Cin * k * k) ** 0.5).to(cuda)
    bias = torch.randn(Cout, generator=g).to(cuda)
    OH, OW = (H + 2 * (k // 2) - k) // stride + 1, (W + 2 * (k // 2) - k) // stride + 1
    res = _bf16(torch.randn(B, OH, OW, Cout, generator=g)).to(cuda) if use_res else None
    wt = w.permute(0, 2, 3, 1).contiguous().reshape(Cout, k * k * Cin)      # [Cout][kh][kw][Cin]
    out = torch.zeros(B, OH, OW, Cout, device=cuda, dtype=_act_dtype())
    check(lib.rmb_conv2d_nhwc(ptr(x), ptr(wt), ptr(bias), ptr(res), ptr(out), B, H, W, Cin, Cout, k, stride,
                                   relu, cur_stream()), "conv")
    torch.cuda.synchronize()
    ref = _conv_ref(x, w, bias, res, stride, relu)
    err, scale = _report(f"conv B{B} {H}x{W} {Cin}->{Cout} k{k} s{stride}", out, ref)
    assert err <= 2e-2 * max(scale, 1.0)


def _stem_weight(w_oihw):
    """[64,3,7,7] -> [64, 7, 8, 4] with slot 0 / channel 3 zero (gemm_plan.h plan_conv1)."""
    co = w_oihw.shape[0]
    r = torch.zeros(co, 7, 8, 4, dtype=w_oihw.dtype, device=w_oihw.device)
    r[:, :, 1:, :3] = w_oihw.permute(0, 2, 3, 1)
    return r.reshape(co, 7 * 32).contiguous()


@pytest.mark.parametrize("B,H,W", [(2, 64, 96), (1, 480, 640)])
def test_preprocess_and_stem(lib, cuda, B, H, W):
    g = torch.Generator(device="cpu").manual_seed(H)
    img = torch.randint(0, 256, (B, H, W, 3), generator=g, dtype=torch.uint8).to(cuda)
    x0 = torch.full((B, H, W + 8, 4), 7.0, device=cuda, dtype=_act_dtype())
    check(lib.rmb_preprocess_images_u8(ptr(img), ptr(x0), B, H, W, cur_stream()), "preprocess_u8")
    mean = torch.tensor([0.485, 0.456, 0.406], device=cuda)
    std = torch.tensor([0.229, 0.224, 0.225], device=cuda)
    xn = (img.float() / 255.0 - mean) / std
    torch.cuda.synchronize()
    assert (x0[:, :, :4] == 0).all() and (x0[:, :, -4:] == 0).all() and (x0[..., 3] == 0).all()
    err, _ = _report("preprocess_u8", x0[:, :, 4:-4, :3], xn)
    assert err <= 2e-2
    # the f32 CHW entry point must produce the same buffer
    x0b = torch.empty_like(x0)
    chw = (img.float() / 255.0).permute(0, 3, 1, 2).contiguous()
    check(lib.rmb_preprocess_images_f32(ptr(chw), ptr(x0b), B, H, W, cur_stream()), "preprocess_f32")
    torch.cuda.synchronize()
    assert (x0b.float() - x0.float()).abs().max().item() <= 2e-2

    w = _bf16(torch.randn(64, 3, 7, 7, generator=g) / 147 ** 0.5).to(cuda)
    bias = torch.randn(64, generator=g).to(cuda)
    out = torch.zeros(B, H // 2, W // 2, 64, device=cuda, dtype=_act_dtype())
    check(lib.rmb_conv1_stem(ptr(x0), ptr(_stem_weight(w)), ptr(bias), ptr(out), B, H, W, cur_stream()), "stem")
    torch.cuda.synchronize()
    ref = F.conv2d(x0[:, :, 4:-4, :3].float().permute(0, 3, 1, 2), w.float(), bias, stride=2, padding=3).relu()
    err, scale = _report(f"stem {H}x{W}", out, ref.permute(0, 2, 3, 1))
    assert err <= 2e-2 * max(scale, 1.0)


@pytest.mark.parametrize("B,H,W", [(2, 64, 96), (1, 96, 160), (5, 256, 192), (3, 480, 640), (300, 64, 64)])
def test_stem_pool_fused_is_bit_identical_to_stem_then_pool(lib, cuda, B, H, W):
    """stem_pool_kernel (the launch the engine uses) == conv1 stem kernel followed by the max-pool kernel, bit for bit:
    ragged last strip (W/2 = 80 = 2.5 strips), one band per image and several (warm-up tile rows, carries across
    strips), more bands than epilogue groups (300 images), and vs the fp32 torch reference."""
    g = torch.Generator(device="cpu").manual_seed(H + W)
    img = torch.randint(0, 256, (B, H, W, 3), generator=g, dtype=torch.uint8).to(cuda)
    x0 = torch.empty(B, H, W + 8, 4, device=cuda, dtype=_act_dtype())
    check(lib.rmb_preprocess_images_u8(ptr(img), ptr(x0), B, H, W, cur_stream()), "preprocess_u8")
    w = _bf16(torch.randn(64, 3, 7, 7, generator=g) / 147 ** 0.5).to(cuda)
    bias = torch.randn(64, generator=g).to(cuda)
    wt = _stem_weight(w)
    c1 = torch.zeros(B, H // 2, W // 2, 64, device=cuda, dtype=_act_dtype())
    check(lib.rmb_conv1_stem(ptr(x0), ptr(wt), ptr(bias), ptr(c1), B, H, W, cur_stream()), "stem")
    pooled = torch.zeros(B, H // 4, W // 4, 64, device=cuda, dtype=_act_dtype())
    check(lib.rmb_maxpool3x3s2_nhwc(ptr(c1), ptr(pooled), B, H // 2, W // 2, 64, cur_stream()), "maxpool")
    fused = torch.full((B, H // 4, W // 4, 64), -7.0, device=cuda, dtype=_act_dtype())
    check(lib.rmb_stem_pool(ptr(x0), ptr(wt), ptr(bias), ptr(fused), B, H, W, cur_stream()), "stem_pool")
    torch.cuda.synchronize()
    assert torch.equal(fused.view(torch.int16), pooled.view(torch.int16))
    if B <= 5:
        ref = F.conv2d(x0[:, :, 4:-4, :3].float().permute(0, 3, 1, 2), w.float(), bias, stride=2, padding=3).relu()
        ref = F.max_pool2d(ref, 3, 2, 1).permute(0, 2, 3, 1)
        err, scale = _report(f"stem+pool {B}x{H}x{W}", fused, ref)
        assert err <= 4e-3 * max(scale, 1.0)


def test_maxpool(lib, cuda):
    x = _bf16(torch.randn(3, 24, 32, 64)).to(cuda)
    out = torch.empty(3, 12, 16, 64, device=cuda, dtype=_act_dtype())
    check(lib.rmb_maxpool3x3s2_nhwc(ptr(x), ptr(out), 3, 24, 32, 64, cur_stream()), "maxpool")
    torch.cuda.synchronize()
    ref = F.max_pool2d(x.float().permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1)
    assert torch.equal(out.float(), ref)


@pytest.fixture(params=[1, 2, 3, 4], ids=["v1", "v2", "v3", "v4"])
def attn_variant(lib, request):
    """Every attention kernel variant (csrc/act_engine.cu::launch_attention); restores the library default afterwards."""
    default = lib.rmb_set_attention_variant(0)
    assert lib.rmb_set_attention_variant(request.param) == request.param
    yield request.param
    lib.rmb_set_attention_variant(default)


@pytest.mark.parametrize("B,Lq,Lk,shared", [(2, 128, 128, 0), (2, 302, 302, 0), (3, 602, 602, 0), (3, 100, 602, 1),
                                             (1, 902, 902, 0), (2, 77, 45, 0), (40, 602, 602, 0)])
def test_attention(lib, cuda, attn_variant, B, Lq, Lk, shared):
    g = torch.Generator(device="cpu").manual_seed(Lq + Lk)
    H, D = 8, 512
    qrows = Lq if shared else B * Lq
    QK = _bf16(torch.randn(max(qrows, B * Lk), 2 * D, generator=g)).to(cuda)   # q at cols 0.., k at cols 512..
    Qs = _bf16(torch.randn(qrows, D, generator=g)).to(cuda)
    V = _bf16(torch.randn(B * Lk, D, generator=g)).to(cuda)
    out = torch.zeros(B * Lq, D, device=cuda, dtype=_act_dtype())
    if shared:
        check(lib.rmb_attention(ptr(Qs), D, 0, 1, ptr(QK), 2 * D, D, ptr(V), D, 0, ptr(out), D, B, H, Lq, Lk,
                                     cur_stream()), "attention")
        q = Qs.float().view(1, Lq, H, 64).expand(B, Lq, H, 64)
    else:
        check(lib.rmb_attention(ptr(QK), 2 * D, 0, 0, ptr(QK), 2 * D, D, ptr(V), D, 0, ptr(out), D, B, H, Lq, Lk,
                                     cur_stream()), "attention")
        q = QK[:B * Lq, :D].float().view(B, Lq, H, 64)
    torch.cuda.synchronize()
    k = QK[:B * Lk, D:].float().view(B, Lk, H, 64)
    v = V.float().view(B, Lk, H, 64)
    s = torch.einsum("bqhd,bkhd->bhqk", q, k) / 8.0
    ref = torch.einsum("bhqk,bkhd->bqhd", s.softmax(-1), v).reshape(B * Lq, D)
    err, scale = _report(f"attention B{B} Lq{Lq} Lk{Lk}", out, ref)
    assert err <= 2e-2 * max(scale, 1.0)


@pytest.mark.parametrize("B,Lq,Lk", [(2, 602, 602), (40, 602, 602), (3, 100, 602), (2, 130, 257), (300, 20, 602)])
def test_attention_online_softmax_reference_point_moves(lib, cuda, attn_variant, B, Lq, Lk):
    """One-sweep attention keeps a reference maximum per row and only moves it (rescaling O and the row sum in TMEM)
    when a key block exceeds it by more than 2^8. Here the logits GROW along the keys (later blocks are up to 6x larger),
    so every block forces the move; plus a flat case mix: more work items than streams (40 x 8 x 5 = 1600 items on 296
    streams), a single query tile with dead warps (Lq = 20, 100), ragged key blocks (Lk = 257: one valid key in the last
    block). P is 16-bit, so errors are relative to the softmax's own scale."""
    g = torch.Generator(device="cpu").manual_seed(B + Lq + Lk)
    H, D = 8, 512
    Q = _bf16(torch.randn(B * Lq, D, generator=g) * 1.5).to(cuda)
    grow = (1.0 + 5.0 * torch.arange(Lk).float() / Lk).repeat(B).unsqueeze(1)
    K = _bf16(torch.randn(B * Lk, D, generator=g) * grow).to(cuda)
    V = _bf16(torch.randn(B * Lk, D, generator=g)).to(cuda)
    out = torch.full((B * Lq, D), 7.0, device=cuda, dtype=_act_dtype())
    check(lib.rmb_attention(ptr(Q), D, 0, 0, ptr(K), D, 0, ptr(V), D, 0, ptr(out), D, B, H, Lq, Lk, cur_stream()), "attention")
    torch.cuda.synchronize()
    q = Q.float().view(B, Lq, H, 64)
    k = K.float().view(B, Lk, H, 64)
    v = V.float().view(B, Lk, H, 64)
    s = torch.einsum("bqhd,bkhd->bhqk", q, k) / 8.0
    assert (s.amax(-1) - s[..., :128].amax(-1)).max().item() * 1.4427 > 8.0      # the move really is exercised
    ref = torch.einsum("bhqk,bkhd->bqhd", s.softmax(-1), v).reshape(B * Lq, D)
    err, scale = _report(f"attention (growing logits) B{B} Lq{Lq} Lk{Lk}", out, ref)
    assert torch.isfinite(out.float()).all() and err <= 1e-2 * max(scale, 1.0)


def test_layernorm(lib, cuda):
    rows, L = 1204, 602
    x = (torch.randn(rows, 512) * 3 + 1).to(cuda)
    gamma, beta = torch.randn(512).to(cuda), torch.randn(512).to(cuda)
    pos = torch.randn(L, 512).to(cuda)
    y = torch.empty(rows, 512, device=cuda)
    yb = torch.empty(rows, 512, device=cuda, dtype=_act_dtype())
    ypb = torch.empty(rows, 512, device=cuda, dtype=_act_dtype())
    check(lib.rmb_layernorm512(ptr(x), ptr(gamma), ptr(beta), ptr(pos), L, ptr(y), ptr(yb), ptr(ypb), rows, 1,
                               cur_stream()), "layernorm")
    torch.cuda.synchronize()
    ref = F.layer_norm(x, (512,), gamma, beta, 1e-5)
    assert (y - ref).abs().max().item() <= 1e-4
    assert (yb.float() - ref).abs().max().item() <= 4e-2
    assert (ypb.float() - (ref + pos.repeat(2, 1))).abs().max().item() <= 6e-2
    check(lib.rmb_layernorm512(ptr(x), None, None, ptr(pos), L, ptr(y), None, ptr(ypb), rows, 0, cur_stream()), "cast")
    torch.cuda.synchronize()
    assert torch.equal(y, x)
    assert (ypb.float() - (x + pos.repeat(2, 1))).abs().max().item() <= 8e-2


def test_layernorm_16bit_rows(lib, cuda):
    """layernorm512_kernel_t<act_t> (the encoder's LayerNorms in the fp16 build read 16-bit pre-norm sums): against
    F.layer_norm of the SAME 16-bit values in fp32, so only the arithmetic is compared (<= 1e-4 on the fp32 output)."""
    rows, L = 2411, 602           # odd row count: the two-rows-in-flight loop ends on a single row in some warps
    x16 = (torch.randn(rows, 512) * 3 + 1).to(cuda).to(_act_dtype())
    gamma, beta = torch.randn(512).to(cuda), torch.randn(512).to(cuda)
    pos = torch.randn(L, 512).to(cuda)
    y = torch.empty(rows, 512, device=cuda)
    yb = torch.empty(rows, 512, device=cuda, dtype=_act_dtype())
    ypb = torch.empty(rows, 512, device=cuda, dtype=_act_dtype())
    check(lib.rmb_layernorm512_act(ptr(x16), ptr(gamma), ptr(beta), ptr(pos), L, ptr(y), ptr(yb), ptr(ypb), rows, 1,
                                   cur_stream()), "layernorm_act")
    torch.cuda.synchronize()
    ref = F.layer_norm(x16.float(), (512,), gamma, beta, 1e-5)
    posr = pos.repeat((rows + L - 1) // L, 1)[:rows]
    assert (y - ref).abs().max().item() <= 1e-4
    assert (yb.float() - ref).abs().max().item() <= 4e-2
    assert (ypb.float() - (ref + posr)).abs().max().item() <= 6e-2
    y2 = torch.empty_like(y)
    check(lib.rmb_layernorm512(ptr(x16.float().contiguous()), ptr(gamma), ptr(beta), ptr(pos), L, ptr(y2), None, None, rows, 1,
                               cur_stream()), "layernorm")
    torch.cuda.synchronize()
    assert torch.equal(y, y2)     # same arithmetic as the fp32-row instance on the same values


@pytest.mark.parametrize("norm", ["gaussian", "limits"])
def test_ensemble_bit_exact(lib, cuda, norm):
    """Ring positions, step counters and float64 actions must equal the numpy oracle exactly (C=7, wraps 3 times)."""
    from oracle.rollout_oracle import ActRolloutOracle
    from robomanipbaselines_b200.common.ensemble import make_weight_table

    B, C, Da, steps = 5, 7, 3, 25
    rng = np.random.default_rng(3)
    stats = {"mean": rng.normal(size=Da), "std": rng.uniform(0.5, 2.0, size=Da), "min": rng.normal(size=Da),
             "range": rng.uniform(0.5, 2.0, size=Da)}
    if norm == "limits":
        stats["norm_config"] = {"type": "limits", "out_min": -1.0, "out_max": 1.0}
        p0 = stats["range"] / (1.0 - (-1.0))
        p1, out_min, ntype = stats["min"], -1.0, 1
    else:
        p0, p1, out_min, ntype = stats["std"], stats["mean"], 0.0, 0
    oracles = [ActRolloutOracle(C, Da, stats) for _ in range(B)]
    ring = torch.zeros(B, C, C, Da, device=cuda)
    step = torch.zeros(B, dtype=torch.int32, device=cuda)
    wtab = torch.tensor(make_weight_table(C), device=cuda)
    d_p0, d_p1 = torch.tensor(p0, device=cuda), torch.tensor(p1, device=cuda)
    act = torch.zeros(B, Da, dtype=torch.float64, device=cuda)
    actn = torch.zeros(B, Da, dtype=torch.float64, device=cuda)
    n_calls = [0] * B
    for t in range(steps):
        chunks = rng.normal(size=(B, C, Da)).astype(np.float32)
        reset = np.zeros(B, dtype=np.uint8)
        if t == 11:
            reset[2] = 1          # episode 2 restarts mid-way
            oracles[2].reset()
            n_calls[2] = 0
        d_chunks = torch.tensor(chunks, device=cuda)
        d_reset = torch.tensor(reset, device=cuda)
        check(lib.rmb_ensemble_step(ptr(ring), ptr(d_chunks), ptr(step), ptr(d_reset), ptr(wtab), ntype, ptr(d_p0),
                                    ptr(d_p1), out_min, ptr(act), None, ptr(actn), B, C, Da, cur_stream()), "ensemble")
        torch.cuda.synchronize()
        for b in range(B):
            ref = oracles[b].step(chunks[b])
            n_calls[b] += 1
            assert np.array_equal(act[b].cpu().numpy(), ref), (t, b)
            assert np.array_equal(actn[b].cpu().numpy(), oracles[b].normalized_action), (t, b)
            oring, filled = oracles[b].ring_view(n_calls[b])
            assert np.array_equal(ring[b].cpu().numpy()[filled], oring[filled]), (t, b)
        assert step.cpu().tolist() == n_calls


def test_normalize_bit_exact(lib, cuda):
    from oracle.rollout_oracle import denormalize_data, normalize_data

    rng = np.random.default_rng(5)
    rows, D = 33, 7
    x = rng.normal(size=(rows, D))
    stats = {"mean": rng.normal(size=D), "std": rng.uniform(0.1, 2, size=D), "min": rng.normal(size=D),
             "range": rng.uniform(0.1, 2, size=D)}
    dx = torch.tensor(x, device=cuda)
    y32 = torch.empty(rows, D, device=cuda)
    y64 = torch.empty(rows, D, device=cuda, dtype=torch.float64)
    mean, std = torch.tensor(stats["mean"], device=cuda), torch.tensor(stats["std"], device=cuda)
    check(lib.rmb_normalize(ptr(dx), 0, ptr(mean), ptr(std), 0.0, ptr(y32), rows, D, cur_stream()), "normalize")
    check(lib.rmb_denormalize(ptr(dx), 0, ptr(std), ptr(mean), 0.0, ptr(y64), rows, D, cur_stream()), "denormalize")
    torch.cuda.synchronize()
    assert np.array_equal(y32.cpu().numpy(), normalize_data(x, stats).astype(np.float32))
    assert np.array_equal(y64.cpu().numpy(), denormalize_data(x, stats))
    lim = dict(stats, norm_config={"type": "limits", "out_min": -1.0, "out_max": 1.0})
    scale_n = torch.tensor(2.0 / stats["range"], device=cuda)
    scale_d = torch.tensor(stats["range"] / 2.0, device=cuda)
    mn = torch.tensor(stats["min"], device=cuda)
    check(lib.rmb_normalize(ptr(dx), 1, ptr(mn), ptr(scale_n), -1.0, ptr(y32), rows, D, cur_stream()), "normalize")
    check(lib.rmb_denormalize(ptr(dx), 1, ptr(scale_d), ptr(mn), -1.0, ptr(y64), rows, D, cur_stream()), "denorm")
    torch.cuda.synchronize()
    assert np.array_equal(y32.cpu().numpy(), normalize_data(x, lim).astype(np.float32))
    assert np.array_equal(y64.cpu().numpy(), denormalize_data(x, lim))
