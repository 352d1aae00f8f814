"""GPU parity of the full ACT engine (through ACTPolicy -> C ABI) against the golden vectors the reference produced,
and of BatchedRollout against the rollout oracle. Tolerance from BASELINE.json north_star: actions max-abs <= 1e-2 in
normalised action space; ensembling indices / buffer positions bit-exact."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

from cases import ACT_CASES, act_inputs, policy_args  # noqa: E402
from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats, synthetic_state_dict  # noqa: E402
from robomanipbaselines_b200.policy.act import ACTPolicy  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(HERE, "golden")
ACTION_TOL = 1e-2   # north_star: max-abs in normalised action space


def _policy(case, max_batch=1):
    pol = ACTPolicy(policy_args(case), state_dim=case["state_dim"], action_dim=case["action_dim"], max_batch=max_batch)
    shapes = {k: tuple(v.shape) for k, v in pol.state_dict().items()}
    pol.load_state_dict(synthetic_state_dict(shapes, seed=case["weight_seed"]), strict=True)
    return pol


@pytest.mark.parametrize("name", ["small", "c1_1cam", "c1_2cam"])
def test_act_forward_matches_reference_golden(cuda, name):
    case = ACT_CASES[name]
    gold = np.load(os.path.join(GOLD, f"act_golden_{name}.npz"))
    pol = _policy(case)
    qpos, img_u8 = act_inputs(case)
    d_qpos = torch.from_numpy(qpos).to(cuda)
    d_img = torch.from_numpy(img_u8).to(cuda)
    a_u8 = pol(d_qpos, d_img)                                              # env-layout uint8 fast path
    torch.cuda.synchronize()
    eng = pol.engine_for(case["batch"], case["img_h"], case["img_w"], cuda)
    mem = eng.debug_read(1, case["batch"]).cpu().numpy()
    d_f32 = (d_img.permute(0, 1, 4, 2, 3).float() * (1.0 / 255.0)).contiguous()   # reference layout (B,Ncam,3,H,W)
    a_f32 = pol(d_qpos, d_f32)
    torch.cuda.synchronize()
    err_u8 = np.abs(a_u8.cpu().numpy() - gold["a_hat"]).max()
    err_f32 = np.abs(a_f32.cpu().numpy() - gold["a_hat"]).max()
    if "memory" in gold:
        merr = np.abs(mem - gold["memory"]).max()
    else:
        merr = np.abs(mem.mean(-1) - gold["memory_token_mean"]).max()
    print(f"[act {name}] a_hat max-abs err: u8 path {err_u8:.3e}, f32 path {err_f32:.3e} (|a|max {np.abs(gold['a_hat']).max():.2f});"
          f" encoder memory err {merr:.3e}")
    assert np.isfinite(a_u8.cpu().numpy()).all()
    assert err_u8 <= ACTION_TOL and err_f32 <= ACTION_TOL


def test_batch_invariance_and_partial_batch(cuda):
    """An episode's result must not depend on which other episodes share the batch (or on max_batch)."""
    case = dict(ACT_CASES["small"], batch=3)
    pol = _policy(case, max_batch=8)
    qpos, img = act_inputs(case)
    dq, di = torch.from_numpy(qpos).to(cuda), torch.from_numpy(img).to(cuda)
    full = pol(dq, di).clone()
    one = pol(dq[1:2].contiguous(), di[1:2].contiguous()).clone()
    rev = pol(dq.flip(0).contiguous(), di.flip(0).contiguous()).clone()
    torch.cuda.synchronize()
    assert torch.equal(full[1:2], one)
    assert torch.equal(full, rev.flip(0))


def test_full_size_batch_invariance_across_kernel_variants(cuda):
    """BASELINE configs[1] shapes (2 cameras 480x640, ff 3200, chunk 100) at 24 episodes per launch: every GEMM / conv
    then runs as CTA pairs (>= 148 tiles) with the TMA-store epilogue, while the 2-episode launch of the golden case
    runs the single-CTA kernels with other pixel boxes. Size-independent properties: an episode's chunk is bit-identical
    whatever batch it rides in, and the two golden episodes embedded in the large batch still match the reference."""
    case = ACT_CASES["c1_2cam"]
    gold = np.load(os.path.join(GOLD, "act_golden_c1_2cam.npz"))
    qpos, img = act_inputs(case)                                   # the golden inputs (batch 2)
    rng = np.random.default_rng(99)
    B = 24
    big_q = rng.standard_normal((B, case["state_dim"])).astype(np.float32)
    big_i = rng.integers(0, 256, size=(B,) + img.shape[1:], dtype=np.uint8)
    big_q[5], big_i[5] = qpos[0], img[0]
    big_q[17], big_i[17] = qpos[1], img[1]
    pol = _policy(case, max_batch=B)
    dq, di = torch.from_numpy(big_q).to(cuda), torch.from_numpy(big_i).to(cuda)
    full = pol(dq, di).clone()
    pair = pol(dq[[5, 17]].contiguous(), di[[5, 17]].contiguous()).clone()
    one = pol(dq[9:10].contiguous(), di[9:10].contiguous()).clone()
    torch.cuda.synchronize()
    assert torch.isfinite(full).all()
    assert torch.equal(full[[5, 17]], pair)
    assert torch.equal(full[9:10], one)
    err = np.abs(full[[5, 17]].cpu().numpy() - gold["a_hat"]).max()
    print(f"[act full-size] golden episodes inside a {B}-episode batch: max-abs err {err:.3e}")
    assert err <= ACTION_TOL


def test_batched_rollout_matches_oracle(cuda):
    """B episodes, 25 steps with C = 10 (ring wraps twice): ring slots / counters bit-exact; float64 actions equal the
    numpy oracle fed with the engine's own chunks; chunks within tolerance of the fp32 oracle network."""
    from oracle import act_oracle, rollout_oracle
    from robomanipbaselines_b200.common.batched_rollout import BatchedRollout

    case = dict(ACT_CASES["small"], batch=4)
    pol = _policy(case, max_batch=4)
    B, C, Da, Ds = 4, case["chunk_size"], case["action_dim"], case["state_dim"]
    meta = {"data": {"chunk_size": C}, "state": synthetic_norm_stats(Ds, 1), "action": synthetic_norm_stats(Da, 2, "limits")}
    br = BatchedRollout(pol, B, meta, device="cuda:0")
    oracles = [rollout_oracle.ActRolloutOracle(C, Da, meta["action"]) for _ in range(B)]
    sd = {k: v.cpu() for k, v in pol.state_dict().items()}
    cfg = policy_args(case)
    rng = np.random.default_rng(5)
    n_calls = [0] * B
    br.reset()
    for t in range(25):
        imgs = rng.integers(0, 256, size=(B, case["num_cams"], case["img_h"], case["img_w"], 3), dtype=np.uint8)
        raw = rng.standard_normal(size=(B, Ds))
        if t == 13:
            br.reset([False, True, False, False])
            oracles[1].reset()
            n_calls[1] = 0
        act = br.infer(torch.from_numpy(imgs).to(cuda), torch.from_numpy(raw).to(cuda))
        torch.cuda.synchronize()
        chunks = br.chunks.cpu().numpy()
        qpos_ref = np.concatenate([rollout_oracle.get_state(raw[b], meta["state"]) for b in range(B)])
        assert np.array_equal(br.qpos.cpu().numpy(), qpos_ref)                     # normalize_data bit-exact
        if t in (0, 24):                                                           # network parity (fp32 oracle is slow)
            image = torch.from_numpy(np.concatenate([rollout_oracle.get_images(list(imgs[b])) for b in range(B)]))
            ref_chunks = act_oracle.act_forward(sd, cfg, torch.from_numpy(qpos_ref), image, live_only=True).numpy()
            err = np.abs(chunks - ref_chunks).max()
            print(f"[rollout t={t}] chunk max-abs err vs fp32 oracle {err:.3e}")
            assert err <= ACTION_TOL
        for b in range(B):
            ref = oracles[b].step(chunks[b])
            n_calls[b] += 1
            assert np.array_equal(act[b].cpu().numpy(), ref), (t, b)
            oring, filled = oracles[b].ring_view(n_calls[b])
            assert np.array_equal(br.ring[b].cpu().numpy()[filled], oring[filled]), (t, b)
        assert br.step.cpu().tolist() == n_calls


def test_engine_rejects_bad_input(cuda):
    case = ACT_CASES["small"]
    pol = _policy(case)
    qpos, img = act_inputs(case)
    with pytest.raises(RuntimeError):
        pol(torch.from_numpy(qpos), torch.from_numpy(img))                       # CPU tensors: no fallback
    with pytest.raises(NotImplementedError):
        pol(torch.from_numpy(qpos).to(cuda), torch.from_numpy(img).to(cuda), actions=torch.zeros(1))
    eng = pol.engine_for(case["batch"], case["img_h"], case["img_w"], cuda)
    with pytest.raises(ValueError):
        eng.forward(torch.zeros(3, 2, 64, 96, 4, dtype=torch.uint8, device=cuda), torch.from_numpy(qpos).to(cuda))


def test_no_temp_ensem_fifo_is_per_episode(cuda):
    """--no_temp_ensem with STAGGERED resets (RolloutAct.py:70,86-87 keeps one policy_action_buf per episode): an episode
    that is reset gets a fresh inference, the others keep popping their own chunk. Actions bit-exact vs the pinned
    oracle fed with the engine's chunks; the policy runs only for the episodes whose chunk is used up."""
    from oracle import rollout_oracle
    from robomanipbaselines_b200.common.batched_rollout import BatchedRollout

    case = dict(ACT_CASES["small"], batch=3)
    pol = _policy(case, max_batch=3)
    B, C, Da, Ds = 3, case["chunk_size"], case["action_dim"], case["state_dim"]
    meta = {"data": {"chunk_size": C}, "state": synthetic_norm_stats(Ds, 1), "action": synthetic_norm_stats(Da, 2)}
    br = BatchedRollout(pol, B, meta, device="cuda:0", no_temp_ensem=True)
    oracles = [rollout_oracle.ActRolloutOracle(C, Da, meta["action"], no_temp_ensem=True) for _ in range(B)]
    rng = np.random.default_rng(8)
    br.reset()
    expected_inferences = 0
    for t in range(27):
        if t == 4:
            br.reset([False, True, False]); oracles[1].reset()
        if t == 17:
            br.reset([True, False, False]); oracles[0].reset()
        imgs = rng.integers(0, 256, size=(B, case["num_cams"], case["img_h"], case["img_w"], 3), dtype=np.uint8)
        raw = rng.standard_normal(size=(B, Ds))
        need = [o.needs_inference() for o in oracles]
        expected_inferences += int(any(need))
        d_img, d_raw = torch.from_numpy(imgs).to(cuda), torch.from_numpy(raw).to(cuda)
        act = br.infer(d_img, d_raw).cpu().numpy()
        if any(need):
            # the chunk an episode received must be what the engine computes for that episode alone (batch invariance)
            qpos = torch.from_numpy(np.concatenate([rollout_oracle.get_state(raw[b], meta["state"]) for b in range(B)])).to(cuda)
            solo = pol(qpos, d_img).cpu().numpy()
        for b in range(B):
            if need[b]:
                assert np.array_equal(br.fifo[b].cpu().numpy(), solo[b].astype(np.float64)), (t, b)
            assert np.array_equal(act[b], oracles[b].step(solo[b] if need[b] else None)), (t, b)
        assert np.array_equal(br.fifo_pos.cpu().numpy(), br._host_pos)
    assert br.n_inferences == expected_inferences


def test_bench_batch_matches_reference_golden(cuda, monkeypatch):
    """BASELINE configs[1] exactly as bench.py runs it: 256 episodes, 2 cameras 480x640. The two golden episodes sit in
    rows 0 and 255 (first tile of the first CTA, ragged last wave): chunk within tolerance of the reference's DETRVAE
    and bit-identical to the 2-episode launch (many tiles per persistent CTA, both TMEM accumulators in flight).
    Run on BOTH encoder paths: the default (residual GEMM + layernorm512_kernel) and the opt-in fused residual + LayerNorm
    epilogue (RMB_LN_FUSE=always, gemm_ln.cuh); each is bit-invariant to the batch size, the two agree to round-off."""
    results = {}
    for mode in ("never", "always"):
        monkeypatch.setenv("RMB_LN_FUSE", mode)
        results[mode] = _bench_batch_case(cuda)
    diff = float((results["never"] - results["always"]).abs().max())
    print(f"[act bench batch] fused-LayerNorm path vs default path on the golden episodes: max-abs diff {diff:.3e}")
    assert diff <= 2e-3         # both are within 1e-2 of the reference; 16-bit copies of the fp32 stream may round apart


def _bench_batch_case(cuda):
    case = ACT_CASES["c1_2cam"]
    gold = np.load(os.path.join(GOLD, "act_golden_c1_2cam.npz"))
    qpos, img = act_inputs(case)
    B = 256
    g = torch.Generator(device="cuda").manual_seed(5)
    di = torch.randint(0, 256, (B,) + img.shape[1:], device=cuda, dtype=torch.uint8, generator=g)
    dq = torch.randn(B, case["state_dim"], device=cuda, generator=g)
    di[0], di[255] = torch.from_numpy(img[0]).to(cuda), torch.from_numpy(img[1]).to(cuda)
    dq[0], dq[255] = torch.from_numpy(qpos[0]).to(cuda), torch.from_numpy(qpos[1]).to(cuda)
    pol = _policy(case, max_batch=B)
    full = pol(dq, di).clone()
    pair = pol(dq[[0, 255]].contiguous(), di[[0, 255]].contiguous()).clone()
    mid = pol(dq[100:103].contiguous(), di[100:103].contiguous()).clone()
    torch.cuda.synchronize()
    assert torch.isfinite(full).all()
    err = np.abs(full[[0, 255]].cpu().numpy() - gold["a_hat"]).max()
    print(f"[act bench batch] golden episodes at rows 0 / 255 of a 256-episode batch: max-abs err {err:.3e}")
    assert err <= ACTION_TOL
    assert torch.equal(full[[0, 255]], pair)
    assert torch.equal(full[100:103], mid)
    return pair


def test_downsample_in_conv1_launch_is_bit_identical(cuda, monkeypatch):
    """Opt-in (RMB_DOWN_FUSE=1; measured slower): the 1x1 stride-2 downsample convs ride in the launch of
    their block's conv1 as side column tiles restricted to the centre tap (gemm_tc.cuh, GemmParams::n_main_ct): same
    products in the same order as their own launches, so the chunks are bit-identical -- at a batch that takes CTA
    pairs and at one that does not."""
    case = ACT_CASES["c1_2cam"]
    qpos, img = act_inputs(case)
    out = {}
    for B in (2, 24):
        g = torch.Generator(device="cuda").manual_seed(11)
        di = torch.randint(0, 256, (B,) + img.shape[1:], device=cuda, dtype=torch.uint8, generator=g)
        dq = torch.randn(B, case["state_dim"], device=cuda, generator=g)
        di[0], dq[0] = torch.from_numpy(img[0]).to(cuda), torch.from_numpy(qpos[0]).to(cuda)
        for mode in ("0", "1"):
            monkeypatch.setenv("RMB_DOWN_FUSE", mode)
            pol = _policy(case, max_batch=B)
            out[(B, mode)] = pol(dq, di).clone()
            torch.cuda.synchronize()
            n = pol.engine_for(B, case["img_h"], case["img_w"], torch.device("cuda:0")).num_launches
            assert n == (61 if mode == "1" else 64), n
        assert torch.equal(out[(B, "0")], out[(B, "1")]), B
    gold = np.load(os.path.join(GOLD, "act_golden_c1_2cam.npz"))
    err = np.abs(out[(24, "0")][0].cpu().numpy() - gold["a_hat"][0]).max()
    print(f"[act downsample fused] bit-identical to the separate launches; golden episode max-abs err {err:.3e}")
    assert err <= ACTION_TOL


def test_cuda_graph_recaptured_when_engine_is_rebuilt(cuda):
    """A captured policy-step graph pins the engine's workspace. When another user of the same ACTPolicy makes it
    rebuild its engine (larger batch), the next infer() must recapture instead of replaying into freed memory."""
    from robomanipbaselines_b200.common.batched_rollout import BatchedRollout

    case = dict(ACT_CASES["small"], batch=2)
    pol = _policy(case, max_batch=2)
    B, C, Da, Ds = 2, case["chunk_size"], case["action_dim"], case["state_dim"]
    meta = {"data": {"chunk_size": C}, "state": synthetic_norm_stats(Ds, 1), "action": synthetic_norm_stats(Da, 2)}
    br = BatchedRollout(pol, B, meta, device="cuda:0", use_cuda_graph=True)
    ref = BatchedRollout(pol, B, meta, device="cuda:0", use_cuda_graph=False)
    rng = np.random.default_rng(2)
    br.reset(); ref.reset()
    for t in range(4):
        imgs = torch.from_numpy(rng.integers(0, 256, size=(B, case["num_cams"], case["img_h"], case["img_w"], 3), dtype=np.uint8)).to(cuda)
        raw = torch.from_numpy(rng.standard_normal(size=(B, Ds))).to(cuda)
        if t == 2:      # someone else grows the policy's engine: the old one (and its workspace) is destroyed
            big_q = torch.zeros(5, Ds, device=cuda)
            big_i = torch.zeros((5,) + tuple(imgs.shape[1:]), device=cuda, dtype=torch.uint8)
            pol(big_q, big_i)
        a = br.infer(imgs, raw).clone()
        b = ref.infer(imgs, raw).clone()
        assert torch.equal(a, b), t
