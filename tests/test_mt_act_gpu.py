"""MT-ACT on the engine vs oracle/act_oracle.py's task-conditioned branch. PARITY UNPINNED: the task-conditioned
DETRVAE is in the un-vendored third_party/roboagent, so both sides implement the same ARCHITECTURE HYPOTHESIS
(robomanipbaselines_b200/policy/mt_act/policy.py); what the reference pins is the call signature and shapes
(policy/mt_act/RolloutMtAct.py:79-120). Tolerance: max-abs <= 1e-2 in normalised action space."""
import os
import sys
import types

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

from cases import ACT_CASES, act_inputs, policy_args  # noqa: E402
from oracle import act_oracle, rollout_oracle  # noqa: E402
from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats, synthetic_state_dict  # noqa: E402
from robomanipbaselines_b200.policy.mt_act import ACTPolicy, RolloutMtAct  # noqa: E402

pytestmark = pytest.mark.gpu
TOL = 1e-2


def _mt_policy(case, task_token, max_batch=1, zero_film=False):
    args = dict(policy_args(case), task_token=task_token)
    pol = ACTPolicy(args, state_dim=case["state_dim"], action_dim=case["action_dim"], max_batch=max_batch)
    sd = synthetic_state_dict({k: tuple(v.shape) for k, v in pol.state_dict().items()}, seed=case["weight_seed"])
    for k in sd:
        if "film_models" in k:
            sd[k] = torch.zeros_like(sd[k]) if zero_film else sd[k] * 0.5     # gamma, beta ~ N(0, 0.5): a visible modulation
    pol.load_state_dict(sd, strict=True)
    return pol, sd, args


def _task_emb(n, seed=4):
    g = torch.Generator().manual_seed(seed)
    e = torch.randn(n, 384, generator=g)
    return e / e.norm(dim=1, keepdim=True)            # sentence-transformers embeddings are L2-normalised


@pytest.mark.parametrize("task_token", [False, True])
def test_mt_act_small_matches_oracle(cuda, task_token):
    case = ACT_CASES["small"]
    pol, sd, args = _mt_policy(case, task_token)
    qpos, img = act_inputs(case)
    B = case["batch"]
    te = _task_emb(B)
    image = torch.from_numpy(np.concatenate([rollout_oracle.get_images(list(img[b])) for b in range(B)]))
    want = act_oracle.act_forward(sd, args, torch.from_numpy(qpos), image, live_only=True, task_emb=te, task_token=task_token)
    dq, di = torch.from_numpy(qpos).to(cuda), torch.from_numpy(img).to(cuda)
    got = pol(dq, di, task_emb=te.to(cuda)).cpu()
    err = float((got - want).abs().max())
    # the conditioning is live: another task moves the actions by far more than the tolerance
    other = pol(dq, di, task_emb=_task_emb(B, seed=9).to(cuda)).cpu()
    moved = float((other - got).abs().max())
    # one (1, 384) embedding for all episodes (RolloutMtAct passes exactly that) == the same row repeated
    one = pol(dq, di, task_emb=te[:1].to(cuda)).cpu()
    rep = pol(dq, di, task_emb=te[:1].expand(B, -1).contiguous().to(cuda)).cpu()
    print(f"[mt-act small token={task_token}] max-abs err vs oracle {err:.3e}; a different task moves the chunk by {moved:.3e}")
    assert torch.isfinite(got).all() and err <= TOL
    assert moved > TOL
    assert torch.equal(one, rep)
    with pytest.raises(ValueError):
        pol(dq, di)                                     # MT-ACT without task_emb


def test_mt_act_zero_film_reproduces_the_act_reference_golden(cuda):
    """With gamma = beta = 0 and no task token the hypothesised network IS the ACT network: the reference's own DETRVAE
    output (act_golden_small.npz) pins everything except the FiLM term."""
    case = ACT_CASES["small"]
    gold = np.load(os.path.join(HERE, "golden", "act_golden_small.npz"))
    pol, sd, args = _mt_policy(case, task_token=False, zero_film=True)
    qpos, img = act_inputs(case)
    got = pol(torch.from_numpy(qpos).to(cuda), torch.from_numpy(img).to(cuda), task_emb=_task_emb(1).to(cuda)).cpu().numpy()
    err = np.abs(got - gold["a_hat"]).max()
    print(f"[mt-act zero FiLM] max-abs err vs the ACT reference golden {err:.3e}")
    assert err <= TOL


def test_mt_act_full_size_and_batch_invariance(cuda):
    """MujocoUR5eCable shapes with MT-ACT's chunk size 20 (TrainMtAct.py:36-38): 2 episodes vs the oracle, and the same
    episodes inside a 24-episode batch (CTA-pair kernels, several images per tile in layers 3-4, i.e. per-ROW FiLM
    parameters inside one tile) bit-identical."""
    case = dict(ACT_CASES["c1_2cam"], chunk_size=20)
    pol, sd, args = _mt_policy(case, task_token=False, max_batch=24)
    qpos, img = act_inputs(case)
    te = _task_emb(2)
    image = torch.from_numpy(np.concatenate([rollout_oracle.get_images(list(img[b])) for b in range(2)]))
    want = act_oracle.act_forward(sd, args, torch.from_numpy(qpos), image, live_only=True, task_emb=te)
    got = pol(torch.from_numpy(qpos).to(cuda), torch.from_numpy(img).to(cuda), task_emb=te.to(cuda)).clone()
    err = float((got.cpu() - want).abs().max())
    print(f"[mt-act full size] max-abs err vs oracle {err:.3e}")
    assert err <= TOL
    rng = np.random.default_rng(3)
    B = 24
    bq = rng.standard_normal((B, case["state_dim"])).astype(np.float32)
    bi = rng.integers(0, 256, size=(B,) + img.shape[1:], dtype=np.uint8)
    bt = _task_emb(B, seed=11)
    bq[3], bi[3], bt[3] = torch.from_numpy(qpos[0]).numpy(), img[0], te[0]
    bq[20], bi[20], bt[20] = qpos[1], img[1], te[1]
    full = pol(torch.from_numpy(bq).to(cuda), torch.from_numpy(bi).to(cuda), task_emb=bt.to(cuda)).clone()
    assert torch.equal(full[[3, 20]], got)


def test_rollout_mt_act_class(cuda):
    """RolloutMtAct with a precomputed-embedding table: get_task_emb caches per text (RolloutMtAct.py:115-120), actions
    equal BatchedRollout(B = 1) fed with the same embedding, unknown task descriptions are rejected."""
    from robomanipbaselines_b200.common.batched_rollout import BatchedRollout
    from test_rollout_classes_gpu import _Env, _MotionManager

    case = ACT_CASES["small"]
    Ds, Da, C = case["state_dim"], case["action_dim"], case["chunk_size"]
    cams = [f"cam{i}" for i in range(case["num_cams"])]
    meta = {"data": {"chunk_size": C, "skip": 1}, "state": dict(synthetic_norm_stats(Ds, 1), keys=["k0", "k1"]),
            "action": dict(synthetic_norm_stats(Da, 2), keys=["command_joint_pos"]), "image": {"camera_names": cams},
            "policy": {"args": policy_args(case)}}
    table = {"pick the cable": _task_emb(1, seed=1)[0].numpy(), "open the door": _task_emb(1, seed=2)[0].numpy()}
    mm = _MotionManager()
    r = RolloutMtAct(_Env(n_arm=Da - 1), model_meta_info=meta, motion_manager=mm, device=str(cuda),
                     task_desc="pick the cable", task_emb_table=table)
    sd = synthetic_state_dict({k: tuple(v.shape) for k, v in r.policy.state_dict().items()}, seed=case["weight_seed"])
    r.policy.load_state_dict(sd, strict=True)
    ref = BatchedRollout(r.policy, 1, meta, device=str(cuda), use_cuda_graph=False)
    ref.set_task_embedding(torch.from_numpy(table["pick the cable"][None]).to(cuda))
    ref.reset()
    rng = np.random.default_rng(6)
    for t in range(4):
        mm.raw = rng.normal(size=Ds)
        frames = {c: rng.integers(0, 256, size=(case["img_h"], case["img_w"], 3), dtype=np.uint8) for c in cams}
        r.info = {"rgb_images": frames}
        r.policy_step()
        want = ref.infer(torch.from_numpy(np.stack([frames[c] for c in cams])[None]).to(cuda),
                         torch.from_numpy(mm.raw[None]).to(cuda))[0].cpu().numpy()
        assert np.array_equal(r.policy_action, want), t
    assert list(r.task_text_emb_map) == ["pick the cable"]
    r.args.task_desc = "fold the towel"
    with pytest.raises(KeyError):
        r.policy_step()
