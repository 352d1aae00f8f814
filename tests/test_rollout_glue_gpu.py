"""GPU parity of the rollout glue (observation ring buffers, action FIFO, cv2-exact image front end, no_temp_ensem FIFO)
against golden vectors produced by EXECUTING THE REFERENCE's RolloutDiffusionPolicy(3d).infer_policy / get_state /
get_images / get_pointcloud (tests/golden/make_dp_rollout_golden.py) and against the pinned oracles. Everything here is
bit-exact: buffers, positions and float64 actions."""
import os

import numpy as np
import pytest
import torch

from oracle import rollout_oracle as ror

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _stats(g, prefix, norm, example_dim):
    s = {k: g[f"{prefix}_{k}"] for k in ("mean", "std", "min", "range")}
    s["norm_config"] = {"type": "limits", "out_min": -1.0, "out_max": 1.0} if norm == "limits" else {"type": "gaussian"}
    s["example"] = np.zeros(example_dim)
    return s


class _RecordingPolicy:
    """Stands in for predict_action exactly as the golden run's fake did: returns the k-th stored action block for
    every episode of the sub-batch and records what it was fed."""

    def __init__(self, fake_actions, horizon_pad=2):
        self.fake, self.calls, self.pad = fake_actions, [], horizon_pad

    def predict_action(self, obs, x_init=None, step_noise=None):
        k = len(self.calls)
        self.calls.append({key: v.clone() for key, v in obs.items()})
        n = obs["state"].shape[0]
        a = torch.from_numpy(self.fake[k]).to(obs["state"].device)
        # (n, T, Da) trajectory of which [pad : pad + Na] is the action block (a strided view, as the real policy returns)
        traj = torch.zeros(n, a.shape[0] + 2 * self.pad, a.shape[1], device=a.device)
        traj[:, self.pad:self.pad + a.shape[0]] = a
        return {"action": traj[:, self.pad:self.pad + a.shape[0]], "action_pred": traj}


@pytest.mark.parametrize("norm", ["gaussian", "limits"])
def test_batched_dp_rollout_matches_reference_golden(cuda, norm):
    """uint8 env frames -> DpImageFrontend (cv2-exact resize, *2-1) -> ring buffers -> FIFO -> denormalise, for two
    episodes that replay the golden episode; policy inputs at every inference and all actions bit-for-bit."""
    from robomanipbaselines_b200.common.batched_diffusion_rollout import BatchedDiffusionRollout, DpImageFrontend

    g = np.load(os.path.join(GOLD, "dp_rollout_golden.npz"))
    p = f"dp_{norm}"
    raw, imgs, fake = g[f"{p}_raw_states"], g[f"{p}_images_u8"], g[f"{p}_fake_actions"]
    To, Ds = g[f"{p}_in_state"].shape[1:]
    Na, Da = fake.shape[1:]
    h, w = g[f"{p}_in_front"].shape[-2:]
    meta = {"data": {"n_obs_steps": To, "n_action_steps": Na, "image_size": (w, h)},
            "state": _stats(g, f"{p}_state", norm, Ds), "action": _stats(g, f"{p}_action", norm, Da)}
    pol = _RecordingPolicy(fake)
    B = 2
    br = BatchedDiffusionRollout(pol, B, meta, device=str(cuda), obs_keys=("front_rgb_image", "side_rgb_image"),
                                 frontend=DpImageFrontend(["front", "side"], (w, h)))
    br.reset()
    for t in range(raw.shape[0]):
        rs = torch.from_numpy(np.stack([raw[t]] * B)).to(cuda)
        fr = torch.from_numpy(np.stack([imgs[t]] * B)).to(cuda)
        act = br.infer_env(rs, rgb_u8=fr).cpu().numpy()
        for b in range(B):
            assert np.array_equal(act[b], g[f"{p}_actions"][t]), (t, b)
    assert [i for i in range(raw.shape[0]) if i % Na == 0] == list(g[f"{p}_infer_steps"])
    assert len(pol.calls) == len(g[f"{p}_infer_steps"])
    for k, c in enumerate(pol.calls):
        for b in range(B):
            assert np.array_equal(c["state"][b].cpu().numpy(), g[f"{p}_in_state"][k])
            assert np.array_equal(c["front_rgb_image"][b].cpu().numpy(), g[f"{p}_in_front"][k])
            assert np.array_equal(c["side_rgb_image"][b].cpu().numpy(), g[f"{p}_in_side"][k])


@pytest.mark.parametrize("norm", ["gaussian", "limits"])
def test_batched_dp3_rollout_matches_reference_golden(cuda, norm):
    """Point-cloud ring buffer + state buffer + FIFO vs the reference's RolloutDiffusionPolicy3d run. The clouds fed
    are the ones the reference produced (its FPS was a stand-in in the golden run; the device pipeline's own parity is
    tests/test_pointcloud_gpu.py); between inferences garbage is fed, which must never reach the buffers."""
    from robomanipbaselines_b200.common.batched_diffusion_rollout import BatchedDiffusionRollout

    g = np.load(os.path.join(GOLD, "dp_rollout_golden.npz"))
    p = f"dp3_{norm}"
    raw, fake, in_pc = g[f"{p}_raw_states"], g[f"{p}_fake_actions"], g[f"{p}_in_pc"]
    To, Ds = g[f"{p}_in_state"].shape[1:]
    Na, Da = fake.shape[1:]
    meta = {"data": {"n_obs_steps": To, "n_action_steps": Na},
            "state": _stats(g, f"{p}_state", norm, Ds), "action": _stats(g, f"{p}_action", norm, Da)}
    pol = _RecordingPolicy(fake)
    B = 3
    br = BatchedDiffusionRollout(pol, B, meta, device=str(cuda), obs_keys=("point_cloud",))
    br.reset()
    steps = list(g[f"{p}_infer_steps"])
    rng = np.random.default_rng(0)
    for t in range(raw.shape[0]):
        pc = in_pc[steps.index(t), -1] if t in steps else rng.normal(size=in_pc.shape[2:]).astype(np.float32)
        act = br.infer(torch.from_numpy(np.stack([raw[t]] * B)).to(cuda),
                       point_cloud=torch.from_numpy(np.stack([pc] * B)).to(cuda)).cpu().numpy()
        for b in range(B):
            assert np.array_equal(act[b], g[f"{p}_actions"][t]), (t, b)
    assert len(pol.calls) == len(steps)
    for k, c in enumerate(pol.calls):
        for b in range(B):
            assert np.array_equal(c["state"][b].cpu().numpy(), g[f"{p}_in_state"][k])
            assert np.array_equal(c["point_cloud"][b].cpu().numpy(), in_pc[k])


def test_obs_ring_push_and_gather_kernels(cuda, lib):
    """rmb_obs_ring_push modes 0/1/2 and rmb_gather_rows on odd row sizes (4-byte path) and 16-byte rows."""
    from robomanipbaselines_b200._lib import check, cur_stream, ptr

    rng = np.random.default_rng(3)
    for E in (7, 64, 1031):
        B, To = 5, 3
        buf = rng.normal(size=(B, To, E)).astype(np.float32)
        new = rng.normal(size=(B, E)).astype(np.float32)
        mode = np.array([0, 1, 2, 2, 1], dtype=np.uint8)
        want = buf.copy()
        for b in range(B):
            if mode[b] == 1:
                want[b] = new[b]
            elif mode[b] == 2:
                want[b] = np.concatenate([buf[b, 1:], new[b][None]])
        d_buf, d_new, d_mode = (torch.from_numpy(a).to(cuda) for a in (buf, new, mode))
        check(lib.rmb_obs_ring_push(ptr(d_buf), ptr(d_new), ptr(d_mode), B, To, E * 4, cur_stream()), "push")
        assert np.array_equal(d_buf.cpu().numpy(), want)
        idx = np.array([4, 0, 0, 2], dtype=np.int32)
        d_idx = torch.from_numpy(idx).to(cuda)
        d_dst = torch.zeros(4, To, E, device=cuda)
        check(lib.rmb_gather_rows(ptr(d_buf), ptr(d_idx), ptr(d_dst), 4, To * E * 4, cur_stream()), "gather")
        assert np.array_equal(d_dst.cpu().numpy(), want[idx])


def test_rollout_rejects_wrong_dtype_and_shape(cuda):
    """The C ABI takes raw pointers: float32 / non-contiguous / wrongly shaped inputs must raise, not be reinterpreted."""
    from robomanipbaselines_b200.common.batched_diffusion_rollout import BatchedDiffusionRollout
    from robomanipbaselines_b200.common.synthetic import synthetic_norm_stats

    meta = {"data": {"n_obs_steps": 2, "n_action_steps": 4}, "state": synthetic_norm_stats(5, 1),
            "action": synthetic_norm_stats(3, 2)}
    br = BatchedDiffusionRollout(_RecordingPolicy(np.zeros((4, 4, 3), np.float32)), 2, meta, device=str(cuda), obs_keys=())
    with pytest.raises(TypeError):
        br.infer(torch.zeros(2, 5, device=cuda, dtype=torch.float32))
    with pytest.raises(ValueError):
        br.infer(torch.zeros(3, 5, device=cuda, dtype=torch.float64))
    with pytest.raises(ValueError):
        br.infer(torch.zeros(2, 5, dtype=torch.float64))
    # a non-contiguous view is made contiguous, not read with the wrong strides
    raw = torch.arange(20, device=cuda, dtype=torch.float64).view(2, 10)[:, ::2]
    br.infer(raw)
    want = ror.normalize_data(raw.cpu().numpy(), meta["state"]).astype(np.float32)
    assert np.array_equal(br.obs_buf["state"][:, -1].cpu().numpy(), want)
