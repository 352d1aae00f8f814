"""TEST INFRASTRUCTURE ONLY. Runs the REFERENCE's own ACT rollout step on the CPU: the unmodified ACTPolicy / DETRVAE
(third_party/act/policy.py:9-38, third_party/act/detr/models/detr_vae.py:234-259) and the unmodified
RolloutAct.infer_policy / RolloutBase.get_state / get_images (robo_manip_baselines/policy/act/RolloutAct.py:68-101,
common/base/RolloutBase.py:463-490), imported from oracle/_ref (see oracle/make_ref.py) or from /root/reference, with
the import shims of tests/golden/ref_loader.py (stub modules for packages the inference path never executes).
Used by bench.py's CPU legs only."""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def reference_root():
    for cand in (os.environ.get("RMB_REFERENCE_ROOT"), os.path.join(HERE, "_ref"), "/root/reference"):
        if cand and os.path.isfile(os.path.join(cand, "third_party", "act", "policy.py")) and \
                os.path.isfile(os.path.join(cand, "robo_manip_baselines", "policy", "act", "RolloutAct.py")):
            return cand
    return None


class ReferenceActRollout:
    """B single-episode rollouts of the reference, stepped one after the other as the reference would run them."""

    def __init__(self, policy_args, state_dim, action_dim, state_dict, model_meta_info, camera_names, batch):
        root = reference_root()
        if root is None:
            raise RuntimeError("reference sources not found (oracle/_ref is made by oracle/make_ref.py in the build container)")
        os.environ["RMB_REFERENCE_ROOT"] = root
        sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
        import ref_loader

        ref_loader.REF_ROOT = root
        ref_loader.install()
        from robo_manip_baselines.common.base.RolloutBase import RolloutBase
        from robo_manip_baselines.policy.act.RolloutAct import RolloutAct
        from torchvision.transforms import v2

        self.root = root
        self.RolloutAct, self.RolloutBase = RolloutAct, RolloutBase
        pol = ref_loader.ReferenceACTPolicy(policy_args, state_dim, action_dim)
        pol.load_state_dict(state_dict)
        self.policy = pol
        self.episodes = []
        for _ in range(batch):
            me = types.SimpleNamespace(
                args=types.SimpleNamespace(no_temp_ensem=False), state_keys=["state"], obs=None,
                device=torch.device("cpu"), camera_names=list(camera_names), model_meta_info=model_meta_info,
                action_dim=action_dim, image_transforms=v2.ToDtype(torch.float32, scale=True), policy=pol,
                policy_action_buf=[], policy_action_buf_history=[], policy_action_list=np.empty((0, action_dim)))
            me.motion_manager = types.SimpleNamespace(get_data=lambda key, obs, me=me: me.raw_state)
            me.get_state = lambda me=me: RolloutBase.get_state(me)
            me.get_images = lambda me=me: RolloutBase.get_images(me)
            self.episodes.append(me)

    def step(self, images_u8, raw_state):
        """images_u8 (B, Ncam, H, W, 3) uint8, raw_state (B, Ds) float64 -> (B, Da) float64 actions."""
        out = []
        for b, me in enumerate(self.episodes):
            me.raw_state = raw_state[b]
            me.info = {"rgb_images": {c: images_u8[b, i] for i, c in enumerate(me.camera_names)}}
            self.RolloutAct.infer_policy(me)
            out.append(me.policy_action)
        return np.stack(out)
