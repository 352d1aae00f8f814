"""RolloutAct: drop-in for robo_manip_baselines.policy.act.RolloutAct (policy/act/RolloutAct.py:15-134) on the B200
engine: same class name (bin/Rollout.py:72-95 discovers `Rollout<Policy>` in the policy package), same
setup_policy / reset_variables / infer_policy / draw_plot contract, one episode per object (B = 1 through the same
kernels that run B = 256)."""
import numpy as np
import torch

from ...common.batched_rollout import BatchedRollout
from ...common.rollout_base import RolloutBase
from .policy import ACTPolicy, DETRVAE


class RolloutAct(RolloutBase):
    policy_name = "Act"

    def setup_policy(self):
        """RolloutAct.py:23-46."""
        self.print_policy_info()
        print(f"  - chunk size: {self.model_meta_info['data']['chunk_size']}, temporal ensembling: {not self.args.no_temp_ensem}")
        DETRVAE.set_state_dim(self.state_dim)
        DETRVAE.set_action_dim(self.action_dim)
        self.policy = self.make_policy()
        # The reference registers a forward hook on every encoder layer's self_attn that stores the head-averaged
        # attention map of episode 0 as `correlation_mat` (:36-43). The engine never materialises that map; when plots
        # are enabled its opt-in debug output fills the same attribute after each inference.
        for layer in self.policy.model.transformer.encoder.layers:
            layer.self_attn.correlation_mat = None
        if self.args.checkpoint is not None:
            self.load_ckpt()
        self.rollout = None

    def make_policy(self):
        return ACTPolicy(self.model_meta_info["policy"]["args"])

    def reset_variables(self):
        """:62-66 -- the chunk buffer and its history live on the device (BatchedRollout's ring buffer / FIFO)."""
        super().reset_variables()
        if getattr(self, "rollout", None) is not None:
            self.rollout.reset()

    def _rollout_for(self, images):
        if self.rollout is None:
            self.rollout = BatchedRollout(self.policy, 1, self.model_meta_info, device=str(self.device),
                                          no_temp_ensem=self.args.no_temp_ensem,
                                          use_cuda_graph=getattr(self.args, "use_cuda_graph", None))
            self.rollout.reset()
        if not self.args.no_plot:
            eng = self.policy.engine_for(1, images.shape[2], images.shape[3], self.device)
            eng.attention_debug(True, episode=0)
            self.rollout.use_cuda_graph = False       # the debug launches are not part of the captured step
        return self.rollout

    def policy_extra_inputs(self):
        return {}

    def infer_policy(self):
        """:68-101: get_state -> get_images -> policy -> chunk history -> temporal ensembling (or FIFO pop) ->
        denormalize_data, all on the device; the denormalised float64 action comes back to the host."""
        state, images = self.get_state(), self.get_images()
        rollout = self._rollout_for(images)
        action = rollout.infer(images, state, **self.policy_extra_inputs())
        self.store_action(action[0].cpu().numpy())
        if not self.args.no_plot:
            eng = self.policy.engine_for(1, images.shape[2], images.shape[3], self.device)
            maps = eng.attention_maps().cpu().numpy()
            for layer, m in zip(self.policy.model.transformer.encoder.layers, maps):
                layer.self_attn.correlation_mat = m

    def draw_plot(self):
        """:103-134, headless: the data of every panel (camera images, action history, one attention image per encoder
        layer = correlation_mat[n_prefix:, 1] reshaped to (15, 20 * Ncam) for 480x640 cameras) goes to self.plot_data."""
        images = self.plot_images()
        h, w = next(iter(images.values())).shape[:2]
        attention_shape = (h // 32, (w // 32) * len(self.camera_names))
        n_prefix = self.policy.model.additional_pos_embed.weight.shape[0]
        attn = {}
        for layer_idx, layer in enumerate(self.policy.model.transformer.encoder.layers):
            if layer.self_attn.correlation_mat is None:
                continue
            attn[layer_idx] = layer.self_attn.correlation_mat[n_prefix:, 1].reshape(attention_shape)
        self.plot_data = {"images": images, "action": self.plot_action(), "attention": attn}
        return self.plot_data
