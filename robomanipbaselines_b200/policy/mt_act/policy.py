"""Drop-in for the MT-ACT `ACTPolicy` (third_party/roboagent, called as
`self.policy(state, images, task_emb=task_emb)`, policy/mt_act/RolloutMtAct.py:84-85; constructed from the same arg
dict as ACT, policy/mt_act/TrainMtAct.py:62-79, default chunk size 20, :36-38).

ARCHITECTURE HYPOTHESIS -- PARITY UNPINNED. third_party/roboagent (github.com/mmurooka/roboagent @ b60fca9) is an
un-vendored submodule and the MiniLM text encoder's weights are absent; what the mount pins is only the call
signature, the (1, 384) embedding shape (RolloutMtAct.py:115-120) and, through draw_plot's
`correlation_mat[2:, 1].reshape(15, 20 * Ncam)` (:147-155), that the encoder sequence has 2 prefix tokens. Built here:
  * proj_text_emb = Linear(384, hidden) on the sentence embedding;
  * FiLM in the shared ResNet-18: layers 2..4, every BasicBlock: out = (1 + gamma) * bn2(conv2(.)) + beta before the
    identity add, [gamma | beta] from backbones.0.0.film_models.{1,2,3} = Linear(hidden, n_blocks * 2 * planes);
  * optionally (args["task_token"] = True) the projected embedding as a third prefix token (additional_pos_embed 3 rows).
The state-dict key names of these extra tensors are part of the hypothesis.
"""
import torch
from torch import nn

from ..act.policy import ACTPolicy as _ActPolicy
from ..act.policy import DETRVAE  # noqa: F401  (RolloutMtAct.setup_policy sets the class-level dims, RolloutMtAct.py:34-36)

TASK_EMB_DIM = 384      # sentence-transformers/all-MiniLM-L6-v2 (RolloutMtAct.py:54-56)


class ACTPolicy(_ActPolicy):
    def __init__(self, args_override, state_dim=None, action_dim=None, max_batch=1):
        super().__init__(args_override, state_dim, action_dim, max_batch)
        d = self.args["hidden_dim"]
        self.task_token = bool(self.args.get("task_token", False))
        self.model.proj_text_emb = nn.Linear(TASK_EMB_DIM, d)
        self.model.additional_pos_embed = nn.Embedding(3 if self.task_token else 2, d)
        planes = [64, 128, 256, 512]
        self.model.backbones[0][0].film_models = nn.ModuleList(
            [None] + [nn.Linear(d, 2 * 2 * planes[i]) for i in (1, 2, 3)])
        self._task_key = None

    def engine_kwargs(self):
        return dict(multi_task=True, task_emb_dim=TASK_EMB_DIM, task_token=self.task_token)

    def _invalidate(self):
        super()._invalidate()
        self._task_key = None

    def set_task_embedding(self, task_emb, batch, img_h, img_w, device):
        """(B or 1, 384) float32 -> engine-resident token + FiLM parameters (skipped when unchanged)."""
        eng = self.engine_for(batch, img_h, img_w, device)
        # "unchanged" = the very same tensor object at the same version (a reference is kept, so its storage cannot be
        # recycled for another embedding behind our back), for the same batch on the same engine
        key = (task_emb, task_emb._version, batch, eng)
        old = self._task_key
        if old is None or old[0] is not key[0] or old[1:3] != key[1:3] or old[3] is not eng:
            te = task_emb.to(device=device, dtype=torch.float32)
            if te.dim() != 2 or te.shape[1] != TASK_EMB_DIM or te.shape[0] not in (1, batch):
                raise ValueError(f"task_emb must be (B or 1, {TASK_EMB_DIM}), got {tuple(task_emb.shape)}")
            eng.set_task_embedding(te.expand(batch, -1).contiguous())
            self._task_key = key
        return eng

    def __call__(self, qpos, image, actions=None, is_pad=None, task_emb=None):
        if actions is not None:
            raise NotImplementedError("training branch is out of scope of the rollout engine")
        if task_emb is None:
            raise ValueError("MT-ACT needs task_emb (policy/mt_act/RolloutMtAct.py:84-85)")
        if not image.is_cuda:
            raise RuntimeError("ACTPolicy runs on the B200 engine only: move inputs to cuda (no CPU fallback)")
        if image.dtype == torch.uint8:
            img_h, img_w = image.shape[2], image.shape[3]
        else:
            img_h, img_w = image.shape[3], image.shape[4]
        eng = self.set_task_embedding(task_emb, qpos.shape[0], img_h, img_w, image.device)
        return eng.forward(image.contiguous(), qpos.to(torch.float32).contiguous())

    forward = __call__
