"""RolloutMtAct: drop-in for robo_manip_baselines.policy.mt_act.RolloutMtAct (policy/mt_act/RolloutMtAct.py:16-166) on
the B200 engine. Same as RolloutAct plus the per-text embedding cache of get_task_emb (:115-120).

The reference builds SentenceTransformer("sentence-transformers/all-MiniLM-L6-v2") in setup_policy (:54-56); its
weights are not available here (no network), so the text encoder is INJECTED: pass `text_encoder=` (any object with
`.encode([text], convert_to_tensor=True, device="cpu") -> (1, 384)`, e.g. the real SentenceTransformer) or
`task_emb_table={task_desc: (384,) array}` of precomputed embeddings. The network behind it is an architecture
hypothesis (policy/mt_act/policy.py)."""
import numpy as np
import torch

from ..act.rollout import RolloutAct
from .policy import TASK_EMB_DIM, ACTPolicy


class _TableEncoder:
    def __init__(self, table):
        self.table = {k: np.asarray(v, dtype=np.float32).reshape(1, TASK_EMB_DIM) for k, v in table.items()}

    def encode(self, texts, convert_to_tensor=True, device="cpu"):
        missing = [t for t in texts if t not in self.table]
        if missing:
            raise KeyError(f"no precomputed sentence embedding for task description(s) {missing}")
        return torch.from_numpy(np.concatenate([self.table[t] for t in texts]))


class RolloutMtAct(RolloutAct):
    require_task_desc = True
    policy_name = "MtAct"

    def setup_policy(self):
        """:28-56."""
        super().setup_policy()
        enc = getattr(self.args, "text_encoder", None)
        table = getattr(self.args, "task_emb_table", None)
        if enc is None and table is None:
            raise ValueError("RolloutMtAct needs text_encoder= or task_emb_table= (the MiniLM weights are not shipped)")
        self.text_encoder = enc if enc is not None else _TableEncoder(table)
        self.task_text_emb_map = {}
        self._task_dev = {}

    def make_policy(self):
        return ACTPolicy(self.model_meta_info["policy"]["args"])

    def reset_variables(self):
        """:73-77."""
        super().reset_variables()
        self.task_text_emb_map = {}
        self._task_dev = {}

    def get_task_emb(self, task_text):
        """:115-120."""
        if task_text not in self.task_text_emb_map:
            self.task_text_emb_map[task_text] = self.text_encoder.encode([task_text], convert_to_tensor=True, device="cpu")
        return self.task_text_emb_map[task_text]

    def _rollout_for(self, images):
        rollout = super()._rollout_for(images)
        text = self.args.task_desc
        if text not in self._task_dev:
            self._task_dev[text] = self.get_task_emb(text).to(self.device, torch.float32)           # :83
        rollout.set_task_embedding(self._task_dev[text])
        return rollout
