"""Drop-in for the reference's `ACTPolicy` / `DETRVAE` at inference time.

Mirrors (names, argument meaning, error behaviour, state-dict layout) of
  * ACTPolicy(args_override)(qpos, image)             -- third_party/act/policy.py:9-38
  * DETRVAE.set_state_dim / set_action_dim            -- third_party/act/detr/models/detr_vae.py:37-46
while the arithmetic runs in the sm_100a engine behind the C ABI (no PyTorch ops on the hot path).

The module tree below exists only so that `load_state_dict(torch.load(ckpt), strict=True)`, `.parameters()`,
`.buffers()` (RolloutBase.calc_model_size, common/base/RolloutBase.py:563-571) and the attention-hook registration in
RolloutAct.setup_policy (policy/act/RolloutAct.py:36-43) behave as with the reference: every one of the reference's
state-dict keys is present with the same shape, including the ones inference never reads (CVAE encoder,
is_pad_head, decoder layers 1..6).
"""
import torch
from torch import nn

from ...engine import ActEngine


class DETRVAE:
    """Carrier of the reference's class-level dims (detr_vae.py:37-46). ACTPolicy snapshots them per instance, so two
    policies with different dims can coexist in one process (the reference's class attributes collide)."""

    STATE_DIM = 7
    ACTION_DIM = 7

    @classmethod
    def set_state_dim(cls, state_dim):
        cls.STATE_DIM = state_dim

    @classmethod
    def set_action_dim(cls, action_dim):
        cls.ACTION_DIM = action_dim


class _FrozenBN(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.register_buffer("weight", torch.ones(n))
        self.register_buffer("bias", torch.zeros(n))
        self.register_buffer("running_mean", torch.zeros(n))
        self.register_buffer("running_var", torch.ones(n))

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        state_dict.pop(prefix + "num_batches_tracked", None)     # tolerated like backbone.py:36-45
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)


class _Block(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = _FrozenBN(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = _FrozenBN(cout)
        if stride != 1 or cin != cout:
            self.downsample = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), _FrozenBN(cout))


class _ResNet18Body(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = _FrozenBN(64)
        chans = [64, 64, 128, 256, 512]
        for i in range(1, 5):
            stride = 1 if i == 1 else 2
            setattr(self, f"layer{i}", nn.Sequential(_Block(chans[i - 1], chans[i], stride), _Block(chans[i], chans[i], 1)))


class _BackboneHolder(nn.Module):
    def __init__(self):
        super().__init__()
        self.body = _ResNet18Body()


class _EncLayer(nn.Module):
    def __init__(self, d, nhead, ff):
        super().__init__()
        self.self_attn = nn.MultiheadAttention(d, nhead)
        self.linear1 = nn.Linear(d, ff)
        self.linear2 = nn.Linear(ff, d)
        self.norm1 = nn.LayerNorm(d)
        self.norm2 = nn.LayerNorm(d)


class _DecLayer(nn.Module):
    def __init__(self, d, nhead, ff):
        super().__init__()
        self.self_attn = nn.MultiheadAttention(d, nhead)
        self.multihead_attn = nn.MultiheadAttention(d, nhead)
        self.linear1 = nn.Linear(d, ff)
        self.linear2 = nn.Linear(ff, d)
        self.norm1 = nn.LayerNorm(d)
        self.norm2 = nn.LayerNorm(d)
        self.norm3 = nn.LayerNorm(d)


class _Stack(nn.Module):
    def __init__(self, layers, norm=None):
        super().__init__()
        self.layers = nn.ModuleList(layers)
        if norm is not None:
            self.norm = norm


class _Transformer(nn.Module):
    def __init__(self, d, nhead, ff, n_enc, n_dec):
        super().__init__()
        self.encoder = _Stack([_EncLayer(d, nhead, ff) for _ in range(n_enc)])
        self.decoder = _Stack([_DecLayer(d, nhead, ff) for _ in range(n_dec)], nn.LayerNorm(d))
        self.d_model, self.nhead = d, nhead


class _DetrVaeParams(nn.Module):
    """Parameter container with the reference DETRVAE's state-dict keys (344 tensors at the default config)."""

    def __init__(self, args, state_dim, action_dim):
        super().__init__()
        d, nhead, ff = args["hidden_dim"], args["nheads"], args["dim_feedforward"]
        self.num_queries = args["num_queries"]
        self.camera_names = list(args["camera_names"])
        self.register_buffer("pos_table", torch.zeros(1, 2 + self.num_queries, d))
        self.transformer = _Transformer(d, nhead, ff, args["enc_layers"], args["dec_layers"])
        self.encoder = _Stack([_EncLayer(d, nhead, ff) for _ in range(args["enc_layers"])])
        self.action_head = nn.Linear(d, action_dim)
        self.is_pad_head = nn.Linear(d, 1)
        self.query_embed = nn.Embedding(self.num_queries, d)
        self.input_proj = nn.Conv2d(512, d, kernel_size=1)
        self.backbones = nn.ModuleList([nn.Sequential(_BackboneHolder())])
        self.input_proj_robot_state = nn.Linear(state_dim, d)
        self.latent_dim = 32
        self.cls_embed = nn.Embedding(1, d)
        self.encoder_action_proj = nn.Linear(action_dim, d)
        self.encoder_joint_proj = nn.Linear(state_dim, d)
        self.latent_proj = nn.Linear(d, self.latent_dim * 2)
        self.latent_out_proj = nn.Linear(self.latent_dim, d)
        self.additional_pos_embed = nn.Embedding(2, d)


_DEFAULT_ARGS = {  # third_party/act/detr/main.py:12-66 defaults that policy["args"] may omit
    "hidden_dim": 256, "dim_feedforward": 2048, "enc_layers": 4, "dec_layers": 6, "nheads": 8, "num_queries": 400,
    "backbone": "resnet18", "camera_names": [], "position_embedding": "sine", "pre_norm": False, "dilation": False,
    "masks": False,
}


class ACTPolicy(nn.Module):
    """ACTPolicy(args_override)(qpos, image) -> a_hat, as third_party/act/policy.py:9-38 (inference branch only).

    args_override is model_meta_info["policy"]["args"] (policy/act/TrainAct.py:45-57). State / action dims are taken
    from DETRVAE.STATE_DIM / ACTION_DIM at construction (as RolloutAct.setup_policy sets them, RolloutAct.py:31-33)
    unless given explicitly.
    """

    def __init__(self, args_override, state_dim=None, action_dim=None, max_batch=1):
        super().__init__()
        args = dict(_DEFAULT_ARGS)
        args.update(args_override)
        if args["backbone"] != "resnet18" or args["position_embedding"] not in ("sine", "v2"):
            raise NotImplementedError("only the resnet18 + sine position embedding configuration of RMB is supported")
        if args["pre_norm"] or args["dilation"] or args["masks"]:
            raise NotImplementedError("pre_norm / dilation / masks are not used by RoboManipBaselines")
        self.args = args
        self.state_dim = DETRVAE.STATE_DIM if state_dim is None else state_dim
        self.action_dim = DETRVAE.ACTION_DIM if action_dim is None else action_dim
        self.kl_weight = args.get("kl_weight")
        self.model = _DetrVaeParams(args, self.state_dim, self.action_dim)
        self.max_batch = max_batch
        self._engine = None
        self._engine_key = None

    # ---- engine management -------------------------------------------------------------------------------------
    def _invalidate(self):
        if self._engine is not None:
            self._engine.close()
        self._engine, self._engine_key = None, None

    def load_state_dict(self, state_dict, strict=True, **kw):
        out = super().load_state_dict(state_dict, strict=strict, **kw)
        self._invalidate()
        return out

    def engine_kwargs(self):
        """Extra ActEngine options of subclasses (MT-ACT)."""
        return {}

    def engine_for(self, batch, img_h, img_w, device):
        """(Re)build the device engine for this image size; the workspace grows to the largest batch seen."""
        dev_index = device.index if device.index is not None else torch.cuda.current_device()
        max_batch = max(self.max_batch, batch)
        key = (img_h, img_w, dev_index)
        if self._engine is None or self._engine_key != key or self._engine.max_batch < batch:
            self._invalidate()
            a = self.args
            self._engine = ActEngine(max_batch=max_batch, num_cams=len(a["camera_names"]), img_h=img_h, img_w=img_w,
                                     state_dim=self.state_dim, action_dim=self.action_dim, hidden_dim=a["hidden_dim"],
                                     dim_feedforward=a["dim_feedforward"], nheads=a["nheads"],
                                     enc_layers=a["enc_layers"], dec_layers=a["dec_layers"],
                                     chunk_size=a["num_queries"], device=dev_index, **self.engine_kwargs())
            self._engine.load_state_dict(self.state_dict())
            self._engine_key = key
            self.max_batch = max_batch
        return self._engine

    # ---- the reference call ------------------------------------------------------------------------------------
    def __call__(self, qpos, image, actions=None, is_pad=None):
        if actions is not None:
            raise NotImplementedError("training branch (policy.py:24-35) is out of scope of the rollout engine")
        if not image.is_cuda:
            raise RuntimeError("ACTPolicy runs on the B200 engine only: move inputs to cuda (no CPU fallback)")
        if image.dtype == torch.uint8:      # extension: env-layout uint8 (B, Ncam, H, W, 3), skips the f32 round trip
            img_h, img_w = image.shape[2], image.shape[3]
        else:                               # reference layout: float32 (B, Ncam, 3, H, W) in [0, 1]
            img_h, img_w = image.shape[3], image.shape[4]
        eng = self.engine_for(qpos.shape[0], img_h, img_w, image.device)
        return eng.forward(image.contiguous(), qpos.to(torch.float32).contiguous())

    forward = __call__

    def configure_optimizers(self):
        raise NotImplementedError("training is out of scope of the rollout engine")
