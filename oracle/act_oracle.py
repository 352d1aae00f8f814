"""TEST INFRASTRUCTURE ONLY (parity checker; never on the product path).

Plain-PyTorch fp32 CPU restatement of the reference's ACT inference forward, written against the checkpoint
state-dict keys (no reference code is imported):
  * ACTPolicy.__call__ inference branch      -- third_party/act/policy.py:18-38
  * DETRVAE.forward inference branch         -- third_party/act/detr/models/detr_vae.py:89-98,122-150
  * Backbone / FrozenBatchNorm2d / Joiner    -- third_party/act/detr/models/backbone.py:21-57,86-112 (torchvision resnet18)
  * PositionEmbeddingSine(normalize=True)    -- third_party/act/detr/models/position_encoding.py:30-52
  * Transformer / encoder / decoder (post-norm, return_intermediate) -- third_party/act/detr/models/transformer.py:49-250

Parity pinned: tests/golden/make_golden.py runs the reference's own DETRVAE (imported from /root/reference in the
build container) on the same synthetic weights and inputs and stores its outputs in tests/golden/act_golden_*.npz;
tests/test_oracle.py checks this restatement against them (max-abs <= 2e-5, fp32 round-off of a different op order).
"""
import math

import torch
import torch.nn.functional as F

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def frozen_bn(x, sd, prefix):
    """backbone.py:47-57 (eps = 1e-5 added before rsqrt)."""
    w = sd[prefix + ".weight"].reshape(1, -1, 1, 1)
    b = sd[prefix + ".bias"].reshape(1, -1, 1, 1)
    rv = sd[prefix + ".running_var"].reshape(1, -1, 1, 1)
    rm = sd[prefix + ".running_mean"].reshape(1, -1, 1, 1)
    scale = w * (rv + 1e-5).rsqrt()
    return x * scale + (b - rm * scale)


def resnet18_trunk(x, sd, prefix, film=None):
    """torchvision resnet18 up to layer4 (IntermediateLayerGetter return_layers={'layer4': '0'}, backbone.py:66-71).

    film (MT-ACT HYPOTHESIS, parity unpinned -- third_party/roboagent is not in the mount): {layer index 2..4:
    (B, 2, n_blocks, planes)} per-sample FiLM; block `bi` of that layer applies (1 + gamma) * bn2(conv2(.)) + beta with
    gamma = film[:, 0, bi], beta = film[:, 1, bi] before the identity branch is added."""
    x = F.conv2d(x, sd[prefix + "conv1.weight"], None, stride=2, padding=3)
    x = F.relu(frozen_bn(x, sd, prefix + "bn1"))
    x = F.max_pool2d(x, kernel_size=3, stride=2, padding=1)
    for li in range(1, 5):
        for bi in range(2):
            p = f"{prefix}layer{li}.{bi}."
            stride = 2 if (li > 1 and bi == 0) else 1
            identity = x
            out = F.conv2d(x, sd[p + "conv1.weight"], None, stride=stride, padding=1)
            out = F.relu(frozen_bn(out, sd, p + "bn1"))
            out = F.conv2d(out, sd[p + "conv2.weight"], None, stride=1, padding=1)
            out = frozen_bn(out, sd, p + "bn2")
            if film is not None and li in film:
                gamma, beta = film[li][:, 0, bi], film[li][:, 1, bi]
                out = (1 + gamma[:, :, None, None]) * out + beta[:, :, None, None]
            if (p + "downsample.0.weight") in sd:
                identity = F.conv2d(x, sd[p + "downsample.0.weight"], None, stride=stride)
                identity = frozen_bn(identity, sd, p + "downsample.1")
            x = F.relu(out + identity)
    return x


def sine_position_embedding(h, w, num_pos_feats, temperature=10000.0):
    """position_encoding.py:30-52 with normalize=True, scale=2*pi. Returns (1, 2*num_pos_feats, h, w)."""
    ones = torch.ones(1, h, w)
    y_embed = ones.cumsum(1, dtype=torch.float32)
    x_embed = ones.cumsum(2, dtype=torch.float32)
    eps, scale = 1e-6, 2 * math.pi
    y_embed = y_embed / (y_embed[:, -1:, :] + eps) * scale
    x_embed = x_embed / (x_embed[:, :, -1:] + eps) * scale
    dim_t = torch.arange(num_pos_feats, dtype=torch.float32)
    dim_t = temperature ** (2 * (dim_t // 2) / num_pos_feats)
    pos_x = x_embed[:, :, :, None] / dim_t
    pos_y = y_embed[:, :, :, None] / dim_t
    pos_x = torch.stack((pos_x[:, :, :, 0::2].sin(), pos_x[:, :, :, 1::2].cos()), dim=4).flatten(3)
    pos_y = torch.stack((pos_y[:, :, :, 0::2].sin(), pos_y[:, :, :, 1::2].cos()), dim=4).flatten(3)
    return torch.cat((pos_y, pos_x), dim=3).permute(0, 3, 1, 2)


def multihead_attention(query, key, value, sd, prefix, nheads):
    """nn.MultiheadAttention forward, eval mode, no masks; tensors are (L, B, E). Returns (out, head-averaged weights)."""
    Lq, B, E = query.shape
    Lk = key.shape[0]
    w, b = sd[prefix + ".in_proj_weight"], sd[prefix + ".in_proj_bias"]
    q = F.linear(query, w[:E], b[:E])
    k = F.linear(key, w[E:2 * E], b[E:2 * E])
    v = F.linear(value, w[2 * E:], b[2 * E:])
    hd = E // nheads
    q = q.reshape(Lq, B * nheads, hd).transpose(0, 1) * (hd ** -0.5)
    k = k.reshape(Lk, B * nheads, hd).transpose(0, 1)
    v = v.reshape(Lk, B * nheads, hd).transpose(0, 1)
    attn = torch.softmax(q @ k.transpose(1, 2), dim=-1)
    out = (attn @ v).transpose(0, 1).reshape(Lq, B, E)
    out = F.linear(out, sd[prefix + ".out_proj.weight"], sd[prefix + ".out_proj.bias"])
    return out, attn.reshape(B, nheads, Lq, Lk).mean(1)


def layer_norm(x, sd, prefix):
    return F.layer_norm(x, (x.shape[-1],), sd[prefix + ".weight"], sd[prefix + ".bias"], 1e-5)


def encoder_layer(src, pos, sd, p, nheads):
    """TransformerEncoderLayer.forward_post (transformer.py:166-179)."""
    q = k = src + pos
    src2, attn = multihead_attention(q, k, src, sd, p + "self_attn", nheads)
    src = layer_norm(src + src2, sd, p + "norm1")
    src2 = F.linear(F.relu(F.linear(src, sd[p + "linear1.weight"], sd[p + "linear1.bias"])),
                    sd[p + "linear2.weight"], sd[p + "linear2.bias"])
    return layer_norm(src + src2, sd, p + "norm2"), attn


def decoder_layer(tgt, memory, pos, query_pos, sd, p, nheads):
    """TransformerDecoderLayer.forward_post (transformer.py:229-250)."""
    q = k = tgt + query_pos
    tgt2, _ = multihead_attention(q, k, tgt, sd, p + "self_attn", nheads)
    tgt = layer_norm(tgt + tgt2, sd, p + "norm1")
    tgt2, _ = multihead_attention(tgt + query_pos, memory + pos, memory, sd, p + "multihead_attn", nheads)
    tgt = layer_norm(tgt + tgt2, sd, p + "norm2")
    tgt2 = F.linear(F.relu(F.linear(tgt, sd[p + "linear1.weight"], sd[p + "linear1.bias"])),
                    sd[p + "linear2.weight"], sd[p + "linear2.bias"])
    return layer_norm(tgt + tgt2, sd, p + "norm3")


@torch.no_grad()
def act_forward(sd, cfg, qpos, image, live_only=False, taps=None, task_emb=None, task_token=False):
    """a_hat (B, num_queries, action_dim) for qpos (B, Ds) f32 and image (B, Ncam, 3, H, W) f32 in [0, 1].

    sd: checkpoint state dict with the reference's `model.` key prefix (fp32 CPU tensors).
    cfg: dict with hidden_dim, nheads, enc_layers, dec_layers, num_queries, camera_names.
    live_only: skip decoder layers >= 1, whose outputs the reference computes and discards (detr_vae.py:142 takes [0]).
    taps: optional dict filled with intermediates ('src_tokens', 'memory', 'attn_l<i>') for debugging.
    task_emb: MT-ACT only (policy/mt_act/RolloutMtAct.py:84-85): (B or 1, 384) sentence embedding. ARCHITECTURE
      HYPOTHESIS, PARITY UNPINNED (the task-conditioned DETRVAE is in the un-vendored third_party/roboagent):
      e = proj_text_emb(task_emb); layers 2..4 of the shared ResNet get per-sample FiLM from
      backbones.0.0.film_models.{1,2,3} = Linear(512, n_blocks * 2 * planes) viewed as (B, 2, n_blocks, planes);
      task_token=True additionally makes e a third prefix token (additional_pos_embed then has 3 rows). The default
      (no token) is what RolloutMtAct.draw_plot implies: correlation_mat[2:, 1] reshaped to (15, 20 * Ncam)
      (RolloutMtAct.py:147-155) needs L = 2 + 300 * Ncam.
    """
    sd = {k[len("model."):]: v.float() for k, v in sd.items() if k.startswith("model.")}
    nheads, hidden = cfg["nheads"], cfg["hidden_dim"]
    ncam = len(cfg["camera_names"])
    bs = qpos.shape[0]
    mean = torch.tensor(IMAGENET_MEAN).view(1, 1, 3, 1, 1)
    std = torch.tensor(IMAGENET_STD).view(1, 1, 3, 1, 1)
    image = (image - mean) / std                                               # policy.py:20-22
    latent_sample = torch.zeros(bs, 32)                                        # detr_vae.py:124-125
    latent_input = F.linear(latent_sample, sd["latent_out_proj.weight"], sd["latent_out_proj.bias"])
    film, task_tok = None, None
    if task_emb is not None:
        task_tok = F.linear(task_emb.float().expand(bs, -1), sd["proj_text_emb.weight"], sd["proj_text_emb.bias"])
        film = {}
        for i, planes in ((1, 128), (2, 256), (3, 512)):
            g = F.linear(task_tok, sd[f"backbones.0.0.film_models.{i}.weight"], sd[f"backbones.0.0.film_models.{i}.bias"])
            film[i + 1] = g.view(bs, 2, 2, planes)
    feats, poss = [], []
    for cam in range(ncam):                                                    # detr_vae.py:131-136 (shared backbone)
        f = resnet18_trunk(image[:, cam], sd, "backbones.0.0.body.", film)
        poss.append(sine_position_embedding(f.shape[2], f.shape[3], hidden // 2))
        feats.append(F.conv2d(f, sd["input_proj.weight"], sd["input_proj.bias"]))
    proprio_input = F.linear(qpos, sd["input_proj_robot_state.weight"], sd["input_proj_robot_state.bias"])
    src = torch.cat(feats, dim=3)                                              # cameras folded along width
    pos = torch.cat(poss, dim=3)
    # Transformer.forward (transformer.py:49-77)
    src = src.flatten(2).permute(2, 0, 1)
    pos_embed = pos.flatten(2).permute(2, 0, 1).repeat(1, bs, 1)
    query_embed = sd["query_embed.weight"].unsqueeze(1).repeat(1, bs, 1)
    additional_pos = sd["additional_pos_embed.weight"].unsqueeze(1).repeat(1, bs, 1)
    pos_embed = torch.cat([additional_pos, pos_embed], dim=0)
    prefix = [latent_input, proprio_input] + ([task_tok] if (task_token and task_tok is not None) else [])
    src = torch.cat([torch.stack(prefix, dim=0), src], dim=0)
    if taps is not None:
        taps["src_tokens"] = src.permute(1, 0, 2).contiguous()
    tgt = torch.zeros_like(query_embed)
    memory = src
    for l in range(cfg["enc_layers"]):
        memory, attn = encoder_layer(memory, pos_embed, sd, f"transformer.encoder.layers.{l}.", nheads)
        if taps is not None:
            taps[f"attn_l{l}"] = attn
    if taps is not None:
        taps["memory"] = memory.permute(1, 0, 2).contiguous()
    out = tgt
    inter = []
    n_dec = 1 if live_only else cfg["dec_layers"]
    for l in range(n_dec):
        out = decoder_layer(out, memory, pos_embed, query_embed, sd, f"transformer.decoder.layers.{l}.", nheads)
        inter.append(layer_norm(out, sd, "transformer.decoder.norm"))          # transformer.py:129-139
    hs = torch.stack(inter).transpose(1, 2)[0]                                 # detr_vae.py:142: [0] = layer-0 output
    return F.linear(hs, sd["action_head.weight"], sd["action_head.bias"])       # detr_vae.py:148
