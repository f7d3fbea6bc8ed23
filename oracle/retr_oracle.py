"""CPU oracle for the RE⫶TR hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A plain-PyTorch fp32 restatement of the reference's arithmetic, written as pure functions of
a ``state_dict`` (the reference's own key names).  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import it; nothing under
``retr_b200/`` does.  Every function cites the reference lines it restates.

Pinned: ``tests/golden/*.pt`` were produced by ``oracle/gen_golden.py`` importing the
UNMODIFIED reference from /root/reference (``models.backbone.is_main_process`` patched to
skip the ImageNet download) in the build container; ``tests/test_oracle_golden.py`` checks
this file against them.  The reference repo itself has no tests or golden vectors for this
path, so those fixtures are the only pin there is.

The reference's arithmetic lives in torch/torchvision library modules (nn.MultiheadAttention,
nn.LayerNorm, nn.Linear, nn.Conv2d, torchvision ResNet; torch 2.11.0+cu128, torchvision
0.26.0+cu128, un-pinned by the reference).  The restatement uses torch.nn.functional primitives
(conv2d, linear, layer_norm, softmax, max_pool2d) and spells the module logic out.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]

RESNET_BLOCKS = {"ResNet18": (2, 2, 2, 2), "ResNet34": (3, 4, 6, 3), "ResNet50": (3, 4, 6, 3), "ResNet101": (3, 4, 23, 3), "ResNet152": (3, 8, 36, 3)}


# --------------------------------------------------------------------------------------------
# backbone
# --------------------------------------------------------------------------------------------
def frozen_bn(x, sd: SD, p: str):
    """FrozenBatchNorm2d.forward — models/backbone.py:41-51 (eps 1e-5, x*scale+bias)."""
    scale = sd[p + ".weight"] * (sd[p + ".running_var"] + 1e-5).rsqrt()
    bias = sd[p + ".bias"] - sd[p + ".running_mean"] * scale
    return x * scale.reshape(1, -1, 1, 1) + bias.reshape(1, -1, 1, 1)


def bottleneck(x, sd: SD, p: str, stride: int, dilation: int):
    """torchvision Bottleneck.forward — torchvision/models/resnet.py:143-165; stride and
    dilation sit on the 3x3 conv."""
    out = F.relu(frozen_bn(F.conv2d(x, sd[p + ".conv1.weight"]), sd, p + ".bn1"))
    out = F.conv2d(out, sd[p + ".conv2.weight"], stride=stride, padding=dilation, dilation=dilation)
    out = F.relu(frozen_bn(out, sd, p + ".bn2"))
    out = frozen_bn(F.conv2d(out, sd[p + ".conv3.weight"]), sd, p + ".bn3")
    if p + ".downsample.0.weight" in sd:
        x = frozen_bn(F.conv2d(x, sd[p + ".downsample.0.weight"], stride=stride), sd, p + ".downsample.1")
    return F.relu(out + x)


def basic_block(x, sd: SD, p: str, stride: int):
    """torchvision BasicBlock.forward — torchvision/models/resnet.py:88-105 (ResNet-18/34; stride on the first 3x3)."""
    out = F.relu(frozen_bn(F.conv2d(x, sd[p + ".conv1.weight"], stride=stride, padding=1), sd, p + ".bn1"))
    out = frozen_bn(F.conv2d(out, sd[p + ".conv2.weight"], padding=1), sd, p + ".bn2")
    if p + ".downsample.0.weight" in sd:
        x = frozen_bn(F.conv2d(x, sd[p + ".downsample.0.weight"], stride=stride), sd, p + ".downsample.1")
    return F.relu(out + x)


def resnet_layer4(img, sd: SD, arch: str = "ResNet101", dilation: bool = True, p: str = "backbone.body"):
    """conv1/bn1/relu/maxpool + layer1..4 (torchvision/models/resnet.py:225-264 up to layer4, as kept
    by IntermediateLayerGetter, models/backbone.py:61-65); layer4's stride becomes dilation 2 with the
    first block at dilation 1 (replace_stride_with_dilation, models/backbone.py:89-91)."""
    x = F.conv2d(img, sd[p + ".conv1.weight"], stride=2, padding=3)
    x = F.relu(frozen_bn(x, sd, p + ".bn1"))
    x = F.max_pool2d(x, kernel_size=3, stride=2, padding=1)
    dil = 1
    for li, nblocks in enumerate(RESNET_BLOCKS[arch]):
        stride = 1 if li == 0 else 2
        first_dil = dil
        if li == 3 and dilation:
            dil *= stride
            stride = 1
        for bi in range(nblocks):
            bp = f"{p}.layer{li + 1}.{bi}"
            if bp + ".conv3.weight" in sd:
                x = bottleneck(x, sd, bp, stride if bi == 0 else 1, first_dil if bi == 0 else dil)
            else:
                x = basic_block(x, sd, bp, stride if bi == 0 else 1)
    return x


def backbone(img, mask, sd: SD, cfg):
    """BackboneBase.forward — models/backbone.py:68-77: features + nearest-interpolated mask."""
    feat = resnet_layer4(img, sd, cfg.backbone, cfg.dilation)
    m = F.interpolate(mask[None].float(), size=feat.shape[-2:]).to(torch.bool)[0]
    return feat, m


# --------------------------------------------------------------------------------------------
# transformer pieces
# --------------------------------------------------------------------------------------------
def sine_table(sd: SD, length: int):
    """PositionalEncoding buffer rows 0..length-1 — models/position_encoding.py:13-35. [length, d]"""
    k = "transformer.positional_encoding."
    if k + "pe" in sd:
        return sd[k + "pe"][:length, 0]
    # learned variant ('v3'/'learned'), models/position_encoding.py:38-63 in eval mode (dropout = identity)
    return F.layer_norm(sd[k + "pos_embed.weight"][:length], (sd[k + "pos_embed.weight"].shape[1],),
                        sd[k + "LayerNorm.weight"], sd[k + "LayerNorm.bias"], 1e-5)


def sine_table_fresh(length: int, d: int):
    """The same table recomputed from its definition (models/position_encoding.py:16-21)."""
    pos = torch.arange(length).unsqueeze(1)
    div = torch.exp(torch.arange(0, d, 2) * (-math.log(10000.0) / d))
    pe = torch.zeros(length, d)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe


def mha(q_in, k_in, v_in, sd: SD, p: str, nheads: int, key_padding_mask=None, causal=False,
        drop_p: float = 0.0):
    """nn.MultiheadAttention forward as the reference calls it (need_weights=True, separate q/k/v
    inputs) — torch/nn/functional.py:6244-6664.  Batch-major: q_in [B,Lq,d], k_in/v_in [B,Lk,d];
    key_padding_mask [B,Lk] True = ignore.  Returns (out [B,Lq,d], head-averaged P [B,Lq,Lk])."""
    w, b = sd[p + ".in_proj_weight"], sd[p + ".in_proj_bias"]
    d = w.shape[1]
    dh = d // nheads
    q = F.linear(q_in, w[:d], b[:d])
    k = F.linear(k_in, w[d:2 * d], b[d:2 * d])
    v = F.linear(v_in, w[2 * d:], b[2 * d:])
    B, Lq, _ = q.shape
    Lk = k.shape[1]
    q = q.view(B, Lq, nheads, dh).transpose(1, 2) * (1.0 / math.sqrt(dh))   # q scaled before QK^T
    k = k.view(B, Lk, nheads, dh).transpose(1, 2)
    v = v.view(B, Lk, nheads, dh).transpose(1, 2)
    s = q @ k.transpose(-1, -2)                                              # [B,H,Lq,Lk]
    if causal:
        s = s + torch.full((Lq, Lk), float("-inf"), dtype=s.dtype, device=s.device).triu(1)
    if key_padding_mask is not None:
        s = s.masked_fill(key_padding_mask[:, None, None, :], float("-inf"))
    pr = torch.softmax(s, dim=-1)
    pr_used = F.dropout(pr, drop_p) if drop_p > 0 else pr
    o = (pr_used @ v).transpose(1, 2).reshape(B, Lq, d)
    o = F.linear(o, sd[p + ".out_proj.weight"], sd[p + ".out_proj.bias"])
    return o, pr.mean(dim=1)


def layer_norm(x, sd: SD, p: str, eps: float = 1e-5):
    return F.layer_norm(x, (x.shape[-1],), sd[p + ".weight"], sd[p + ".bias"], eps)


def ffn(x, sd: SD, p: str):
    """feed_forward — models/transformer_modules.py:6-11."""
    h = F.relu(F.linear(x, sd[p + ".0.weight"], sd[p + ".0.bias"]))
    return F.linear(h, sd[p + ".2.weight"], sd[p + ".2.bias"])


def _drop(x, p):
    return F.dropout(x, p) if p > 0 else x


def encoder(src, mask, pos, sd: SD, cfg, drop_p=0.0):
    """TransformerEncoder/-Layer — models/ConcatTransformer.py:85-108,171-194 with SelfAttResidual
    (transformer_modules.py:22-46) and FFResidual (:77-97).  src [B,S,d], mask [B,S], pos [S,d]."""
    x, atts = src, []
    for l in range(cfg.enc_layers):
        p = f"transformer.encoder.layers.{l}"
        n = layer_norm(x, sd, p + ".self_attn.norm")
        a, w = mha(n + pos, n + pos, n, sd, p + ".self_attn.sublayer", cfg.nheads, mask, False, drop_p)
        x = x + _drop(a, drop_p)
        x = x + _drop(ffn(layer_norm(x, sd, p + ".ff.norm"), sd, p + ".ff.sublayer"), drop_p)
        atts.append(w)
    if cfg.pre_norm:                                  # encoder_norm exists only then (:24)
        x = layer_norm(x, sd, "transformer.encoder.norm")
    return x, torch.stack(atts)


def decoder_embeddings(ids, sd: SD, cfg, drop_p=0.0):
    """DecoderEmbeddings — models/transformer_modules.py:113-129."""
    p = "transformer.embeddings"
    T = ids.shape[1]
    # padding_idx only zeroes the gradient of that row (the row itself is xavier-initialised, not 0)
    e = F.embedding(ids, sd[p + ".word_embeddings.weight"], padding_idx=cfg.pad_token_id)
    e = e + sd[p + ".position_embeddings.weight"][:T]
    return _drop(layer_norm(e, sd, p + ".LayerNorm", cfg.layer_norm_eps), drop_p)


def decoder(tgt, memory, mem_mask, tgt_mask, pos, sd: SD, cfg, drop_p=0.0):
    """TransformerDecoder/-Layer — models/ConcatTransformer.py:119-149,220-257.  query_pos is the
    whole learned position table (:64-65), so T must equal max_position_embeddings."""
    qpos = sd["transformer.embeddings.position_embeddings.weight"]          # [T,d]
    y, self_atts, cross_atts = tgt, [], []
    for l in range(cfg.dec_layers):
        p = f"transformer.decoder.layers.{l}"
        n = layer_norm(y, sd, p + ".tgt_self_attn.norm")
        a, w = mha(n + qpos, n + qpos, n, sd, p + ".tgt_self_attn.sublayer", cfg.nheads, tgt_mask, True, drop_p)
        y = y + _drop(a, drop_p)
        self_atts.append(w)
        n = layer_norm(y, sd, p + ".tgt_src_cross_attn.norm")
        a, w = mha(n + qpos, memory + pos, memory, sd, p + ".tgt_src_cross_attn.sublayer", cfg.nheads,
                   mem_mask, False, drop_p)
        y = y + _drop(a, drop_p)
        cross_atts.append(w)
        y = y + _drop(ffn(layer_norm(y, sd, p + ".ff.norm"), sd, p + ".ff.sublayer"), drop_p)
    y = layer_norm(y, sd, "transformer.decoder.norm")
    return y, torch.stack(self_atts), torch.stack(cross_atts)


def mlp_head(x, sd: SD):
    """MLP — models/caption.py:161-174."""
    h = F.relu(F.linear(x, sd["mlp.layers.0.weight"], sd["mlp.layers.0.bias"]))
    h = F.relu(F.linear(h, sd["mlp.layers.1.weight"], sd["mlp.layers.1.bias"]))
    return F.linear(h, sd["mlp.layers.2.weight"], sd["mlp.layers.2.bias"])


# --------------------------------------------------------------------------------------------
# model forward
# --------------------------------------------------------------------------------------------
def image_tokens(img, mask, sd: SD, cfg):
    """backbone -> input_proj -> flatten (models/caption.py:117-123). -> [B,196,d], [B,196]"""
    feat, m = backbone(img, mask, sd, cfg)
    src = F.conv2d(feat, sd["input_proj.weight"], sd["input_proj.bias"])
    return src.flatten(2).transpose(1, 2), m.flatten(1)


def encoder_inputs(sd: SD, cfg, t_img, t_mask, g_img=None, g_mask=None, loc=None):
    """Token/mask assembly of the three model classes (models/caption.py:23-47, 64-95, 112-158;
    concat at models/ConcatTransformer.py:48-51).  Returns src [B,S,d], mask [B,S]."""
    src, mask = image_tokens(t_img, t_mask, sd, cfg)
    B = src.shape[0]
    if loc is not None:
        w, b = sd["loc_proj.weight"], sd["loc_proj.bias"]
        if w.shape[1] == 1:                       # CaptionGlobalLoc: one token per scalar
            loc_tok = F.linear(loc.unsqueeze(2), w, b)
        else:                                     # CaptionLoc: whole vector -> one token
            loc_tok = F.linear(loc, w, b).unsqueeze(1)
        src = torch.cat([src, loc_tok], 1)
        mask = torch.cat([mask, torch.zeros(B, loc_tok.shape[1], dtype=torch.bool, device=mask.device)], 1)
    if g_img is not None:
        g_src, g_m = image_tokens(g_img, g_mask, sd, cfg)
        assert bool((~g_m).any(1).all()), "all-masked context row: ensure_unmasked_values path (models/utils.py:60-89)"
        src = torch.cat([src, g_src], 1)
        mask = torch.cat([mask, g_m], 1)
    return src, mask


def encode(sd: SD, cfg, t_img, t_mask, g_img=None, g_mask=None, loc=None, drop_p=0.0):
    src, mask = encoder_inputs(sd, cfg, t_img, t_mask, g_img, g_mask, loc)
    pos = sine_table(sd, src.shape[1])
    memory, enc_att = encoder(src, mask, pos, sd, cfg, drop_p)
    return memory, mask, pos, enc_att


def decode_logits(sd: SD, cfg, memory, mask, pos, ids, id_mask, drop_p=0.0):
    tgt = decoder_embeddings(ids, sd, cfg, drop_p)
    hs, self_att, cross_att = decoder(tgt, memory, mask, id_mask, pos, sd, cfg, drop_p)
    return mlp_head(hs, sd), self_att, cross_att


def forward(sd: SD, cfg, t_img, t_mask, g_img, g_mask, loc, ids, id_mask, return_attention=False,
            drop_p=0.0):
    """Caption*.forward (models/caption.py:112-158): logits [B,T,V] (+ the three stacked,
    head-averaged attention maps, models/ConcatTransformer.py:72-74)."""
    memory, mask, pos, enc_att = encode(sd, cfg, t_img, t_mask, g_img, g_mask, loc, drop_p)
    logits, self_att, cross_att = decode_logits(sd, cfg, memory, mask, pos, ids, id_mask, drop_p)
    if return_attention:
        return logits, {"enc_tc_self_att": enc_att, "dec_exp_self_att": self_att,
                        "dec_exp_tc_cross_att": cross_att}
    return logits


# --------------------------------------------------------------------------------------------
# greedy decode
# --------------------------------------------------------------------------------------------
def create_caption_and_mask(start_token, max_length, batch_size=1):
    """eval_utils/decode.py:20-27."""
    cap = torch.zeros((batch_size, max_length), dtype=torch.long)
    msk = torch.ones((batch_size, max_length), dtype=torch.bool)
    cap[:, 0] = start_token
    msk[:, 0] = False
    return cap, msk


def greedy(sd: SD, cfg, t_img, t_mask, g_img, g_mask, loc, max_len=128, bos_token=101, eos_token=102,
           literal=True, max_steps: Optional[int] = None):
    """eval_utils/decode.py:53-81.  ``literal=True`` re-runs the FULL model (both ResNet passes)
    every step exactly like the reference; ``literal=False`` encodes once and re-runs only the
    decoder over the whole 128-long buffer (same numbers: position i depends on ids 0..i only).
    ``max_steps`` bounds the loop for timing samples.  Returns (caption [B,max_len], steps run)."""
    B = t_img.shape[0]
    cap, msk = create_caption_and_mask(bos_token, max_len, B)
    cap, msk = cap.to(t_img.device), msk.to(t_img.device)      # decode.py:62 moves the samples; ids live beside them
    finished = torch.zeros(B, dtype=torch.bool, device=t_img.device)
    enc = None if literal else encode(sd, cfg, t_img, t_mask, g_img, g_mask, loc)
    steps = 0
    for i in range(max_len - 1):
        if max_steps is not None and steps >= max_steps:
            break
        if literal:
            logits = forward(sd, cfg, t_img, t_mask, g_img, g_mask, loc, cap, msk)
        else:
            logits = decode_logits(sd, cfg, enc[0], enc[1], enc[2], cap, msk)[0]
        steps += 1
        nxt = logits[:, i, :].argmax(-1)
        finished |= nxt == eos_token
        if bool(finished.all()):
            return cap, steps
        cap[:, i + 1] = nxt
        msk[:, i + 1] = False
    return cap, steps


def prune_cap_ids(idx_seqs, clean=True, pad_token=0, bos_token=1, eos_token=2):
    """eval_utils/decode.py:84-101."""
    out = []
    for seq in idx_seqs:
        cut = []
        for t in seq:
            cut.append(t)
            if t == eos_token:
                break
        if clean:
            cut = [t for t in cut if t not in (pad_token, bos_token, eos_token)]
        out.append(cut)
    return out


# --------------------------------------------------------------------------------------------
# training step
# --------------------------------------------------------------------------------------------
def train_loss(sd: SD, cfg, t_img, t_mask, g_img, g_mask, loc, caps, cap_masks, drop_p=0.0):
    """engine.py:70-71: teacher-forced logits on caps[:, :-1]; CE vs caps[:, 1:], mean over ALL
    B*(T) positions, pads included (models/caption.py:210, no ignore_index)."""
    logits = forward(sd, cfg, t_img, t_mask, g_img, g_mask, loc, caps[:, :-1], cap_masks[:, :-1],
                     drop_p=drop_p)
    return F.cross_entropy(logits.permute(0, 2, 1), caps[:, 1:])
