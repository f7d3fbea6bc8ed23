"""Margin study (TEST INFRASTRUCTURE): which stages of the hot path tolerate bf16 operands?

Runs on CPU.  The oracle's fp32 forward is compared with copies of itself in which the GEMM
OPERANDS of one stage group (backbone / encoder / cross K,V caches / decoder / head) are
rounded to bf16 (fp32 accumulation, fp32 residual stream, fp32 LayerNorm and softmax
statistics — what the CUDA engine does).  The comparison is teacher-forced on the fp32 greedy
ids: a sequence is "exact" when the argmax agrees at every one of the 127 positions, which is
exactly the condition under which a free-running greedy decode reproduces the fp32 sequence.

    python oracle/margin_study.py [n_samples]

Results of the run that shaped the precision plan are in DESIGN.md §2.
"""
import contextlib
import io
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import retr_oracle as O  # noqa: E402
from retr_b200 import synthetic  # noqa: E402
from retr_b200.configuration import Config  # noqa: E402
from retr_b200.models import caption  # noqa: E402


def r(x):
    return x.to(torch.bfloat16).to(torch.float32)


def split2(x):
    """x ~ hi + lo with both bf16 (16 mantissa bits together)."""
    hi = r(x)
    return hi + r(x - hi)


class Q:
    """rounding policy per stage group; identity unless listed"""
    def __init__(self, stages, fn=r):
        self.stages, self.fn = set(stages), fn

    def __call__(self, x, stage):
        return self.fn(x) if stage in self.stages else x


def backbone_q(img, sd, cfg, q):
    st = "backbone"
    def cbr(x, w, bn, relu=True, **kw):
        y = O.frozen_bn(F.conv2d(q(x, st), q(sd[w], st), **kw), sd, bn)
        return F.relu(y) if relu else y
    p = "backbone.body"
    x = cbr(img, p + ".conv1.weight", p + ".bn1", stride=2, padding=3)
    x = F.max_pool2d(x, 3, 2, 1)
    dil = 1
    for li, nb in enumerate(O.RESNET_BLOCKS[cfg.backbone]):
        stride = 1 if li == 0 else 2
        fd = dil
        if li == 3 and cfg.dilation:
            dil *= stride; stride = 1
        for bi in range(nb):
            b = f"{p}.layer{li + 1}.{bi}"
            s_, d_ = (stride, fd) if bi == 0 else (1, dil)
            o = cbr(x, b + ".conv1.weight", b + ".bn1")
            o = cbr(o, b + ".conv2.weight", b + ".bn2", stride=s_, padding=d_, dilation=d_)
            o = cbr(o, b + ".conv3.weight", b + ".bn3", relu=False)
            idn = q(x, st)                      # the engine keeps activations in bf16 between layers
            if b + ".downsample.0.weight" in sd:
                idn = cbr(x, b + ".downsample.0.weight", b + ".downsample.1", relu=False, stride=s_)
            x = F.relu(o + idn)
    return q(x, st)


def lin(x, w, b, q, st):
    return F.linear(q(x, st), q(w, st), b)


def mha_q(q_in, k_in, v_in, sd, p, H, mask, causal, q, st, kv_stage=None):
    w, b = sd[p + ".in_proj_weight"], sd[p + ".in_proj_bias"]
    d = w.shape[1]; dh = d // H
    # positions enter as an fp32 per-row bias in the engine: W(x+p)+b = Wx + (Wp+b)
    kvs = kv_stage or st
    qq = lin(q_in[0], w[:d], b[:d], q, st) + F.linear(q_in[1], w[:d])
    kk = lin(k_in[0], w[d:2 * d], b[d:2 * d], q, kvs) + F.linear(k_in[1], w[d:2 * d])
    vv = lin(v_in, w[2 * d:], b[2 * d:], q, kvs)
    qq, kk, vv = q(qq, st), q(kk, kvs), q(vv, kvs)        # stored operands of the attention kernel
    B, Lq, _ = qq.shape; Lk = kk.shape[1]
    qh = qq.view(B, Lq, H, dh).transpose(1, 2) * dh ** -0.5
    kh = kk.view(B, Lk, H, dh).transpose(1, 2)
    vh = vv.view(B, Lk, H, dh).transpose(1, 2)
    s = qh @ kh.transpose(-1, -2)
    if causal:
        s = s + torch.full((Lq, Lk), float("-inf")).triu(1)
    s = s.masked_fill(mask[:, None, None, :], float("-inf"))
    pr = q(torch.softmax(s, -1), st + "_p")
    o = (pr @ vh).transpose(1, 2).reshape(B, Lq, d)
    return lin(o, sd[p + ".out_proj.weight"], sd[p + ".out_proj.bias"], q, st)


def ffn_q(x, sd, p, q, st):
    h = F.relu(lin(x, sd[p + ".0.weight"], sd[p + ".0.bias"], q, st))
    return lin(h, sd[p + ".2.weight"], sd[p + ".2.bias"], q, st)


def forward_q(sd, cfg, data, ids, id_mask, q):
    toks, masks = [], []
    for k in ("t", "g"):
        feat = backbone_q(data[k + "_img"], sd, cfg, q)
        m = F.interpolate(data[k + "_mask"][None].float(), size=feat.shape[-2:]).bool()[0]
        src = F.conv2d(q(feat, "encoder"), q(sd["input_proj.weight"], "encoder"), sd["input_proj.bias"])
        toks.append(src.flatten(2).transpose(1, 2)); masks.append(m.flatten(1))
    B = toks[0].shape[0]
    loc_tok = F.linear(data["loc"].unsqueeze(2), sd["loc_proj.weight"], sd["loc_proj.bias"])
    x = torch.cat([toks[0], loc_tok, toks[1]], 1)
    mask = torch.cat([masks[0], torch.zeros(B, loc_tok.shape[1], dtype=torch.bool), masks[1]], 1)
    pos = O.sine_table(sd, x.shape[1])
    for l in range(cfg.enc_layers):
        p = f"transformer.encoder.layers.{l}"
        n = O.layer_norm(x, sd, p + ".self_attn.norm")
        x = x + mha_q((n, pos), (n, pos), n, sd, p + ".self_attn.sublayer", cfg.nheads, mask, False, q, "encoder")
        x = x + ffn_q(O.layer_norm(x, sd, p + ".ff.norm"), sd, p + ".ff.sublayer", q, "encoder")
    memory = O.layer_norm(x, sd, "transformer.encoder.norm")
    qpos = sd["transformer.embeddings.position_embeddings.weight"]
    y = O.decoder_embeddings(ids, sd, cfg)
    zero = torch.zeros_like(pos)
    for l in range(cfg.dec_layers):
        p = f"transformer.decoder.layers.{l}"
        n = O.layer_norm(y, sd, p + ".tgt_self_attn.norm")
        y = y + mha_q((n, qpos), (n, qpos), n, sd, p + ".tgt_self_attn.sublayer", cfg.nheads, id_mask, True, q, "decoder")
        n = O.layer_norm(y, sd, p + ".tgt_src_cross_attn.norm")
        y = y + mha_q((n, qpos), (memory, pos), memory, sd, p + ".tgt_src_cross_attn.sublayer", cfg.nheads, mask,
                      False, q, "decoder", kv_stage="crosskv")
        y = y + ffn_q(O.layer_norm(y, sd, p + ".ff.norm"), sd, p + ".ff.sublayer", q, "decoder")
    hs = O.layer_norm(y, sd, "transformer.decoder.norm")
    h = F.relu(lin(hs, sd["mlp.layers.0.weight"], sd["mlp.layers.0.bias"], q, "head"))
    h = F.relu(lin(h, sd["mlp.layers.1.weight"], sd["mlp.layers.1.bias"], q, "head"))
    return lin(h, sd["mlp.layers.2.weight"], sd["mlp.layers.2.bias"], q, "head")


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    calibrated = os.environ.get("CALIBRATED", "1") == "1"
    cfg = Config(use_global_features=True, use_location_features=True)
    torch.manual_seed(42)
    with contextlib.redirect_stdout(io.StringIO()):
        model, _ = caption.build_model(cfg)
    sd = model.state_dict()
    if calibrated:
        sd = synthetic.calibrate_frozen_bn(sd)
    sd = {k: v.clone() for k, v in sd.items()}
    data = synthetic.make_batch(n, shard=3)
    with torch.no_grad():
        enc = (data["t_img"], data["t_mask"], data["g_img"], data["g_mask"], data["loc"])
        ids, _ = O.greedy(sd, cfg, *enc, literal=False)
        id_mask = torch.zeros_like(ids, dtype=torch.bool)
        ref = forward_q(sd, cfg, data, ids, id_mask, Q([]))
        chk = O.forward(sd, cfg, *enc, ids, id_mask)
        print("restatement vs oracle max diff", (ref - chk).abs().max().item())
        top2 = ref[:, :127].topk(2, -1).values
        margin = (top2[..., 0] - top2[..., 1])
        print(f"logit std {ref.std():.4f} max {ref.abs().max():.4f}; margin min {margin.min():.2e} "
              f"median {margin.median():.2e}; per-seq min margin median {margin.min(1).values.median():.2e}")
        am = ref[:, :127].argmax(-1)
        plans = {
            "backbone": ["backbone"], "encoder": ["encoder", "encoder_p"], "crosskv": ["crosskv"],
            "decoder": ["decoder", "decoder_p"], "head": ["head"],
            "all": ["backbone", "encoder", "encoder_p", "crosskv", "decoder", "decoder_p", "head"],
            "all-but-backbone": ["encoder", "encoder_p", "crosskv", "decoder", "decoder_p", "head"],
            "backbone+encoder": ["backbone", "encoder", "encoder_p"],
        }
        for name, stages in plans.items():
            for fn_name, fn in (("bf16", r), ("split2", split2)):
                if fn_name == "split2" and name not in ("all", "backbone", "backbone+encoder"):
                    continue
                out = forward_q(sd, cfg, data, ids, id_mask, Q(stages, fn))
                err = (out - ref).abs()
                agree = out[:, :127].argmax(-1) == am
                print(f"{name:18s} {fn_name:6s} max|dlogit| {err.max():.2e} rms {err.pow(2).mean().sqrt():.2e}  "
                      f"step agreement {agree.float().mean() * 100:.2f}%  sequences exact "
                      f"{int(agree.all(1).sum())}/{n}", flush=True)


if __name__ == "__main__":
    main()
