"""Generate tests/golden/*.pt from the UNMODIFIED reference (run in the build container only).

    python oracle/gen_golden.py            # needs /root/reference; writes tests/golden/

Imports /root/reference read-only, patches ``models.backbone.is_main_process`` (the only change:
it skips the ImageNet-weights download at models/backbone.py:87-88, there is no network), builds
the reference model with ``torch.manual_seed(42)`` and records, for seeded synthetic inputs:
  * a fingerprint of the reference state_dict (per-tensor float64 sums for a key sample + global
    sum) — pins ``retr_b200.models.caption.build_model``'s init replay;
  * logits / attention slices of ``model(*samples, caption, cap_mask, return_attention=True)``;
  * ``eval_utils.decode.greedy`` token rows;
  * loss + a few gradient norms of the ``engine.train_one_epoch`` arithmetic (dropout 0).
Fixtures are kilobytes; the 336 MB of weights are never stored (they are re-derived from the seed).
"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def ref_modules():
    sys.path.insert(0, REF)
    import models.backbone as ref_bb
    ref_bb.is_main_process = lambda: False
    from configuration_template import Config as RefConfig
    from models import caption as ref_caption
    from models.utils import NestedTensor as RefNested
    from eval_utils import decode as ref_decode
    return RefConfig, ref_caption, RefNested, ref_decode


def fingerprint(sd):
    keys = sorted(sd.keys())
    sample = keys[:: max(1, len(keys) // 48)]
    return {
        "n_entries": len(keys),
        "keys": keys,
        "shapes": {k: tuple(sd[k].shape) for k in keys},
        "sums": {k: float(sd[k].double().sum()) for k in sample},
        "abs_sum_total": float(sum(sd[k].double().abs().sum() for k in keys)),
    }


def main():
    from retr_b200 import synthetic
    RefConfig, ref_caption, RefNested, ref_decode = ref_modules()
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)

    for kind, flags in (("CaptionGlobalLoc", (True, True)), ("Caption", (False, False)),
                        ("CaptionLoc", (False, True))):
        cfg = RefConfig()
        cfg.use_global_features, cfg.use_location_features = flags
        torch.manual_seed(42)
        model, criterion = ref_caption.build_model(cfg)
        model.eval()
        sd = model.state_dict()
        B = 2
        data = synthetic.make_batch(B, shard=0, train=True)
        t = RefNested(data["t_img"], data["t_mask"])
        gl = RefNested(data["g_img"], data["g_mask"])
        loc7 = torch.cat([data["loc"], data["loc"][:, :2] * 0.5 + 0.25], 1)
        samples = {"CaptionGlobalLoc": (t, gl, data["loc"]), "Caption": (t,), "CaptionLoc": (t, loc7)}[kind]
        caps, cap_masks = data["caps"], data["cap_masks"]
        fx = {"kind": kind, "batch": B, "fingerprint": fingerprint(sd)}
        with torch.no_grad():
            t0 = time.perf_counter()
            logits, att = model(*samples, caps[:, :-1], cap_masks[:, :-1], return_attention=True)
            fx["forward_seconds"] = time.perf_counter() - t0
        fx["logits_shape"] = tuple(logits.shape)
        # slices: first 12 positions x a strided vocab sample + per-position argmax/top2 margin
        vidx = torch.arange(0, logits.shape[-1], 97)
        fx["vocab_idx"] = vidx
        fx["logits_slice"] = logits[:, :12, :][:, :, vidx].clone()
        top2 = logits.topk(2, dim=-1)
        fx["argmax"] = top2.indices[..., 0].clone()
        fx["margin"] = (top2.values[..., 0] - top2.values[..., 1]).clone()
        fx["logits_absmax"] = float(logits.abs().max())
        fx["logits_rowsum"] = logits.double().sum(-1).float()
        fx["att_shapes"] = {k: tuple(v.shape) for k, v in att.items()}
        fx["att_slices"] = {k: v[:, :, :6, :].clone() if k != "enc_tc_self_att" else v[:, :, 190:204, :].clone()
                            for k, v in att.items()}
        if kind == "CaptionGlobalLoc":
            # greedy, the literal reference loop (eval_utils/decode.py:53-81), first 6 steps timed
            with torch.no_grad():
                cap8 = ref_decode.greedy(samples, _StepLimited(model, 6), max_len=128, device="cpu",
                                         bos_token=101, eos_token=102)
            fx["greedy_first6"] = cap8[:, :7].clone()
            # train arithmetic (engine.py:70-83) with dropout disabled
            cfg0 = RefConfig(); cfg0.use_global_features = cfg0.use_location_features = True
            cfg0.dropout = 0.0
            torch.manual_seed(42)
            m0, crit = ref_caption.build_model(cfg0)
            m0.train()
            out = m0(*samples, caps[:, :-1], cap_masks[:, :-1])
            loss = crit(out.permute(0, 2, 1), caps[:, 1:])
            loss.backward()
            fx["train_loss"] = float(loss)
            gn = {}
            for n, p in m0.named_parameters():
                if p.grad is not None and (n.startswith("mlp") or "layers.0." in n or "input_proj" in n
                                           or "layer4.2.conv3" in n or "layer2.0.conv1" in n
                                           or n.startswith("loc_proj") or "embeddings" in n):
                    gn[n] = float(p.grad.double().norm())
            fx["grad_norms"] = gn
            fx["grad_total_norm"] = float(torch.sqrt(sum((p.grad.double() ** 2).sum()
                                                         for p in m0.parameters() if p.grad is not None)))
        path = os.path.join(OUT, f"{kind}.pt")
        torch.save(fx, path)
        print(kind, "->", path, os.path.getsize(path), "bytes; forward", round(fx["forward_seconds"], 2), "s")


def learned_variant():
    """position_embedding='learned' (models/position_encoding.py:38-63): eval-mode logits of the reference."""
    from retr_b200 import synthetic
    RefConfig, ref_caption, RefNested, _ = ref_modules()
    cfg = RefConfig()
    cfg.use_global_features = cfg.use_location_features = True
    cfg.position_embedding = "learned"
    torch.manual_seed(42)
    model, _ = ref_caption.build_model(cfg)
    model.eval()
    data = synthetic.make_batch(2, shard=0, train=True)
    samples = (RefNested(data["t_img"], data["t_mask"]), RefNested(data["g_img"], data["g_mask"]), data["loc"])
    with torch.no_grad():
        logits = model(*samples, data["caps"][:, :-1], data["cap_masks"][:, :-1])
    vidx = torch.arange(0, logits.shape[-1], 97)
    top2 = logits.topk(2, dim=-1)
    fx = {"kind": "CaptionGlobalLoc", "position_embedding": "learned", "batch": 2,
          "fingerprint": fingerprint(model.state_dict()), "vocab_idx": vidx,
          "logits_slice": logits[:, :12, :][:, :, vidx].clone(), "argmax": top2.indices[..., 0].clone(),
          "margin": (top2.values[..., 0] - top2.values[..., 1]).clone(), "logits_absmax": float(logits.abs().max())}
    path = os.path.join(OUT, "CaptionGlobalLoc_learned.pt")
    torch.save(fx, path)
    print("learned ->", path, os.path.getsize(path), "bytes")


def basic_block_variant():
    """backbone='ResNet18', dilation=False (models/backbone.py:86-93: BasicBlock, 512 channels, 7x7 = 49 tokens)."""
    from retr_b200 import synthetic
    RefConfig, ref_caption, RefNested, _ = ref_modules()
    cfg = RefConfig()
    cfg.use_global_features = cfg.use_location_features = True
    cfg.backbone, cfg.dilation = "ResNet18", False
    torch.manual_seed(42)
    model, _ = ref_caption.build_model(cfg)
    model.eval()
    data = synthetic.make_batch(2, shard=0, train=True)
    samples = (RefNested(data["t_img"], data["t_mask"]), RefNested(data["g_img"], data["g_mask"]), data["loc"])
    with torch.no_grad():
        logits = model(*samples, data["caps"][:, :-1], data["cap_masks"][:, :-1])
    vidx = torch.arange(0, logits.shape[-1], 97)
    top2 = logits.topk(2, dim=-1)
    fx = {"kind": "CaptionGlobalLoc", "backbone": "ResNet18", "dilation": False, "batch": 2,
          "fingerprint": fingerprint(model.state_dict()), "vocab_idx": vidx,
          "logits_slice": logits[:, :12, :][:, :, vidx].clone(), "argmax": top2.indices[..., 0].clone(),
          "margin": (top2.values[..., 0] - top2.values[..., 1]).clone(), "logits_absmax": float(logits.abs().max())}
    path = os.path.join(OUT, "CaptionGlobalLoc_resnet18.pt")
    torch.save(fx, path)
    print("resnet18 ->", path, os.path.getsize(path), "bytes")


class _StepLimited:
    """Wraps the reference model; after ``n`` calls raises StopIteration-like exit by making every
    row emit EOS (the reference loop then returns the caption as filled so far)."""

    def __init__(self, model, n):
        self.model, self.n, self.calls = model, n, 0

    def __call__(self, *a):
        self.calls += 1
        out = self.model(*a)
        if self.calls > self.n:
            out = out.clone()
            out[..., 102] = 1e9
        return out


if __name__ == "__main__":
    if "--learned-only" in sys.argv:
        learned_variant()
    elif "--resnet18-only" in sys.argv:
        basic_block_variant()
    else:
        main()
        learned_variant()
        basic_block_variant()
