"""Training-step parity (reference engine.py:52-87): loss, every parameter gradient, clip + AdamW update of
the native path against the CPU oracle's autograd on the same seeded weights and inputs.  GPU only.

fp32 mode: per-parameter gradient error <= 2e-3 of the gradient's norm (accumulation-order noise only).
bf16 mode: global gradient error <= 6e-2 relative, per-parameter cosine similarity >= 0.97 for every
parameter whose reference gradient is not negligible (the tolerance is bf16 rounding of activations and
of the back-propagated signals through 33 bottlenecks + 12 transformer layers).
Dropout cannot be RNG-matched with torch, so parity runs use dropout 0; the dropout
path is checked for determinism, keep-rate and train/eval consistency.
"""
import copy
import os
import sys

import pytest
import torch
import torch.nn.functional as F

from conftest import build_product_model, synthetic_samples
from oracle import retr_oracle as O
from retr_b200 import engine as E
from retr_b200 import synthetic
from retr_b200.models.utils import NestedTensor

pytestmark = pytest.mark.gpu
KINDS = ["CaptionGlobalLoc", "Caption", "CaptionLoc"]


def train_model(kind, precision, dropout=0.0, **cfg_over):
    model, _, cfg = build_product_model(kind, **cfg_over)
    model = copy.deepcopy(model)
    model.load_state_dict(synthetic.calibrate_frozen_bn(model.state_dict()))
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    model.precision = precision
    model.dropout_p = dropout
    return model.to("cuda").train(), sd, cfg


def samples_of(enc_in, device="cuda"):
    t_img, t_mask, g_img, g_mask, loc = enc_in
    s = [NestedTensor(t_img, t_mask).to(device)]
    if g_img is not None:
        s.append(NestedTensor(g_img, g_mask).to(device))
    if loc is not None:
        s.append(loc.to(device))
    return tuple(s)


def oracle_grads(model, sd, cfg, enc_in, caps, cap_masks):
    names = [n for n, p in model.named_parameters() if p.requires_grad]
    sdg = {k: v.clone().requires_grad_(k in names) for k, v in sd.items()}
    loss = O.train_loss(sdg, cfg, *enc_in, caps, cap_masks)
    loss.backward()
    return float(loss), {n: sdg[n].grad for n in names}


def native_grads(model, enc_in, caps, cap_masks):
    model.zero_grad()
    out = model(*samples_of(enc_in), caps[:, :-1].cuda(), cap_masks[:, :-1].cuda())
    loss = F.cross_entropy(out.permute(0, 2, 1), caps[:, 1:].cuda())
    loss.backward()
    return float(loss), {n: p.grad.detach().cpu() for n, p in model.named_parameters() if p.requires_grad}


def compare(ref, got):
    rows = []
    for n, g in ref.items():
        a, b = g.double().flatten(), got[n].double().flatten()
        na = float(a.norm())
        err = float((a - b).norm()) / (na + 1e-30)
        cos = float(torch.dot(a, b) / (na * float(b.norm()) + 1e-30))
        rows.append((n, na, err, cos))
    return rows


@pytest.mark.parametrize("kind", KINDS)
def test_fp32_loss_and_grads_match_oracle(kind):
    model, sd, cfg = train_model(kind, "fp32")
    data, enc_in = synthetic_samples(kind, 2, shard=5)
    ref_loss, ref = oracle_grads(model, sd, cfg, enc_in, data["caps"], data["cap_masks"])
    loss, got = native_grads(model, enc_in, data["caps"], data["cap_masks"])
    assert abs(loss - ref_loss) <= 2e-5 * abs(ref_loss), (loss, ref_loss)
    gmax = max(r[1] for r in compare(ref, got))
    for n, na, err, cos in compare(ref, got):
        if na > 1e-6 * gmax:
            assert err <= 2e-3, (n, na, err, cos)
    # frozen parts carry no gradient (models/backbone.py:58-60)
    assert all(p.grad is None for n, p in model.named_parameters() if not p.requires_grad)


def test_fp32_grads_without_final_encoder_norm():
    """pre_norm=False only drops the final encoder LayerNorm (models/ConcatTransformer.py:24)."""
    kind = "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, "fp32", pre_norm=False)
    assert "transformer.encoder.norm.weight" not in sd
    data, enc_in = synthetic_samples(kind, 2, shard=6)
    ref_loss, ref = oracle_grads(model, sd, cfg, enc_in, data["caps"], data["cap_masks"])
    loss, got = native_grads(model, enc_in, data["caps"], data["cap_masks"])
    assert abs(loss - ref_loss) <= 2e-5 * abs(ref_loss), (loss, ref_loss)
    gmax = max(r[1] for r in compare(ref, got))
    for n, na, err, cos in compare(ref, got):
        if na > 1e-6 * gmax:
            assert err <= 2e-3, (n, na, err, cos)


def test_fp32_grads_basic_block_backbone():
    """ResNet-18 (BasicBlock, dilation=False): layer2-4 gradients through 3x3 (stride 2) + 3x3 blocks vs oracle autograd."""
    kind = "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, "fp32", backbone="ResNet18", dilation=False)
    data, enc_in = synthetic_samples(kind, 2, shard=7)
    ref_loss, ref = oracle_grads(model, sd, cfg, enc_in, data["caps"], data["cap_masks"])
    loss, got = native_grads(model, enc_in, data["caps"], data["cap_masks"])
    assert abs(loss - ref_loss) <= 2e-5 * abs(ref_loss), (loss, ref_loss)
    assert any("backbone.body.layer2" in n for n in got)
    gmax = max(r[1] for r in compare(ref, got))
    for n, na, err, cos in compare(ref, got):
        if na > 1e-6 * gmax:
            assert err <= 2e-3, (n, na, err, cos)
    m16, _, _ = train_model(kind, "bf16", backbone="ResNet18", dilation=False)
    loss16, got16 = native_grads(m16, enc_in, data["caps"], data["cap_masks"])
    assert abs(loss16 - ref_loss) <= 2e-3 * abs(ref_loss)
    rows = compare(ref, got16)
    num = sum((r[1] * r[2]) ** 2 for r in rows) ** 0.5
    den = sum(r[1] ** 2 for r in rows) ** 0.5
    assert num / den <= 6e-2, num / den


def test_fp32_grads_learned_positions():
    """position_embedding='learned' in train mode (models/position_encoding.py:38-63): LN(pos_embed[:S]) feeds every
    encoder q/k input and the cross-attention keys; gradients of the table and its LayerNorm vs oracle autograd
    (dropout 0; the reference's hard-coded 0.1 dropout of the table follows the step's dropout switch here)."""
    kind = "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, "fp32", position_embedding="learned")
    data, enc_in = synthetic_samples(kind, 2, shard=8)
    ref_loss, ref = oracle_grads(model, sd, cfg, enc_in, data["caps"], data["cap_masks"])
    loss, got = native_grads(model, enc_in, data["caps"], data["cap_masks"])
    assert abs(loss - ref_loss) <= 2e-5 * abs(ref_loss), (loss, ref_loss)
    rows = compare(ref, got)
    assert any(n == "transformer.positional_encoding.pos_embed.weight" and na > 0 for n, na, _, _ in rows)
    gmax = max(r[1] for r in rows)
    for n, na, err, cos in rows:
        if na > 1e-6 * gmax:
            assert err <= 2e-3, (n, na, err, cos)
    assert float(got["transformer.positional_encoding.pos_embed.weight"][397:].abs().max()) == 0.0     # rows >= S unused
    # with dropout on, the per-sample table dropout is a pure function of the seed: two steps, same seed, same gradients
    model.dropout_p = 0.1
    s = samples_of(enc_in)
    t, g, loc = model._split_samples(s)
    inputs = model._train_inputs(t, g, loc, data["caps"][:, :-1].cuda(), data["cap_masks"][:, :-1].cuda())
    tr = model.trainer(2)
    outs = []
    for seed in (5, 5, 6):
        tr.loss_and_grads(*inputs, data["caps"][:, 1:].cuda(), dropout_p=0.1, seed=seed)
        torch.cuda.synchronize()
        outs.append(tr.grads["transformer.positional_encoding.pos_embed.weight"].clone())
    assert (outs[0] - outs[1]).abs().max().item() <= 1e-6 * outs[0].abs().max().item()
    assert (outs[0] - outs[2]).abs().max().item() > 1e-3 * outs[0].abs().max().item()


def test_bf16_grads_close_to_oracle():
    kind = "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, "bf16")
    data, enc_in = synthetic_samples(kind, 2, shard=5)
    ref_loss, ref = oracle_grads(model, sd, cfg, enc_in, data["caps"], data["cap_masks"])
    loss, got = native_grads(model, enc_in, data["caps"], data["cap_masks"])
    assert abs(loss - ref_loss) <= 2e-3 * abs(ref_loss), (loss, ref_loss)
    rows = compare(ref, got)
    num = sum((r[1] * r[2]) ** 2 for r in rows) ** 0.5
    den = sum(r[1] ** 2 for r in rows) ** 0.5
    assert num / den <= 6e-2, num / den
    gmax = max(r[1] for r in rows)
    for n, na, err, cos in rows:
        if na > 1e-3 * gmax:
            assert cos >= 0.97, (n, na, err, cos)


def test_fused_step_matches_torch_clip_adamw():
    """retr_train_loss_and_grads + retr_train_clip_adamw == autograd path + clip_grad_norm_ + torch AdamW."""
    kind = "CaptionGlobalLoc"
    data, enc_in = synthetic_samples(kind, 2, shard=6)
    caps, masks = data["caps"], data["cap_masks"]
    a, _, cfg = train_model(kind, "fp32")
    b = copy.deepcopy(a)
    groups = [{"params": [p for n, p in a.named_parameters() if "backbone" not in n and p.requires_grad]},
              {"params": [p for n, p in a.named_parameters() if "backbone" in n and p.requires_grad], "lr": cfg.lr_backbone}]
    opt = torch.optim.AdamW(groups, lr=cfg.lr, weight_decay=cfg.weight_decay)
    for step in range(2):
        out = a(*samples_of(enc_in), caps[:, :-1].cuda(), masks[:, :-1].cuda())
        loss_a = F.cross_entropy(out.permute(0, 2, 1), caps[:, 1:].cuda())
        opt.zero_grad()
        loss_a.backward()
        norm_a = torch.nn.utils.clip_grad_norm_(a.parameters(), cfg.clip_max_norm)
        opt.step()
        loss_b, norm_b = E.train_step_native(b, samples_of(enc_in), caps.cuda(), masks.cuda(), max_norm=cfg.clip_max_norm,
                                             lr=cfg.lr, lr_backbone=cfg.lr_backbone, weight_decay=cfg.weight_decay)
        assert abs(float(loss_a) - float(loss_b)) <= 1e-4 * abs(float(loss_a)), (step, float(loss_a), float(loss_b))
        assert abs(float(norm_a) - float(norm_b)) <= 2e-3 * float(norm_a), (step, float(norm_a), float(norm_b))
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        if pa.requires_grad:
            # two AdamW steps move each weight by <= 2*lr; the two paths must agree to a fraction of that
            assert float((pa - pb).abs().max()) <= 0.2 * cfg.lr, n


def test_dropout_path():
    kind = "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, "bf16", dropout=0.1)
    data, enc_in = synthetic_samples(kind, 2, shard=7)
    args = samples_of(enc_in) + (data["caps"][:, :-1].cuda(), data["cap_masks"][:, :-1].cuda())
    torch.manual_seed(3)
    a = model(*args).detach()
    torch.manual_seed(3)
    b = model(*args).detach()
    torch.manual_seed(4)
    c = model(*args).detach()
    assert torch.equal(a, b)                      # same seed -> same Philox stream -> bit-identical
    assert not torch.equal(a, c)
    assert torch.isfinite(a).all()
    model.dropout_p = 0.0
    d = model(*args).detach()
    model.eval()
    with torch.no_grad():
        e = model(*args)
    scale = float(e.abs().max())
    assert float((d - e).abs().max()) <= 2e-2 * scale        # p = 0 train forward == eval forward (bf16 tolerance)
    # the dropped forward is a perturbation of the clean one, not garbage
    assert float((a - d).abs().max()) <= 0.75 * scale + 1.0


def _fused_inputs(model, enc_in, data):
    t, g, loc = model._split_samples(samples_of(enc_in))
    return model._train_inputs(t, g, loc, data["caps"][:, :-1].cuda(), data["cap_masks"][:, :-1].cuda())


def test_dropout_backward_matches_finite_differences_fp32():
    """With a fixed Philox seed the dropped forward is a deterministic function of the weights, so the backward
    pass (which REGENERATES every mask) must agree with a central finite difference of the loss."""
    kind = "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, "fp32", dropout=0.1)
    data, enc_in = synthetic_samples(kind, 2, shard=8)
    tr = model.trainer(2)
    inputs = _fused_inputs(model, enc_in, data)
    targets = data["caps"][:, 1:].cuda()

    def loss():
        return float(tr.loss_and_grads(*inputs, targets, dropout_p=0.1, seed=1234))

    loss()
    g = {n: tr.grads[n].clone() for n in tr.trainable}
    gnorm = sum(float(v.double().pow(2).sum()) for v in g.values()) ** 0.5
    params = dict(model.named_parameters())
    eps = 2e-2
    with torch.no_grad():
        for n, v in g.items():
            params[n].add_(v, alpha=eps / gnorm)
        lp = loss()
        for n, v in g.items():
            params[n].add_(v, alpha=-2 * eps / gnorm)
        lm = loss()
    fd = (lp - lm) / (2 * eps)
    assert abs(fd - gnorm) <= 0.03 * gnorm, (fd, gnorm)


def test_flash_training_attention_matches_simt_with_dropout():
    """Same seed, same mask function: the tensor-core flash attention (bf16 default) and the stored-probability
    SIMT kernels (RETR_TRAIN_SIMT_ATTENTION=1) must give the same loss and gradients up to bf16 rounding."""
    kind = "CaptionGlobalLoc"
    data, enc_in = synthetic_samples(kind, 2, shard=9)
    targets = data["caps"][:, 1:].cuda()
    out = {}
    for mode in ("simt", "flash"):
        if mode == "simt":
            os.environ["RETR_TRAIN_SIMT_ATTENTION"] = "1"
        try:
            model, sd, cfg = train_model(kind, "bf16", dropout=0.1)
            tr = model.trainer(2)
        finally:
            os.environ.pop("RETR_TRAIN_SIMT_ATTENTION", None)
        l = float(tr.loss_and_grads(*_fused_inputs(model, enc_in, data), targets, dropout_p=0.1, seed=77))
        out[mode] = (l, {n: tr.grads[n].clone().cpu() for n in tr.trainable})
    assert abs(out["simt"][0] - out["flash"][0]) <= 2e-3 * abs(out["simt"][0]), (out["simt"][0], out["flash"][0])
    rows = compare(out["simt"][1], out["flash"][1])
    num = sum((r[1] * r[2]) ** 2 for r in rows) ** 0.5
    den = sum(r[1] ** 2 for r in rows) ** 0.5
    assert num / den <= 5e-2, num / den


class _FakeLoader:
    """The dataloader contract of engine.train_one_epoch (reference engine.py:60-64)."""

    class _DS:
        return_global_context = True
        return_location_features = True

    def __init__(self, batches):
        self.batches, self.dataset = batches, self._DS()

    def __len__(self):
        return len(self.batches)

    def __iter__(self):
        return iter(self.batches)


def test_train_one_epoch_drop_in_reduces_loss():
    kind = "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, "bf16", dropout=0.1)
    batches = []
    for shard in (11, 12):
        data, _ = synthetic_samples(kind, 2, shard=shard)
        batches.append((list(range(2)), data["t_img"], data["t_mask"], data["g_img"], data["g_mask"], data["loc"],
                        data["caps"], data["cap_masks"]))
    loader = _FakeLoader(batches)
    groups = [{"params": [p for n, p in model.named_parameters() if "backbone" not in n and p.requires_grad]},
              {"params": [p for n, p in model.named_parameters() if "backbone" in n and p.requires_grad], "lr": cfg.lr_backbone}]
    opt = torch.optim.AdamW(groups, lr=cfg.lr, weight_decay=cfg.weight_decay)
    crit = torch.nn.CrossEntropyLoss()
    losses = [E.train_one_epoch(model, crit, loader, opt, "cuda", ep, cfg.clip_max_norm) for ep in range(4)]
    assert all(l == l for l in losses)
    assert losses[-1] < losses[0], losses


if __name__ == "__main__":
    # diagnostic table: python tests/test_train_gpu.py [fp32|bf16] [kind]
    prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
    kind = sys.argv[2] if len(sys.argv) > 2 else "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, prec)
    data, enc_in = synthetic_samples(kind, 2, shard=5)
    ref_loss, ref = oracle_grads(model, sd, cfg, enc_in, data["caps"], data["cap_masks"])
    loss, got = native_grads(model, enc_in, data["caps"], data["cap_masks"])
    print("loss", loss, "ref", ref_loss)
    rows = compare(ref, got)
    for n, na, err, cos in rows:
        flag = "  <<<" if err > (2e-3 if prec == "fp32" else 1e-1) else ""
        print(f"{n:70s} |g|={na:10.3e} err={err:9.2e} cos={cos:8.5f}{flag}")


def test_native_nccl_data_parallel():
    """the training step's one real collective.  Needs >= 2 GPUs (skipped on a 1-GPU box); launches
    tests/dp_nccl_check.py under torchrun: bucketed + overlapped native all-reduce == one all-reduce after the pass."""
    import os
    import subprocess
    import sys
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29611", os.path.join(here, "dp_nccl_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "DP_NCCL_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_bf16_grads_at_bench_batch_equal_mean_of_chunks():
    """The training bench runs B = 128 per GPU; parity against the oracle is checked at B = 2.  Samples are independent
    and the loss is the mean over B*T positions, so the gradient of a 128-sample batch must equal the mean of the
    gradients of its four 32-sample chunks (dropout 0): exercises the large-batch tiling / chunking of every kernel of
    the backward pass.  Both sides are the bf16 path, so the tolerance is bf16 accumulation-order noise."""
    kind = "CaptionGlobalLoc"
    model, sd, cfg = train_model(kind, "bf16")
    model.max_train_batch = 128
    B, C = 128, 32
    data = synthetic.make_batch(B, shard=40, train=True)
    caps, masks = data["caps"].cuda(), data["cap_masks"].cuda()

    def grads_of(lo, hi):
        sub = {k: v[lo:hi] for k, v in data.items()}
        samples = tuple(s.to("cuda") for s in synthetic.pack(sub))
        t, g, loc = model._split_samples(samples)
        inputs = model._train_inputs(t, g, loc, caps[lo:hi, :-1], masks[lo:hi, :-1])
        tr = model.trainer(hi - lo)
        loss = tr.loss_and_grads(*inputs, caps[lo:hi, 1:], dropout_p=0.0, seed=1)
        torch.cuda.synchronize()
        return float(loss), tr.gflat.clone()

    loss_all, g_all = grads_of(0, B)
    parts = [grads_of(i, i + C) for i in range(0, B, C)]
    loss_mean = sum(p[0] for p in parts) / len(parts)
    g_mean = torch.stack([p[1] for p in parts]).mean(0)
    assert abs(loss_all - loss_mean) <= 2e-3 * abs(loss_mean), (loss_all, loss_mean)
    rel = float((g_all - g_mean).norm() / g_mean.norm())
    cos = float(torch.dot(g_all, g_mean) / (g_all.norm() * g_mean.norm()))
    print(f"B=128 vs mean of 4x32: relative gradient difference {rel:.3e}, cosine {cos:.6f}")
    assert rel <= 3e-2 and cos >= 0.999, (rel, cos)
