"""Pins (1) the product's init replay and (2) the oracle's arithmetic to fixtures produced from
the unmodified reference (oracle/gen_golden.py).  CPU only."""
import pytest
import torch

from conftest import build_product_model, golden, synthetic_samples
from oracle import retr_oracle as O

KINDS = ["CaptionGlobalLoc", "Caption", "CaptionLoc"]


@pytest.mark.parametrize("kind", KINDS)
def test_init_replay_matches_reference_fingerprint(kind):
    model, _, _ = build_product_model(kind)
    sd = model.state_dict()
    fp = golden(kind)["fingerprint"]
    assert sorted(sd.keys()) == fp["keys"]
    assert len(sd) == fp["n_entries"]
    for k, shape in fp["shapes"].items():
        assert tuple(sd[k].shape) == tuple(shape), k
    for k, s in fp["sums"].items():
        assert float(sd[k].double().sum()) == s, k            # bit-identical init
    total = float(sum(sd[k].double().abs().sum() for k in sorted(sd.keys())))
    assert total == fp["abs_sum_total"]


@pytest.mark.parametrize("kind", KINDS)
def test_oracle_forward_matches_reference(kind):
    model, _, cfg = build_product_model(kind)
    sd = model.state_dict()
    fx = golden(kind)
    data, enc_in = synthetic_samples(kind, fx["batch"])
    caps, cm = data["caps"], data["cap_masks"]
    with torch.no_grad():
        logits, att = O.forward(sd, cfg, *enc_in, caps[:, :-1], cm[:, :-1], return_attention=True)
    assert tuple(logits.shape) == fx["logits_shape"]
    ref = fx["logits_slice"]
    got = logits[:, :12, :][:, :, fx["vocab_idx"]]
    scale = fx["logits_absmax"]
    assert (got - ref).abs().max() <= 2e-5 * scale            # fp32 re-association noise only
    assert torch.equal(logits.argmax(-1)[fx["margin"] > 1e-4 * scale], fx["argmax"][fx["margin"] > 1e-4 * scale])
    assert torch.allclose(logits.double().sum(-1).float(), fx["logits_rowsum"], rtol=0, atol=2e-3 * scale)
    for k, shape in fx["att_shapes"].items():
        assert tuple(att[k].shape) == tuple(shape)
        sl = att[k][:, :, :6, :] if k != "enc_tc_self_att" else att[k][:, :, 190:204, :]
        assert (sl - fx["att_slices"][k]).abs().max() < 1e-5, k


def test_oracle_greedy_matches_reference_loop():
    model, _, cfg = build_product_model("CaptionGlobalLoc")
    sd = model.state_dict()
    fx = golden("CaptionGlobalLoc")
    _, enc_in = synthetic_samples("CaptionGlobalLoc", fx["batch"])
    with torch.no_grad():
        cap, steps = O.greedy(sd, cfg, *enc_in, literal=False, max_steps=6)
    assert steps == 6
    assert torch.equal(cap[:, :7], fx["greedy_first6"])


def test_oracle_cached_equals_literal_one_step():
    """Encode-once decode is the same function as the literal full recompute."""
    model, _, cfg = build_product_model("Caption")
    sd = model.state_dict()
    _, enc_in = synthetic_samples("Caption", 1)
    with torch.no_grad():
        a, _ = O.greedy(sd, cfg, *enc_in, literal=True, max_steps=2)
        b, _ = O.greedy(sd, cfg, *enc_in, literal=False, max_steps=2)
    assert torch.equal(a, b)


def test_oracle_train_loss_and_grads_match_reference():
    model, _, cfg = build_product_model("CaptionGlobalLoc")
    fx = golden("CaptionGlobalLoc")
    sd = {k: v.clone().requires_grad_(v.is_floating_point() and "running" not in k and "bn" not in k
                                      and "downsample.1" not in k and not k.endswith(".pe"))
          for k, v in model.state_dict().items()}
    data, enc_in = synthetic_samples("CaptionGlobalLoc", fx["batch"])
    loss = O.train_loss(sd, cfg, *enc_in, data["caps"], data["cap_masks"])
    assert abs(float(loss) - fx["train_loss"]) < 2e-5 * abs(fx["train_loss"])
    loss.backward()
    for name, gn in fx["grad_norms"].items():
        got = float(sd[name].grad.double().norm())
        assert abs(got - gn) <= 2e-3 * gn + 1e-9, (name, got, gn)


def test_prune_cap_ids_semantics():
    rows = [[101, 5, 6, 102, 7, 0], [101, 9, 0, 0, 0, 0]]
    assert O.prune_cap_ids(rows, True, 0, 101, 102) == [[5, 6], [9]]
    assert O.prune_cap_ids(rows, False, 0, 101, 102) == [[101, 5, 6, 102], [101, 9, 0, 0, 0, 0]]


def test_learned_position_variant_init_and_oracle_forward():
    """position_embedding='learned' (models/position_encoding.py:38-63): init replay and the
    oracle's eval-mode forward against the fixture recorded from the unmodified reference."""
    model, _, cfg = build_product_model("CaptionGlobalLoc", position_embedding="learned")
    sd = model.state_dict()
    fx = golden("CaptionGlobalLoc_learned")
    fp = fx["fingerprint"]
    assert sorted(sd.keys()) == fp["keys"]
    assert "transformer.positional_encoding.pos_embed.weight" in sd
    for k, s in fp["sums"].items():
        assert float(sd[k].double().sum()) == s, k
    assert float(sum(sd[k].double().abs().sum() for k in sorted(sd.keys()))) == fp["abs_sum_total"]
    data, enc_in = synthetic_samples("CaptionGlobalLoc", fx["batch"])
    with torch.no_grad():
        logits = O.forward(sd, cfg, *enc_in, data["caps"][:, :-1], data["cap_masks"][:, :-1])
    scale = fx["logits_absmax"]
    got = logits[:, :12, :][:, :, fx["vocab_idx"]]
    assert (got - fx["logits_slice"]).abs().max() <= 2e-5 * scale
    sure = fx["margin"] > 1e-4 * scale
    assert torch.equal(logits.argmax(-1)[sure], fx["argmax"][sure])


def test_basic_block_backbone_init_and_oracle_forward():
    """backbone='ResNet18', dilation=False (models/backbone.py:86-93): BasicBlock init replay
    (bit-identical state_dict) and the oracle's forward against the fixture recorded from the unmodified reference."""
    model, _, cfg = build_product_model("CaptionGlobalLoc", backbone="ResNet18", dilation=False)
    sd = model.state_dict()
    fx = golden("CaptionGlobalLoc_resnet18")
    fp = fx["fingerprint"]
    assert sorted(sd.keys()) == fp["keys"]
    assert tuple(sd["input_proj.weight"].shape) == (256, 512, 1, 1)
    for k, s in fp["sums"].items():
        assert float(sd[k].double().sum()) == s, k
    assert float(sum(sd[k].double().abs().sum() for k in sorted(sd.keys()))) == fp["abs_sum_total"]
    data, enc_in = synthetic_samples("CaptionGlobalLoc", fx["batch"])
    with torch.no_grad():
        src, mask = O.encoder_inputs(sd, cfg, *enc_in)
        logits = O.forward(sd, cfg, *enc_in, data["caps"][:, :-1], data["cap_masks"][:, :-1])
    assert src.shape[1] == 49 + 5 + 49                      # 7x7 tokens per image without dilation
    scale = fx["logits_absmax"]
    got = logits[:, :12, :][:, :, fx["vocab_idx"]]
    assert (got - fx["logits_slice"]).abs().max() <= 2e-5 * scale
    sure = fx["margin"] > 1e-4 * scale
    assert torch.equal(logits.argmax(-1)[sure], fx["argmax"][sure])
    with pytest.raises(NotImplementedError):                # torchvision's BasicBlock refuses dilation
        build_product_model("CaptionGlobalLoc", backbone="ResNet34", dilation=True)
