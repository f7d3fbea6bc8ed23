"""Offline checkpoint repack (retr_b200/checkpoint.py): a file saved the way the reference's save_ckp does
(train_utils/checkpoints.py:4-21) -> packed file -> model, bit-identical weights; mismatches fail loudly."""
import copy

import pytest
import torch

from conftest import build_product_model
from retr_b200 import checkpoint


def _reference_style_checkpoint(tmp_path, kind="CaptionGlobalLoc"):
    model, _, cfg = build_product_model(kind)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    sd["backbone.body.bn1.num_batches_tracked"] = torch.tensor(0)       # torchvision checkpoints carry these
    path = tmp_path / "checkpoint.pth"
    torch.save({"epoch": 3, "model_state_dict": sd, "optimizer_state_dict": {}, "lr_scheduler_state_dict": {},
                "train_loss": 1.0, "val_loss": 1.0, "cider_score": 0.5}, path)
    return model, cfg, str(path)


def test_pack_and_load_round_trip(tmp_path):
    model, cfg, pth = _reference_style_checkpoint(tmp_path)
    out = str(tmp_path / "model.retr")
    n, size = checkpoint.pack_checkpoint(pth, out, cfg)
    assert n == len(model.state_dict())
    header, base = checkpoint.read_header(out)
    assert header["epoch"] == 3 and header["config"]["use_global_features"] is True and base % 64 == 0
    loaded, cfg2 = checkpoint.load_packed(out)
    assert type(loaded).__name__ == "CaptionGlobalLoc" and cfg2.hidden_dim == cfg.hidden_dim
    a, b = model.state_dict(), loaded.state_dict()
    assert list(a.keys()) == list(b.keys())
    assert all(torch.equal(a[k], b[k]) for k in a)


def test_pack_rejects_a_checkpoint_of_another_model(tmp_path):
    _, cfg, pth = _reference_style_checkpoint(tmp_path, "Caption")
    from retr_b200.configuration import Config
    with pytest.raises(KeyError):
        checkpoint.pack_checkpoint(pth, str(tmp_path / "x.retr"), Config(use_global_features=True, use_location_features=True))
    with pytest.raises(ValueError):
        checkpoint.pack_checkpoint(pth, str(tmp_path / "y.retr"), Config(hidden_dim=512))


@pytest.mark.gpu
def test_packed_model_runs_on_the_gpu(tmp_path):
    from retr_b200 import synthetic
    model, cfg, pth = _reference_style_checkpoint(tmp_path)
    out = str(tmp_path / "model.retr")
    checkpoint.pack_checkpoint(pth, out, cfg)
    loaded, _ = checkpoint.load_packed(out)
    data = synthetic.make_batch(2, shard=1, train=True)
    caps, cm = data["caps"][:, :-1].cuda(), data["cap_masks"][:, :-1].cuda()
    outs = []
    for m in (copy.deepcopy(model), loaded):
        m = m.to("cuda").eval()
        with torch.no_grad():
            outs.append(m(*[s.to("cuda") for s in synthetic.pack(data)], caps, cm))
    assert torch.equal(outs[0], outs[1])
