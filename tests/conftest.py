import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


KIND_FLAGS = {"CaptionGlobalLoc": (True, True), "Caption": (False, False), "CaptionLoc": (False, True)}

_model_cache = {}


def build_product_model(kind="CaptionGlobalLoc", seed=42, **cfg_over):
    """Product model with the reference's seed-42 init (CPU tensors)."""
    from retr_b200.configuration import Config
    from retr_b200.models import caption
    key = (kind, seed, tuple(sorted(cfg_over.items())))
    if key not in _model_cache:
        g, l = KIND_FLAGS[kind]
        cfg = Config(use_global_features=g, use_location_features=l, **cfg_over)
        torch.manual_seed(seed)
        model, crit = caption.build_model(cfg)
        _model_cache[key] = (model, crit, cfg)
    return _model_cache[key]


def golden(kind):
    return torch.load(os.path.join(GOLDEN, f"{kind}.pt"), weights_only=False)


def synthetic_samples(kind, batch, shard=0, train=True):
    from retr_b200 import synthetic
    data = synthetic.make_batch(batch, shard=shard, train=train)
    loc = data["loc"]
    if kind == "CaptionLoc":
        loc = torch.cat([loc, loc[:, :2] * 0.5 + 0.25], 1)
    if kind == "Caption":
        loc = None
    g_img = data["g_img"] if kind == "CaptionGlobalLoc" else None
    g_mask = data["g_mask"] if kind == "CaptionGlobalLoc" else None
    return data, (data["t_img"], data["t_mask"], g_img, g_mask, loc)
