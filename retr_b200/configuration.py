"""Hyper-parameter bag read by ``build_model``.

Mirrors the attribute names of the reference's ``configuration_template.py:4-56`` (the
reference scripts import a user copy of it as ``configuration.Config``).  Only attributes
that the hot path or its callers read are listed; unknown attributes can be added freely
on the instance, exactly as with the reference's plain attribute bag.
"""
from os.path import join


class Config:
    # (name, default) pairs; values equal the reference template so a Config() from either
    # side builds the same model.
    _DEFAULTS = (
        ("prefix", "refcoco"),
        # optimiser recipe (main.py:30-40)
        ("lr_backbone", 1e-5), ("lr", 1e-4), ("epochs", 30), ("lr_drop", 20),
        ("start_epoch", 0), ("weight_decay", 1e-4),
        # backbone
        ("backbone", "ResNet101"), ("position_embedding", "sine"), ("dilation", True),
        # basic
        ("device", "cuda"), ("seed", 42), ("batch_size", 32), ("num_workers", 8),
        ("clip_max_norm", 0.1), ("early_stopping", True),
        ("use_global_features", False), ("use_location_features", False), ("verbose", True),
        # transformer
        ("transformer_type", "Concat"), ("hidden_dim", 256), ("pad_token_id", 0),
        ("max_position_embeddings", 128), ("layer_norm_eps", 1e-12), ("dropout", 0.1),
        ("vocab_size", 30522), ("enc_layers", 6), ("dec_layers", 6),
        ("dim_feedforward", 2048), ("nheads", 8), ("pre_norm", True),
        # dataset (unused by the hot path)
        ("dir", "PATH_TO_COCO"), ("ref_base", "PATH_TO_REF_BASE"), ("limit", -1),
    )

    def __init__(self, **overrides):
        for k, v in self._DEFAULTS:
            setattr(self, k, v)
        self.checkpoint = f"./{self.prefix}_checkpoint.pth"
        self.project_data_path = "./data"
        self.checkpoint_path = join(self.project_data_path, "models", self.prefix)
        self.ref_dir = join(self.ref_base, self.prefix)
        # B200 runtime knobs (no reference analogue)
        self.precision = "bf16"      # "bf16" (tcgen05 path) or "fp32" (strict-parity path)
        for k, v in overrides.items():
            setattr(self, k, v)


def headline_config(**overrides):
    """CaptionGlobalLoc configuration used by BASELINE.json (both feature flags on)."""
    cfg = Config(use_global_features=True, use_location_features=True)
    for k, v in overrides.items():
        setattr(cfg, k, v)
    return cfg
