"""``build_model`` and the three caption models (reference models/caption.py).

Same public surface as the reference: ``build_model(config) -> (model, CrossEntropyLoss())``;
``model(*samples, target_exp, target_exp_mask, return_attention=False) -> logits [B,T,V]``
(or ``(logits, att_dict)``); state_dict keys/shapes identical.

What differs is where the arithmetic runs: ``forward`` hands raw device pointers to the
native sm_100a library through ``retr_b200.runtime.Engine`` (backbone -> token assembly ->
encoder -> decoder -> vocabulary head).  There is no PyTorch/CPU fallback: calling a model
whose inputs are not on a CUDA device, or without the built library, raises.

Token layout per class (reference models/caption.py:23-47, 64-95, 112-158):
  Caption           target 14x14 tokens                                        (S = 196)
  CaptionLoc        target tokens | Linear(7,d)(loc) as ONE token               (S = 197)
  CaptionGlobalLoc  target tokens | Linear(1,d)(loc_k) per scalar | context     (S = 196+n+196)
"""
import torch
from torch import nn

from . import utils
from .backbone import build_backbone
from .transformer import build_transformer
from .utils import NestedTensor, nested_tensor_from_tensor_list


class MLP(nn.Module):
    """Linear/ReLU stack; the head is 256 -> 512 -> 512 -> vocab (reference models/caption.py:161-174)."""

    def __init__(self, input_dim, hidden_dim, output_dim, num_layers):
        super().__init__()
        self.num_layers = num_layers
        dims = [input_dim] + [hidden_dim] * (num_layers - 1) + [output_dim]
        self.layers = nn.ModuleList(nn.Linear(a, b) for a, b in zip(dims[:-1], dims[1:]))


class _CaptionBase(nn.Module):
    n_images = 1          # backbone passes per sample
    loc_in_dim = 0        # 0: no location input

    def __init__(self, backbone, transformer, positional_encoding, hidden_dim, vocab_size):
        super().__init__()
        self.backbone = backbone
        self.positional_encoding = positional_encoding
        self.input_proj = nn.Conv2d(backbone.num_channels, hidden_dim, kernel_size=1)
        if self.loc_in_dim:
            self.loc_proj = nn.Linear(self.loc_in_dim, hidden_dim)
        self.transformer = transformer
        self.mlp = MLP(hidden_dim, 512, vocab_size, 3)
        self._engine = None
        self.precision = "bf16"
        self.layer_norm_eps = 1e-12
        self.pre_norm = True

    # ---- native runtime ------------------------------------------------------------------
    def engine(self, device=None):
        """The native context bound to this model (created on first use)."""
        from ..runtime import Engine
        if device is None:
            device = next(self.parameters()).device
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError(
                "retr_b200 runs only on a CUDA device (B200, sm_100a): model parameters are on "
                f"{device}. Call model.to('cuda') first; there is no CPU path.")
        if self._engine is None or self._engine.device != device or \
                self._engine.precision != self.precision:
            self._engine = Engine(self, device, self.precision)
        return self._engine

    def _apply(self, fn, *a, **kw):
        out = super()._apply(fn, *a, **kw)
        self._engine = None            # parameters moved: pointers held by the context are stale
        return out

    @staticmethod
    def _nested(samples):
        if not isinstance(samples, NestedTensor):
            samples = nested_tensor_from_tensor_list(samples)
        assert samples.mask is not None
        return samples

    def _run(self, t_samples, g_samples, loc, target_exp, target_exp_mask, return_attention):
        eng = self.engine()
        t_samples = self._nested(t_samples)
        if g_samples is not None:
            g_samples = self._nested(g_samples)
            # all-masked context rows get 1% random positions un-masked (models/utils.py:60-89);
            # the check needs the 14x14 mask, which the engine returns lazily
            g_samples = NestedTensor(g_samples.tensors, eng.guard_context_mask(g_samples.mask))
        eng.encode(t_samples, g_samples, loc, need_grad=self.training and torch.is_grad_enabled())
        return eng.teacher_forced(target_exp, target_exp_mask, return_attention,
                                  train=self.training)


class Caption(_CaptionBase):
    def forward(self, samples, target_exp, target_exp_mask, return_attention=False):
        return self._run(samples, None, None, target_exp, target_exp_mask, return_attention)


class CaptionLoc(_CaptionBase):
    loc_in_dim = 7

    def forward(self, t_samples, loc_feats, target_exp, target_exp_mask, return_attention=False):
        return self._run(t_samples, None, loc_feats, target_exp, target_exp_mask, return_attention)


class CaptionGlobalLoc(_CaptionBase):
    n_images = 2
    loc_in_dim = 1

    def forward(self, t_samples, g_samples, loc_feats, target_exp, target_exp_mask,
                return_attention=False):
        return self._run(t_samples, g_samples, loc_feats, target_exp, target_exp_mask,
                         return_attention)


def build_model(config):
    backbone = build_backbone(config)
    transformer = build_transformer(config)
    use_global, use_location = config.use_global_features, config.use_location_features
    print(f"global features: {use_global}, location features: {use_location}")
    if use_global and not use_location:
        raise NotImplementedError()      # no model class for this combination (models/caption.py:198-199)
    Model = {(False, False): Caption, (False, True): CaptionLoc, (True, True): CaptionGlobalLoc}[
        (bool(use_global), bool(use_location))]
    model = Model(backbone, transformer, None, config.hidden_dim, config.vocab_size)
    model.precision = getattr(config, "precision", "bf16")
    model.layer_norm_eps = config.layer_norm_eps
    model.pre_norm = bool(config.pre_norm)
    print(f"Built {model.__class__.__name__} model with {transformer.__class__.__name__}")
    return model, torch.nn.CrossEntropyLoss()
