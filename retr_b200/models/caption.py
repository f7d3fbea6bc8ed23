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


class _NativeTrainForward(torch.autograd.Function):
    """Train-mode forward/backward through the native trainer (reference engine.py:70,80:
    ``outputs = model(...)`` / ``loss.backward()``).  The trainable parameters are inputs of the Function
    only so that autograd routes their gradients; the arithmetic reads them by pointer."""

    @staticmethod
    def forward(ctx, model, inputs, p_drop, seed, names, *params):
        trainer = model.trainer(inputs[0].shape[0])
        if hasattr(trainer, "enable_data_parallel"):
            trainer.enable_data_parallel()     # no-op without an NCCL process group of >= 2 ranks
        ctx.trainer, ctx.names = trainer, names               # names[i] is the state_dict key of params[i]
        return trainer.forward(*inputs, dropout_p=p_drop, seed=seed)

    @staticmethod
    def backward(ctx, dlogits):
        trainer = ctx.trainer
        grads = trainer.backward(dlogits)
        if torch.distributed.is_available() and torch.distributed.is_initialized() \
                and torch.distributed.get_world_size() > 1 and not getattr(trainer, "native_dp", False):
            # data parallel without the native NCCL path (e.g. gloo): one all-reduce of the flat gradient buffer
            torch.distributed.all_reduce(trainer.gflat)
            trainer.gflat.div_(torch.distributed.get_world_size())
        return (None, None, None, None, None) + tuple(grads[n] for n in ctx.names)


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
        self._pipeline = []               # extra contexts of greedy_stream(contexts=2), each with its own stream
        self._trainer = None
        self.dropout_p = 0.1              # config.dropout (configuration_template.py:48)
        self.pad_token_id = 0
        self.precision = "bf16"
        self.max_batch = 8                # initial context capacity; grows on demand
        self.backbone_stride = 16
        self.layer_norm_eps = 1e-12
        self.pre_norm = True
        # reference models/caption.py:141 un-masks random positions of an all-masked context row (models/utils.py:60-89)
        # in BOTH modes; finding such a row costs one host sync per forward, so it can be switched off for inputs
        # known to be sane (real and synthetic RefCOCO samples never have an all-masked context)
        self.check_all_masked_context = True
        self.max_train_batch = 0          # capacity of the first native trainer (grows only before the first step)

    # ---- native runtime ------------------------------------------------------------------
    def engine(self, batch=1):
        """The native context bound to this model (created on first use, rebuilt when the
        parameters moved/changed or a larger batch arrives)."""
        from ..runtime import Engine
        device = next(self.parameters()).device
        if device.type != "cuda":
            raise RuntimeError(
                "retr_b200 runs only on a CUDA device (B200, sm_100a): model parameters are on "
                f"{device}. Call model.to('cuda') first; there is no CPU path.")
        eng = self._engine
        stale = (eng is None or eng.device != device or eng.precision != self.precision
                 or eng.max_batch < batch or eng.weights_version != Engine.version_of(self))
        if stale:
            max_batch = max(batch, self.max_batch, eng.max_batch if eng is not None else 0)
            if eng is not None:
                eng.close()
            self._engine = Engine(self, device, self.precision, max_batch=max_batch)
        return self._engine

    def pipeline_engines(self, batch, n=2):
        """``n`` native contexts of this model, each with its own stream, for keeping ``n`` batches in flight
        (eval_utils.decode.greedy_stream): the encode phase of one overlaps the decode loop of another.  Weights are
        uploaded per context (0.5 GB each)."""
        from ..runtime import Engine
        device = next(self.parameters()).device
        if device.type != "cuda":
            raise RuntimeError("retr_b200 runs only on a CUDA device (B200, sm_100a); there is no CPU path.")
        keep = []
        for eng in self._pipeline:
            if (eng.device == device and eng.precision == self.precision and eng.max_batch >= batch
                    and eng.weights_version == Engine.version_of(self)):
                keep.append(eng)
            else:
                eng.close()
        self._pipeline = keep
        while len(self._pipeline) < n:
            eng = Engine(self, device, self.precision, max_batch=max(batch, self.max_batch))
            eng.stream = torch.cuda.Stream(device=device)
            self._pipeline.append(eng)
        return self._pipeline[:n]

    def trainer(self, batch=1):
        """The native trainer bound (by pointer) to this model's parameters; rebuilt when a parameter
        tensor was replaced, a FrozenBatchNorm2d buffer was rewritten, or a larger batch arrives."""
        from ..runtime import Trainer
        device = next(self.parameters()).device
        if device.type != "cuda":
            raise RuntimeError(
                "retr_b200 runs only on a CUDA device (B200, sm_100a): model parameters are on "
                f"{device}. Call model.to('cuda') first; there is no CPU path.")
        tr = self._trainer
        stale = (tr is None or tr.device != device or tr.precision != self.precision or tr.max_batch < batch
                 or tr.signature != Trainer.signature_of(self))
        if stale:
            max_batch = max(batch, self.max_train_batch, tr.max_batch if tr is not None else 0)
            if tr is not None:
                if tr.step_count > 0 and tr.signature == Trainer.signature_of(self) and tr.device == device \
                        and tr.precision == self.precision:
                    # only the batch grew: the AdamW moments and step count live in the old native arena
                    raise RuntimeError(
                        f"batch {batch} exceeds the native trainer's capacity {tr.max_batch} after {tr.step_count} "
                        "optimiser steps; rebuilding would reset the AdamW state. Set model.max_train_batch to the "
                        "largest batch before the first training step.")
                tr.close()
            self._trainer = Trainer(self, device, self.precision, max_batch=max_batch,
                                    pad_token_id=self.pad_token_id)
        return self._trainer

    def _apply(self, fn, *a, **kw):
        out = super()._apply(fn, *a, **kw)
        self.__dict__.pop("_version_tensors", None)
        if self._engine is not None:       # parameters moved: pointers held by the context are stale
            self._engine.close()
            self._engine = None
        for eng in self._pipeline:
            eng.close()
        self._pipeline = []
        if self._trainer is not None:
            self._trainer.close()
            self._trainer = None
        return out

    def _train_inputs(self, t_samples, g_samples, loc, target_exp, target_exp_mask):
        t_samples = self._nested(t_samples)
        g_img = g_mask = None
        if g_samples is not None:
            g_samples = self._nested(g_samples)
            g_img, g_mask = g_samples.tensors, self._context_mask(g_samples.mask)
        return (t_samples.tensors, t_samples.mask, g_img, g_mask, loc, target_exp, target_exp_mask)

    def _context_mask(self, g_mask):
        """models/caption.py:141 ``ensure_unmasked_values`` on the 14x14 context mask, expressed on the full-size
        mask the native side subsamples (models/backbone.py:75 picks pixel (16i,16j))."""
        if not self.check_all_masked_context:
            return g_mask
        st = self.backbone_stride
        m14 = g_mask[:, ::st, ::st]
        if bool(m14.reshape(m14.shape[0], -1).all(dim=1).any()):
            m14 = utils.ensure_unmasked_values(m14.clone())
            g_mask = m14.repeat_interleave(st, 1).repeat_interleave(st, 2)
        return g_mask

    @staticmethod
    def _nested(samples):
        if not isinstance(samples, NestedTensor):
            samples = nested_tensor_from_tensor_list(samples)
        assert samples.mask is not None
        return samples

    def _encode(self, t_samples, g_samples, loc, att_enc=None):
        t_samples = self._nested(t_samples)
        B = t_samples.tensors.shape[0]
        eng = self.engine(B)
        g_img = g_mask = None
        if g_samples is not None:
            g_samples = self._nested(g_samples)
            g_img, g_mask = g_samples.tensors, self._context_mask(g_samples.mask)
        eng.encode(t_samples.tensors, t_samples.mask, g_img, g_mask, loc, att_enc)
        return eng

    def _run(self, t_samples, g_samples, loc, target_exp, target_exp_mask, return_attention):
        if self.training:
            # train mode: dropout active, activations kept, autograd-capable logits (engine.py:70)
            if return_attention:
                raise NotImplementedError("return_attention=True is served by the eval path (model.eval())")
            inputs = self._train_inputs(t_samples, g_samples, loc, target_exp, target_exp_mask)
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())       # drawn from torch's CPU generator
            named = [(n, p) for n, p in self.named_parameters() if p.requires_grad]
            return _NativeTrainForward.apply(self, inputs, float(self.dropout_p), seed, [n for n, _ in named],
                                             *[p for _, p in named])
        B = target_exp.shape[0]
        att_enc = None
        if return_attention:
            eng = self.engine(B)
            att_enc = torch.empty(eng.n_enc, B, eng.S, eng.S, dtype=torch.float32, device=eng.device)
        eng = self._encode(t_samples, g_samples, loc, att_enc)
        logits, att_s, att_c = eng.teacher_forced(target_exp, target_exp_mask, return_attention)
        if return_attention:
            return logits, {"enc_tc_self_att": att_enc, "dec_exp_self_att": att_s,
                            "dec_exp_tc_cross_att": att_c}
        return logits

    # Device-resident greedy loop (used by retr_b200.eval_utils.decode.greedy)
    def greedy_native(self, samples, max_len, bos_token, eos_token):
        t, g, loc = self._split_samples(samples)
        eng = self._encode(t, g, loc)
        return eng.decode_greedy(bos_token, eos_token, max_len)

    def _split_samples(self, samples):
        samples = tuple(samples)
        if self.n_images == 2:
            return samples[0], samples[1], samples[2]
        if self.loc_in_dim:
            return samples[0], None, samples[1]
        return samples[0], None, None


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
    model.dropout_p = float(config.dropout)
    model.pad_token_id = int(config.pad_token_id)
    model.layer_norm_eps = config.layer_norm_eps
    model.pre_norm = bool(config.pre_norm)
    print(f"Built {model.__class__.__name__} model with {transformer.__class__.__name__}")
    return model, torch.nn.CrossEntropyLoss()
