"""Host-side owner of one native context (``retr_ctx``) bound to a caption model.

Builds the ``retr_model_desc`` from the module tree, streams the model's ``state_dict`` into
the library under the reference's key names, and exposes encode / teacher-forced forward /
greedy decode.  Tensors are only ever passed down as raw device pointers.
"""
import ctypes as C
import weakref

import torch

from . import native
from .models.backbone import RESNET_SPECS

_PRECISION = {"fp32": native.FP32, "bf16": native.BF16}


def _as(t, dtype, device):
    if t.dtype == torch.bool and dtype == torch.uint8:
        t = t.contiguous().view(torch.uint8)        # same bytes: no cast kernel for the padding masks
    return t.to(device=device, dtype=dtype).contiguous()


def model_desc(model, device, precision, max_batch, image_size=224):
    """``retr_model_desc`` of a caption model: the attributes build_model reads from Config."""
    tr = model.transformer
    d = native.ModelDesc()
    d.device = device.index if device.index is not None else torch.cuda.current_device()
    d.precision = _PRECISION[precision]
    d.blocks = (C.c_int32 * 4)(*RESNET_SPECS[model.backbone.arch][0])
    d.dilation = int(model.backbone.dilation)
    d.n_images = model.n_images
    d.loc_in_dim = model.loc_in_dim
    d.n_loc = {0: 0, 1: getattr(model, "n_loc_feats", 5), 7: 7}[model.loc_in_dim]
    d.image_size = image_size
    d.d_model, d.n_heads, d.d_ff = tr.d_model, tr.nhead, tr.dim_feedforward
    d.enc_layers, d.dec_layers = tr.encoder.num_layers, tr.decoder.num_layers
    d.max_pos = tr.embeddings.position_embeddings.num_embeddings
    d.vocab = model.mlp.layers[-1].out_features
    d.pre_norm = int(tr.encoder.norm is not None)
    d.embed_ln_eps = float(tr.embeddings.LayerNorm.eps)
    d.max_batch = int(max_batch)
    return d


def check_loc(desc, loc, B):
    """The native side reads ``B * n_loc`` contiguous floats; a differently shaped tensor would be mis-read
    silently where the reference's ``loc_proj`` Linear raises a shape error (models/caption.py:60,108,126-128)."""
    if desc.loc_in_dim == 0:
        return
    if loc is None:
        raise ValueError("this model class takes location features")
    if tuple(loc.shape) != (B, desc.n_loc):
        raise RuntimeError(f"location features must be [{B},{desc.n_loc}] for this model, got {tuple(loc.shape)} "
                           "(CaptionGlobalLoc: set model.n_loc_feats before the first call to change the count)")


class Engine:
    def __init__(self, model, device, precision="bf16", max_batch=8, image_size=224):
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("retr_b200 has no CPU path: the model must live on a CUDA device")
        if precision not in _PRECISION:
            raise ValueError(f"precision must be one of {sorted(_PRECISION)}")
        self.lib = native.lib()
        self.device, self.precision, self.max_batch, self.image_size = device, precision, int(max_batch), image_size
        d = model_desc(model, device, precision, self.max_batch, image_size)
        self.desc = d
        self.T, self.V, self.d_model = d.max_pos, d.vocab, d.d_model
        self.feat_channels = model.backbone.num_channels
        self.n_enc, self.n_dec = d.enc_layers, d.dec_layers
        handle = C.c_void_p()
        with torch.cuda.device(device):
            native.check(self.lib.retr_create(C.byref(d), C.byref(handle)))
            self.handle = handle
            self._load(model)
        self.S = self.lib.retr_encoder_tokens(self.handle)
        self.weights_version = self.version_of(model)
        self.batch = 0
        self.stream = None

    # ---- weights --------------------------------------------------------------------------
    @staticmethod
    def version_of(model):
        """Changes when any parameter/buffer was written through torch (``_version``) or by the native optimiser
        (``model._weights_epoch``, bumped by Trainer.clip_adamw: it updates the masters through raw pointers, which
        torch's version counters never see).  The tensor list is cached on the model (rebuilt when it moves)."""
        tensors = model.__dict__.get("_version_tensors")
        if tensors is None:
            tensors = list(model.state_dict(keep_vars=True).values())
            model.__dict__["_version_tensors"] = tensors
        return sum(t._version for t in tensors), model.__dict__.get("_weights_epoch", 0)

    def _load(self, model):
        for name, t in model.state_dict().items():
            if not torch.is_floating_point(t):
                continue
            t32 = _as(t.detach(), torch.float32, self.device)
            shape = (C.c_int64 * max(t32.dim(), 1))(*(t32.shape if t32.dim() else (1,)))
            native.check(self.lib.retr_load_weight(self.handle, name.encode(), native.ptr(t32), shape,
                                                   max(t32.dim(), 1), 1))
        native.check(self.lib.retr_finalize_weights(self.handle))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.retr_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- hot path ---------------------------------------------------------------------------
    def _stream(self):
        # a context used next to another one (greedy_stream with contexts=2) owns a stream; otherwise torch's current
        if self.stream is not None:
            return self.stream.cuda_stream
        return native.stream_ptr(self.device)

    def encode(self, t_img, t_mask, g_img=None, g_mask=None, loc=None, att_enc=None):
        B = t_img.shape[0]
        if B > self.max_batch:
            raise RuntimeError(f"batch {B} exceeds the context's max_batch {self.max_batch}")
        if tuple(t_img.shape[1:]) != (3, self.image_size, self.image_size):
            raise ValueError(f"images must be [B,3,{self.image_size},{self.image_size}], got {tuple(t_img.shape)}")
        check_loc(self.desc, loc, B)
        keep = [_as(t_img, torch.float32, self.device), _as(t_mask, torch.uint8, self.device)]
        keep += [None if g_img is None else _as(g_img, torch.float32, self.device),
                 None if g_mask is None else _as(g_mask, torch.uint8, self.device),
                 None if loc is None else _as(loc, torch.float32, self.device)]
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_set_encoder_attention_capture(self.handle, native.ptr(att_enc)))
            native.check(self.lib.retr_encode(self.handle, B, *[native.ptr(k) for k in keep], self._stream()))
            native.check(self.lib.retr_set_encoder_attention_capture(self.handle, None))
        self._keep = keep          # inputs must outlive the asynchronous kernels
        self.batch = B

    def teacher_forced(self, ids, id_mask, return_attention=False):
        B = self.batch
        if tuple(ids.shape) != (B, self.T):
            raise RuntimeError(f"The size of tensor a ({ids.shape[1]}) must match the size of tensor b ({self.T}) "
                               "(caption length must equal max_position_embeddings)")
        ids = _as(ids, torch.int64, self.device)
        id_mask = _as(id_mask, torch.uint8, self.device)
        logits = torch.empty(B, self.T, self.V, dtype=torch.float32, device=self.device)
        att_s = att_c = None
        if return_attention:
            att_s = torch.empty(self.n_dec, B, self.T, self.T, dtype=torch.float32, device=self.device)
            att_c = torch.empty(self.n_dec, B, self.T, self.S, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_forward_teacher_forced(
                self.handle, native.ptr(ids), native.ptr(id_mask), native.ptr(logits), native.ptr(att_s),
                native.ptr(att_c), self._stream()))
        self._keep2 = (ids, id_mask)
        return logits, att_s, att_c

    def decode_greedy(self, bos, eos, max_len):
        out = torch.empty(self.batch, max_len, dtype=torch.int64, device=self.device)
        steps = C.c_int32(0)
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_decode_greedy(self.handle, bos, eos, max_len, native.ptr(out),
                                                     C.byref(steps), self._stream()))
        return out, steps.value

    def generate_host(self, t_img, t_mask, g_img, g_mask, loc, bos, eos, max_len, out=None):
        """Host (pinned) buffers in, host ids out; H2D + encode + decode + D2H in one native call."""
        B = t_img.shape[0]
        for t in (t_img, t_mask, g_img, g_mask, loc):
            if t is not None and (t.device.type != "cpu" or not t.is_contiguous()):
                raise ValueError("generate_host takes contiguous host tensors")
        check_loc(self.desc, loc, B)
        if out is None:
            out = torch.empty(B, max_len, dtype=torch.int64).pin_memory()
        steps = C.c_int32(0)
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_generate_host(
                self.handle, B, native.ptr(t_img), native.ptr(t_mask), native.ptr(g_img), native.ptr(g_mask),
                native.ptr(loc), bos, eos, max_len, native.ptr(out), C.byref(steps), self._stream()))
        self.batch = B
        return out, steps.value

    def encode_host(self, t_img, t_mask, g_img, g_mask, loc):
        """Asynchronous first half of generate_host (copies + backbone + encoder queued on this context's stream)."""
        B = t_img.shape[0]
        for t in (t_img, t_mask, g_img, g_mask, loc):
            if t is not None and (t.device.type != "cpu" or not t.is_contiguous()):
                raise ValueError("encode_host takes contiguous host tensors")
        check_loc(self.desc, loc, B)
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_encode_host(
                self.handle, B, native.ptr(t_img), native.ptr(t_mask), native.ptr(g_img), native.ptr(g_mask),
                native.ptr(loc), self._stream()))
        self._keep_eh = (t_img, t_mask, g_img, g_mask, loc)
        self.batch = B

    def decode_host(self, bos, eos, max_len, out=None):
        """Second half: greedy loop of the batch last encoded here; returns with the ids in `out` (host)."""
        if out is None:
            out = torch.empty(self.batch, max_len, dtype=torch.int64).pin_memory()
        steps = C.c_int32(0)
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_decode_host(self.handle, bos, eos, max_len, native.ptr(out), C.byref(steps),
                                                   self._stream()))
        return out, steps.value

    def set_encode_cta_cap(self, ctas):
        native.check(self.lib.retr_set_encode_cta_cap(self.handle, int(ctas)))

    def prefetch_host(self, t_img, t_mask, g_img, g_mask, loc):
        """Queue the H2D copies of the NEXT batch (pinned host tensors) while the current one decodes; the
        following generate_host call with these same tensors consumes the staged copy."""
        B = t_img.shape[0]
        for t in (t_img, t_mask, g_img, g_mask, loc):
            if t is not None and (t.device.type != "cpu" or not t.is_contiguous()):
                raise ValueError("prefetch_host takes contiguous host tensors")
        check_loc(self.desc, loc, B)
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_prefetch_host(
                self.handle, B, native.ptr(t_img), native.ptr(t_mask), native.ptr(g_img), native.ptr(g_mask),
                native.ptr(loc)))
        self._keep_pf = (t_img, t_mask, g_img, g_mask, loc)

    def read_state(self, what):
        B, S, d = self.batch, self.S, self.d_model
        if what == "mask":
            out = torch.empty(B, S, dtype=torch.uint8, device=self.device)
        elif what in ("memory", "src"):
            out = torch.empty(B, S, d, dtype=torch.float32, device=self.device)
        else:
            g = int(round(((self.S - (self.desc.n_loc if self.desc.loc_in_dim == 1 else min(self.desc.loc_in_dim, 1)))
                           / self.desc.n_images) ** 0.5))
            out = torch.empty(B, g, g, self.feat_channels, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_read_state(self.handle, what.encode(), native.ptr(out),
                                                  out.numel() * out.element_size(), self._stream()))
        return out

    def kernel_launches(self):
        return int(self.lib.retr_kernel_launches(self.handle))

    def phase_ms(self):
        buf = (C.c_float * 3)()
        native.check(self.lib.retr_last_phase_ms(self.handle, buf))
        return {"backbone": buf[0], "encoder": buf[1], "decode": buf[2]}


class Trainer:
    """Host-side owner of one native ``retr_trainer``: the training step of the reference
    (engine.py:52-87).  The model's parameters are bound BY POINTER (the native side reads the fp32
    masters at every step, so an optimiser updating them in place needs no re-upload); gradients
    land in one flat fp32 buffer of which ``self.grads[name]`` are views."""

    def __init__(self, model, device, precision="bf16", max_batch=8, image_size=224, pad_token_id=0):
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("retr_b200 has no CPU path: the model must live on a CUDA device")
        self.lib = native.lib()
        self.device, self.precision, self.max_batch, self.image_size = device, precision, int(max_batch), image_size
        d = model_desc(model, device, precision, self.max_batch, image_size)
        self.desc, self.T, self.V = d, d.max_pos, d.vocab
        sd = model.state_dict(keep_vars=True)
        self.names = [n for n, t in sd.items() if torch.is_floating_point(t)]
        self.tensors = {n: sd[n] for n in self.names}
        for n, t in self.tensors.items():
            if t.dtype != torch.float32 or not t.is_contiguous() or t.device != device:
                raise RuntimeError(f"parameter {n} must be a contiguous fp32 tensor on {device}")
        trainable = [n for n in self.names if isinstance(sd[n], torch.nn.Parameter) and sd[n].requires_grad]
        # flat arenas: backbone group first (its own learning rate, main.py:31-36), then everything else.  The
        # parameters are RE-POINTED into the arena (p.data = view), so the model, any torch optimiser and the
        # native kernels all see the same storage.
        self.trainable = [n for n in trainable if "backbone" in n] + [n for n in trainable if "backbone" not in n]
        total = sum(self.tensors[n].numel() for n in self.trainable)
        self.n_backbone = sum(self.tensors[n].numel() for n in self.trainable if "backbone" in n)
        self.pflat = torch.empty(total, dtype=torch.float32, device=device)
        self.gflat = torch.zeros(total, dtype=torch.float32, device=device)
        self.grads, off = {}, 0
        with torch.no_grad():
            for n in self.trainable:
                t = self.tensors[n]
                k = t.numel()
                view = self.pflat[off:off + k].view(t.shape)
                view.copy_(t.data)
                t.data = view
                self.grads[n] = self.gflat[off:off + k].view(t.shape)
                off += k
        handle = C.c_void_p()
        with torch.cuda.device(device):
            native.check(self.lib.retr_train_create(C.byref(d), int(pad_token_id), C.byref(handle)))
            self.handle = handle
            for n in self.names:
                t = self.tensors[n]
                shape = (C.c_int64 * max(t.dim(), 1))(*(t.shape if t.dim() else (1,)))
                native.check(self.lib.retr_train_bind(self.handle, n.encode(), native.ptr(t.data), native.ptr(self.grads.get(n)),
                                                      shape, max(t.dim(), 1)))
            native.check(self.lib.retr_train_set_arena(self.handle, native.ptr(self.pflat), native.ptr(self.gflat), total,
                                                       self.n_backbone))
            native.check(self.lib.retr_train_finalize(self.handle))
        self.signature = self.signature_of(model)
        self.batch, self.step_count = 0, 0
        self._model = weakref.ref(model)
        self._loss = torch.zeros(1, dtype=torch.float32, device=device)
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=device)

    # ---- data parallel ---------------------------------------------------------------------------------------
    native_dp = False

    def enable_data_parallel(self):
        """Give the native trainer its own NCCL communicator over torch.distributed's ranks: from then on the backward
        pass all-reduces (mean) the gradient buckets itself, overlapped with the rest of the pass.  No-op without an
        initialised NCCL process group or with one rank."""
        import torch.distributed as dist
        if self.native_dp or not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() < 2 \
                or dist.get_backend() != "nccl":
            return self.native_dp
        world, rank = dist.get_world_size(), dist.get_rank()
        ident = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            buf = (C.c_char * 128)()
            native.check(self.lib.retr_nccl_unique_id(buf))
            ident = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        ident = ident.to(self.device)
        dist.broadcast(ident, 0)
        raw = bytes(ident.cpu().numpy().tobytes())
        comm = C.c_void_p()
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_nccl_comm_create(raw, world, rank, C.byref(comm)))
            native.check(self.lib.retr_train_set_comm(self.handle, comm, world))
        self._comm, self.native_dp = comm, True
        return True

    def grad_buckets(self):
        buf = (C.c_int64 * 32)()
        n = self.lib.retr_train_grad_buckets(self.handle, buf, 16)
        return [(buf[2 * i], buf[2 * i + 1]) for i in range(max(n, 0))]

    def allreduce_info(self):
        b = self.grad_buckets()
        return {"native_nccl": bool(self.native_dp), "overlapped_with_backward": bool(self.native_dp),
                "buckets": len(b), "bucket_mb": [round((hi - lo) * 4 / 2 ** 20, 1) for lo, hi in b],
                "order": "reverse execution: head, decoder+embeddings, encoder+projections, layer4, layer3, layer2"}

    @staticmethod
    def signature_of(model):
        """Changes when a bound pointer moved or a folded buffer (FrozenBatchNorm2d) was rewritten."""
        ptrs = tuple(t.data_ptr() for t in model.state_dict(keep_vars=True).values())
        return ptrs, sum(b._version for b in model.buffers())

    def close(self):
        if getattr(self, "handle", None):
            if getattr(self, "_comm", None):
                self.lib.retr_train_set_comm(self.handle, None, 1)
                self.lib.retr_nccl_comm_destroy(self._comm)
                self._comm, self.native_dp = None, False
            self.lib.retr_train_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _inputs(self, t_img, t_mask, g_img, g_mask, loc, ids, id_mask):
        B = t_img.shape[0]
        if B > self.max_batch:
            raise RuntimeError(f"batch {B} exceeds the trainer's max_batch {self.max_batch}")
        if tuple(ids.shape) != (B, self.T):
            raise RuntimeError(f"The size of tensor a ({ids.shape[1]}) must match the size of tensor b ({self.T}) "
                               "(caption length must equal max_position_embeddings)")
        check_loc(self.desc, loc, B)
        keep = [_as(t_img, torch.float32, self.device), _as(t_mask, torch.uint8, self.device),
                None if g_img is None else _as(g_img, torch.float32, self.device),
                None if g_mask is None else _as(g_mask, torch.uint8, self.device),
                None if loc is None else _as(loc, torch.float32, self.device),
                _as(ids, torch.int64, self.device), _as(id_mask, torch.uint8, self.device)]
        self._keep = keep
        self.batch = B
        return B, keep

    def forward(self, t_img, t_mask, g_img, g_mask, loc, ids, id_mask, dropout_p=0.0, seed=0):
        B, keep = self._inputs(t_img, t_mask, g_img, g_mask, loc, ids, id_mask)
        logits = torch.empty(B, self.T, self.V, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_train_forward(self.handle, B, *[native.ptr(k) for k in keep], native.ptr(logits),
                                                     float(dropout_p), int(seed), native.stream_ptr(self.device)))
        return logits

    def backward(self, dlogits):
        dlogits = _as(dlogits, torch.float32, self.device)
        if tuple(dlogits.shape) != (self.batch, self.T, self.V):
            raise RuntimeError(f"dlogits must be [{self.batch},{self.T},{self.V}], got {tuple(dlogits.shape)}")
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_train_backward(self.handle, native.ptr(dlogits), native.stream_ptr(self.device)))
        self._keep_d = dlogits
        return self.grads

    def loss_and_grads(self, t_img, t_mask, g_img, g_mask, loc, ids, id_mask, targets, dropout_p=0.0, seed=0):
        """Fused forward + mean CE + backward; returns the mean loss as a 0-d device tensor (no sync)."""
        B, keep = self._inputs(t_img, t_mask, g_img, g_mask, loc, ids, id_mask)
        targets = _as(targets, torch.int64, self.device)
        self._keep = keep + [targets]
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_train_loss_and_grads(
                self.handle, B, *[native.ptr(k) for k in keep], native.ptr(targets), float(dropout_p), int(seed),
                native.ptr(self._loss), native.stream_ptr(self.device)))
        return self._loss[0] / (B * self.T)

    def clip_adamw(self, max_norm, lr, lr_backbone, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-4):
        """clip_grad_norm_ + AdamW in native kernels; returns the pre-clip gradient norm (0-d device tensor)."""
        self.step_count += 1
        model = self._model()
        if model is not None:       # the masters change under torch's feet: inference contexts must reload
            model.__dict__["_weights_epoch"] = model.__dict__.get("_weights_epoch", 0) + 1
        with torch.cuda.device(self.device):
            native.check(self.lib.retr_train_clip_adamw(self.handle, float(max_norm), float(lr), float(lr_backbone),
                                                        float(betas[0]), float(betas[1]), float(eps), float(weight_decay),
                                                        self.step_count, native.ptr(self._sumsq),
                                                        native.stream_ptr(self.device)))
        return self._sumsq[0].sqrt()

    def kernel_launches(self):
        return int(self.lib.retr_train_kernel_launches(self.handle))
