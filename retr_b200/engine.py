"""Loop shells with the reference's names (engine.py): input packing, the train step loop and
the validation-loss loop.  The model call inside is the native hot path."""
import math
import sys

import torch

from .models.utils import NestedTensor


def pack_encoder_inputs(encoder_input, global_features, location_features, device="cpu"):
    """Dataloader tuple -> the ``samples`` tuple of the model classes (reference engine.py:20-48):
    (t,), (t, g), (t, loc) or (t, g, loc)."""
    encoder_input = list(encoder_input)
    n_expected = 2 + (2 if global_features else 0) + (1 if location_features else 0)
    if len(encoder_input) != n_expected:
        raise ValueError(f"expected {n_expected} encoder inputs, got {len(encoder_input)}")
    out = [NestedTensor(encoder_input[0], encoder_input[1]).to(device)]
    rest = encoder_input[2:]
    if global_features:
        out.append(NestedTensor(rest[0], rest[1]).to(device))
        rest = rest[2:]
    if location_features:
        out.append(rest[0].to(device))
    return tuple(out)


def _flags(data_loader):
    ds = data_loader.dataset
    return ds.return_global_context, ds.return_location_features


def train_one_epoch(model, criterion, data_loader, optimizer, device, epoch, max_norm):
    """Teacher-forced CE step loop — reference engine.py:52-87.  Mean CE over all B*T positions
    (pads included), global-norm clip, optimizer step; exits on a non-finite loss."""
    model.train()
    criterion.train()
    use_global, use_loc = _flags(data_loader)
    total, epoch_loss = len(data_loader), 0.0
    for _ann_ids, *encoder_input, caps, cap_masks in data_loader:
        samples = pack_encoder_inputs(encoder_input, use_global, use_loc, device)
        caps, cap_masks = caps.to(device), cap_masks.to(device)
        outputs = model(*samples, caps[:, :-1], cap_masks[:, :-1])
        loss = criterion(outputs.permute(0, 2, 1), caps[:, 1:])
        loss_value = loss.item()
        epoch_loss += loss_value
        if not math.isfinite(loss_value):
            print(f"Loss is {loss_value}, stopping training")
            sys.exit(1)
        optimizer.zero_grad()
        loss.backward()
        if max_norm > 0:
            torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm)
        optimizer.step()
    return epoch_loss / total


def train_step_native(model, samples, caps, cap_masks, max_norm, lr, lr_backbone, weight_decay=1e-4,
                      betas=(0.9, 0.999), eps=1e-8):
    """One step of the loop body above (engine.py:64-83) entirely in native kernels: train-mode forward,
    fused cross-entropy (bf16 logits never leave the device), backward, NCCL all-reduce of the flat gradient
    buffer when a process group is up, clip_grad_norm_ + AdamW (main.py:30-40 recipe: parameters whose name
    contains "backbone" use lr_backbone).  Returns (mean loss, pre-clip gradient norm) as 0-d device tensors —
    no host synchronisation."""
    import torch.distributed as dist
    t, g, loc = model._split_samples(samples)
    inputs = model._train_inputs(t, g, loc, caps[:, :-1], cap_masks[:, :-1])
    tr = model.trainer(inputs[0].shape[0])
    seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    p = float(model.dropout_p) if model.training else 0.0
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    if multi and hasattr(tr, "enable_data_parallel"):
        tr.enable_data_parallel()          # NCCL group: bucketed all-reduce inside the native backward pass, overlapped
    loss = tr.loss_and_grads(*inputs, caps[:, 1:], dropout_p=p, seed=seed)
    if multi and not getattr(tr, "native_dp", False):
        dist.all_reduce(tr.gflat)          # other backends (gloo in the CPU tests): one bucket after the pass
        tr.gflat.div_(dist.get_world_size())
    norm = tr.clip_adamw(max_norm, lr, lr_backbone, betas=betas, eps=eps, weight_decay=weight_decay)
    return loss, norm


@torch.no_grad()
def evaluate(model, criterion, data_loader, device):
    """Validation loss — reference engine.py:89-114 (same forward, no backward)."""
    model.eval()
    criterion.eval()
    use_global, use_loc = _flags(data_loader)
    total, loss_sum = len(data_loader), 0.0
    for _ann_ids, *encoder_input, caps, cap_masks in data_loader:
        samples = pack_encoder_inputs(encoder_input, use_global, use_loc, device)
        caps, cap_masks = caps.to(device), cap_masks.to(device)
        outputs = model(*samples, caps[:, :-1], cap_masks[:, :-1])
        loss_sum += criterion(outputs.permute(0, 2, 1), caps[:, 1:]).item()
    return loss_sum / total


def normalize_with_tokenizer(sent, tokenizer):
    """reference engine.py:116-121."""
    return tokenizer.decode(tokenizer.encode(sent), skip_special_tokens=True)


def eval_model(model, data_loader, tokenizer, config, metrics_to_omit=(), print_samples=False, compute_metrics=None):
    """Decode a split and score it — the loop shell of reference engine.py:125-187.  The decode is the native
    KV-cached greedy loop (any batch size); the string metrics stay on the host and are DELEGATED: pass
    ``compute_metrics(ref_list=..., hyp_list=...) -> dict`` (e.g. ``nlgeval.NLGEval(...).compute_metrics``, which is
    what the reference constructs at engine.py:135-137).  Without it the metrics dict is empty — nlgeval (java METEOR
    jar) is out of scope and not imported here.
    Returns (metrics_dict, [{'ann_id', 'expression'}, ...]) like the reference."""
    from collections import defaultdict

    from .eval_utils.decode import greedy_decoding
    model.eval()
    annotations = defaultdict(list)
    for a in data_loader.dataset.annot:
        annotations[a[0]].append(a[2])
    hypotheses, ids_hypotheses, references = [], [], []
    pad_id = tokenizer.convert_tokens_to_ids(tokenizer.pad_token)
    bos_id = tokenizer.convert_tokens_to_ids(tokenizer.cls_token)
    eos_id = tokenizer.convert_tokens_to_ids(tokenizer.sep_token)
    use_global, use_loc = _flags(data_loader)
    from collections import deque

    from .eval_utils.decode import greedy_stream, idx2sents, prune_cap_ids
    from .models.caption import _CaptionBase
    max_len = config.max_position_embeddings
    pending_ann = deque()

    def packed():
        for ann_ids, *encoder_input, _caps, _cap_masks in data_loader:
            pending_ann.append(ann_ids)
            yield pack_encoder_inputs(encoder_input, use_global, use_loc)

    def decoded():
        if isinstance(model, _CaptionBase) and next(model.parameters()).device.type == "cuda":
            # native model: the next batch's images cross PCIe while this one is decoded (retr_prefetch_host)
            for ids in greedy_stream(packed(), model, max_len=max_len, bos_token=bos_id, eos_token=eos_id):
                yield idx2sents(prune_cap_ids(ids.numpy().tolist(), clean=True, pad_token=pad_id, bos_token=bos_id,
                                              eos_token=eos_id), tokenizer)
        else:
            for samples in packed():
                yield greedy_decoding(samples, model, tokenizer, max_len=max_len, clean=True, pad_token=pad_id,
                                      bos_token=bos_id, eos_token=eos_id, device="auto")

    with torch.no_grad():
        for hyps in decoded():
            ann_ids = pending_ann.popleft()
            hypotheses += hyps
            ann_list = ann_ids.tolist() if hasattr(ann_ids, "tolist") else list(ann_ids)
            ids_hyps = [{"ann_id": i, "expression": h} for i, h in zip(ann_list, hyps)]
            ids_hypotheses += ids_hyps
            if print_samples:
                print(*ids_hyps, sep="\n")
            references += [[normalize_with_tokenizer(r, tokenizer) for r in annotations[i]] for i in ann_list]
    metrics = {}
    if compute_metrics is not None:
        transposed = list(map(list, zip(*references)))
        metrics = compute_metrics(ref_list=transposed, hyp_list=hypotheses)
    return metrics, ids_hypotheses
