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
    loss = tr.loss_and_grads(*inputs, caps[:, 1:], dropout_p=p, seed=seed)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(tr.gflat)
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
