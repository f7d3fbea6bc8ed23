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
