"""Sample-sharded greedy evaluation across the GPUs of one box (the shape of the
reference's ``engine.eval_model`` loop, engine.py:125-187, without its host-side NLG metrics).

Each expression depends only on its own sample, so rank r decodes the contiguous block
``[r*ceil(N/G), (r+1)*ceil(N/G))`` in batches, with no collective on the data path; the only
exchange is one gather of the pruned id rows at the end.  Parity across shardings is defined on
``prune_cap_ids(clean=True)`` output (raw rows differ after EOS because the early return is per batch).
"""
import math

import torch
import torch.distributed as dist

from .eval_utils.decode import greedy, greedy_stream, prune_cap_ids
from .models.caption import _CaptionBase


def shard_range(n_samples, rank, world):
    per = math.ceil(n_samples / world)
    lo = min(rank * per, n_samples)
    return lo, min(lo + per, n_samples)


def sharded_greedy(make_samples, n_samples, model, batch_size, max_len, device, bos_token, eos_token,
                   pad_token=0, gather=True):
    """``make_samples(lo, hi)`` -> the ``samples`` tuple for samples [lo, hi).  Returns the pruned id
    lists of ALL samples on every rank (``gather=True``) or of this rank's shard."""
    rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
    lo, hi = shard_range(n_samples, rank, world)
    local = []
    spans = [(b0, min(b0 + batch_size, hi)) for b0 in range(lo, hi, batch_size)]
    if isinstance(model, _CaptionBase) and torch.device(device).type == "cuda":
        # native model: batches stream through one context, the next batch's images cross PCIe while this one decodes
        if spans:
            model.max_batch = max(model.max_batch, batch_size)
        rows = greedy_stream((make_samples(b0, b1) for b0, b1 in spans), model, max_len=max_len, bos_token=bos_token,
                             eos_token=eos_token)
    else:
        rows = (greedy(make_samples(b0, b1), model, max_len=max_len, device=device, bos_token=bos_token,
                       eos_token=eos_token).cpu() for b0, b1 in spans)
    for ids in rows:
        local += prune_cap_ids(ids.tolist(), clean=True, pad_token=pad_token, bos_token=bos_token,
                               eos_token=eos_token)
    if not gather or world == 1:
        return local
    parts = [None] * world
    dist.all_gather_object(parts, local)
    return [row for part in parts for row in part]
