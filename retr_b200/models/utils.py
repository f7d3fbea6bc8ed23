"""Host-side helpers with the reference's names (models/utils.py).

``NestedTensor`` is the (images, padding-mask) pair every caller of the model hands in;
``shape`` is the *mask* shape so ``samples[0].shape[0]`` is the batch size
(reference models/utils.py:112-133, relied on by eval_utils/decode.py:60).
"""
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist
from torch import Tensor


class NestedTensor:
    def __init__(self, tensors: Tensor, mask: Optional[Tensor]):
        self.tensors = tensors
        self.mask = mask
        self.shape = mask.shape          # raises on mask=None, like the reference

    def to(self, device):
        m = None if self.mask is None else self.mask.to(device)
        return NestedTensor(self.tensors.to(device), m)

    def pin_memory(self):
        return NestedTensor(self.tensors.pin_memory(), self.mask.pin_memory())

    def decompose(self):
        return self.tensors, self.mask

    def __repr__(self):
        return str(self.tensors)


def nested_tensor_from_tensor_list(tensor_list: List[Tensor]) -> NestedTensor:
    """Pad a list of CxHxW images to a common size; mask is True on padding
    (reference models/utils.py:92-109).  A batched 4-D tensor iterates as a list too."""
    if tensor_list[0].ndim != 3:
        raise ValueError("not supported")
    shapes = torch.tensor([list(t.shape) for t in tensor_list])
    c, h, w = shapes.max(0).values.tolist()
    first = tensor_list[0]
    batch = torch.zeros((len(tensor_list), c, h, w), dtype=first.dtype, device=first.device)
    mask = torch.ones((len(tensor_list), h, w), dtype=torch.bool, device=first.device)
    for i, img in enumerate(tensor_list):
        ci, hi, wi = img.shape
        batch[i, :ci, :hi, :wi].copy_(img)
        mask[i, :hi, :wi] = False
    return NestedTensor(batch, mask)


def generate_square_subsequent_mask(sz: int) -> Tensor:
    """Additive causal mask: 0 on/below the diagonal, -inf above (reference models/utils.py:50-57)."""
    return torch.full((sz, sz), float("-inf")).triu(1)


def with_pos_embed(tensor: Tensor, pos: Optional[Tensor]) -> Tensor:
    return tensor if pos is None else tensor + pos


def ensure_unmasked_values(mask: Tensor, unmasked_ratio: float = 0.01) -> Tensor:
    """If a sample's context mask is entirely True, un-mask a random ``unmasked_ratio`` of
    its positions so attention never sees an all-masked row (reference models/utils.py:60-89;
    same numpy RNG call, same in-place update).  Rare path, runs on the host."""
    b, h, w = mask.shape
    has_unmasked = (~mask.reshape(b, -1)).any(dim=1)
    if not bool(has_unmasked.all()):
        n = h * w
        filler = torch.ones(n, dtype=torch.bool, device=mask.device)
        pick = np.random.choice(np.arange(n), replace=False, size=round(n * unmasked_ratio))
        filler[torch.as_tensor(pick, device=mask.device)] = False
        mask[~has_unmasked] = filler.reshape(h, w)
    return mask


def is_dist_avail_and_initialized() -> bool:
    return dist.is_available() and dist.is_initialized()


def get_rank() -> int:
    return dist.get_rank() if is_dist_avail_and_initialized() else 0


def get_world_size() -> int:
    return dist.get_world_size() if is_dist_avail_and_initialized() else 1


def is_main_process() -> bool:
    return get_rank() == 0
