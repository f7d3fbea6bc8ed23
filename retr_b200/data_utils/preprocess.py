"""Batch preprocessing on the GPU: what ``RefCocoCaption.__getitem__`` does per sample with PIL on
the host (reference data_utils/refcoco.py:147-188, data_utils/utils.py:161-256) as ONE kernel launch per batch:
crop to the bounding box, context image with the target blanked, pad to square, antialiased bilinear resize to 224,
/255 + ImageNet normalisation, the two padding masks and the 5-d location features — bit-identical to the
PIL / torchvision pipeline (tests/test_preprocess.py).  Validation-mode transforms only (the training ColorJitter is a
random augmentation and stays with the caller)."""
import ctypes as C

import numpy as np
import torch

from .. import native

OUT_SIZE = 224


def preprocess_batch(images, boxes, device="cuda", return_global_context=True, return_location_features=True):
    """images: list of uint8 RGB arrays / tensors [H,W,3]; boxes: [B,4] (x, y, w, h) as annotated.
    Returns the dataloader's encoder inputs, batched on ``device``: (t_img [B,3,224,224], t_mask [B,224,224] bool
    [, g_img, g_mask] [, loc [B,5]]) — the tuple ``engine.pack_encoder_inputs`` takes."""
    lib = native.lib()
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("retr_b200 preprocessing runs on a CUDA device only (no CPU path)")
    B = len(images)
    boxes = np.asarray(boxes, dtype=np.float64).reshape(B, 4)
    arrs, hw, offsets, off = [], np.zeros((B, 2), np.int32), np.zeros(B, np.int64), 0
    max_side = lib.retr_preprocess_max_side(OUT_SIZE)
    for b, im in enumerate(images):
        a = im.cpu().numpy() if isinstance(im, torch.Tensor) else np.asarray(im)
        if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 3:
            raise ValueError("images must be uint8 arrays of shape [H, W, 3] (RGB)")
        H, W = a.shape[:2]
        x, y, w, h = (round(float(v)) for v in boxes[b])
        if x < 0 or y < 0:
            raise ValueError(f"bounding box {tuple(boxes[b])} starts outside the image (numpy slicing would wrap)")
        if min(x + w, W) - min(x, W) <= 0 or min(y + h, H) - min(y, H) <= 0:
            raise ValueError(f"bounding box {tuple(boxes[b])} is empty inside a {W}x{H} image")
        if max(H, W) > max_side:
            raise ValueError(f"image side {max(H, W)} exceeds the resampler's limit {max_side}")
        arrs.append(np.ascontiguousarray(a).reshape(-1))
        hw[b] = (H, W)
        offsets[b] = off
        off += a.size
    packed = torch.from_numpy(np.concatenate(arrs)).pin_memory().to(device, non_blocking=True)
    d_off = torch.from_numpy(offsets).to(device)
    d_hw = torch.from_numpy(hw).to(device)
    d_bb = torch.from_numpy(boxes).to(device)
    t_img = torch.empty(B, 3, OUT_SIZE, OUT_SIZE, dtype=torch.float32, device=device)
    t_mask = torch.empty(B, OUT_SIZE, OUT_SIZE, dtype=torch.uint8, device=device)
    g_img = torch.empty_like(t_img) if return_global_context else None
    g_mask = torch.empty_like(t_mask) if return_global_context else None
    loc = torch.empty(B, 5, dtype=torch.float32, device=device) if return_location_features else None
    with torch.cuda.device(device):
        native.check(lib.retr_preprocess(native.ptr(packed), native.ptr(d_off), native.ptr(d_hw), native.ptr(d_bb), B,
                                         OUT_SIZE, native.ptr(t_img), native.ptr(t_mask), native.ptr(g_img),
                                         native.ptr(g_mask), native.ptr(loc), native.stream_ptr(device)))
    out = [t_img, t_mask.bool()]
    if return_global_context:
        out += [g_img, g_mask.bool()]
    if return_location_features:
        out.append(loc)
    return tuple(out)
