"""CPU oracle of the sample preprocessing that precedes the hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates ``RefCocoCaption.__getitem__``'s image half in validation mode (data_utils/refcoco.py:105-188,
``get_transforms('val')`` :14-46) and its helpers (data_utils/utils.py:161-256), calling the same third-party
routines the reference calls for the arithmetic: PIL (``Image.fromarray``, ``ImageOps.pad``, the antialiased bilinear
``Image.resize`` behind torchvision's ``Resize`` on a PIL image) and torchvision (``Resize`` on the bool mask tensor,
``ToTensor``, ``Normalize``).  Pinned by tests/test_preprocess.py against the unmodified reference functions (when
/root/reference is present) and against tests/golden/preprocess.pt.

Two quirks of the reference that the CUDA path must reproduce:
  * the padded IMAGE is pasted at ``round(diff / 2)`` (Python round: half to even, ``ImageOps.pad``) while the padded MASK
    starts at ``floor(diff / 2)`` (``pad_mask_to_max``): for diff = 2 (mod 4) + 1 ... they differ by one pixel;
  * ``Resize`` on the BOOL mask interpolates in float32 (bilinear, antialias) and casts back with ``.to(bool)``: every
    output pixel that receives ANY positive weight from a masked input pixel is masked.
"""
from math import ceil, floor

import numpy as np
import torch
import torch.nn.functional as F
from PIL import Image, ImageOps
from torchvision.transforms import Compose, InterpolationMode, Normalize, Resize, ToTensor

CROP_SIZE = [224]                       # ResNet*_Weights.DEFAULT.transforms().crop_size
MEAN, STD = [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]


def transforms():
    """get_transforms('val', config) — data_utils/refcoco.py:14-46."""
    return {"resize": Resize(size=CROP_SIZE, interpolation=InterpolationMode.BILINEAR),
            "transform": Compose([ToTensor(), Normalize(mean=MEAN, std=STD)])}


def crop_image_to_bb(image, bb):
    """data_utils/utils.py:161-196 with return_context=True."""
    arr = np.array(image)
    x, y, w, h = round(bb[0]), round(bb[1]), round(bb[2]), round(bb[3])
    x_min, x_max, y_min, y_max = x, x + w, y, y + h
    target_region = arr[y_min:y_max, x_min:x_max, :]
    target_mask = np.zeros_like(target_region[:, :, 0]).astype(bool)
    target_image = Image.fromarray(target_region)
    context_mask = np.zeros_like(arr[:, :, 0]).astype(bool)
    arr[y_min:y_max, x_min:x_max, :] = 0
    context_mask[y_min:y_max, x_min:x_max] = True
    return target_image, target_mask, Image.fromarray(arr), context_mask


def compute_position_features(image, bb):
    """data_utils/utils.py:199-229."""
    ih, iw, _ = np.array(image).shape
    x, y, w, h = bb
    return torch.Tensor([x / iw, y / ih, (x + w) / iw, (y + h) / ih, (w * h) / (iw * ih)])


def pad_img_to_max(image):
    """data_utils/utils.py:232-240."""
    m = max(image.size)
    return ImageOps.pad(image, size=(m, m), centering=(0.5, 0.5), color=0)


def pad_mask_to_max(mask):
    """data_utils/utils.py:243-256."""
    mask = torch.from_numpy(mask) if isinstance(mask, np.ndarray) else mask
    x, y = mask.shape
    if x == y:
        return mask
    lo, hi = sorted(mask.shape)
    pad = (hi - lo) / 2
    if x > y:
        return F.pad(mask, (floor(pad), ceil(pad), 0, 0), "constant", True)
    return F.pad(mask, (0, 0, floor(pad), ceil(pad)), "constant", True)


def preprocess_sample(image, bb):
    """image: uint8 HxWx3 array (RGB); bb: (x, y, w, h).  -> t_img [3,224,224] f32, t_mask [224,224] bool, g_img,
    g_mask, loc [5] — the encoder inputs of data_utils/refcoco.py:147-188 (global context + location features)."""
    tf = transforms()
    image = Image.fromarray(np.ascontiguousarray(image))
    t_image, t_mask, c_image, c_mask = crop_image_to_bb(image, bb)
    out = []
    for img, msk in ((t_image, t_mask), (c_image, c_mask)):
        img = tf["transform"](tf["resize"](pad_img_to_max(img)))
        msk = tf["resize"](pad_mask_to_max(msk).unsqueeze(0)).squeeze(0)
        out += [img, msk]
    out.append(compute_position_features(image, bb))
    return tuple(out)


def synthetic_scene(seed):
    """A seeded COCO-shaped uint8 image (noise over smooth gradients: every rounding decision of the fixed-point
    resampler is exercised) and a bounding box inside it."""
    g = np.random.default_rng(seed)
    H, W = int(g.integers(120, 641)), int(g.integers(120, 641))
    yy, xx = np.mgrid[0:H, 0:W]
    base = np.stack([(xx * 255 // max(W - 1, 1)), (yy * 255 // max(H - 1, 1)), ((xx + yy) % 256)], -1)
    img = np.clip(base + g.integers(-60, 61, (H, W, 3)), 0, 255).astype(np.uint8)
    w, h = float(g.uniform(12, W * 0.8)), float(g.uniform(12, H * 0.8))
    x, y = float(g.uniform(0, W - w)), float(g.uniform(0, H - h))
    return img, (x, y, w, h)
