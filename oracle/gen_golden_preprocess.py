"""Generates tests/golden/preprocess.pt from the UNMODIFIED reference (run in the build container, where
/root/reference exists):  crop_image_to_bb / pad_img_to_max / pad_mask_to_max / compute_position_features from
data_utils/utils.py and get_transforms('val') from data_utils/refcoco.py, applied to oracle.preprocess_oracle's
seeded synthetic scenes exactly as RefCocoCaption.__getitem__ applies them (data_utils/refcoco.py:147-188).
Stored per scene: location features, both masks (bit-packed), and for both images a float64 sum plus 256 sampled
values — enough to pin a bit-exact restatement without committing megabytes."""
import os
import sys

import numpy as np
import torch
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
from data_utils import refcoco as RC  # noqa: E402
from data_utils import utils as RU  # noqa: E402
from oracle import preprocess_oracle as P  # noqa: E402


class Cfg:
    backbone = "ResNet101"


def sample_index():
    g = torch.Generator().manual_seed(99)
    return torch.randint(0, 3 * 224 * 224, (256,), generator=g)


def main():
    tf = RC.get_transforms("val", Cfg())
    idx = sample_index()
    scenes = []
    for seed in range(6):
        img, bb = P.synthetic_scene(seed)
        image = Image.fromarray(img)
        t_image, t_mask, c_image, c_mask = RU.crop_image_to_bb(image, bb, return_context=True)
        rec = {"seed": seed, "bb": bb, "shape": img.shape}
        for name, im, mk in (("t", t_image, t_mask), ("g", c_image, c_mask)):
            x = tf["transform"](tf["resize"](RU.pad_img_to_max(im)))
            m = tf["resize"](RU.pad_mask_to_max(mk).unsqueeze(0)).squeeze(0)
            rec[name + "_sum"] = float(x.double().sum())
            rec[name + "_samples"] = x.flatten()[idx].clone()
            rec[name + "_mask_bits"] = torch.from_numpy(np.packbits(m.numpy()))
        rec["loc"] = RU.compute_position_features(image, bb)
        scenes.append(rec)
    out = os.path.join(ROOT, "tests", "golden", "preprocess.pt")
    torch.save({"scenes": scenes, "sample_index": idx}, out)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
