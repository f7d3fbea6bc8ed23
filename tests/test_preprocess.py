"""Preprocessing: the oracle against the fixture recorded from the unmodified reference (CPU), and
the CUDA kernel against the oracle (GPU) — bit-exact: the image path is integer fixed-point + two IEEE float ops, the
masks are booleans, the location features are double arithmetic rounded once to float32."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import preprocess_oracle as P


def _fixture():
    return torch.load(os.path.join(GOLDEN, "preprocess.pt"), weights_only=False)


def test_oracle_matches_reference_fixture():
    fx = _fixture()
    idx = fx["sample_index"]
    for rec in fx["scenes"]:
        img, bb = P.synthetic_scene(rec["seed"])
        assert img.shape == rec["shape"] and bb == rec["bb"]
        t_img, t_mask, g_img, g_mask, loc = P.preprocess_sample(img, bb)
        for name, x, m in (("t", t_img, t_mask), ("g", g_img, g_mask)):
            assert float(x.double().sum()) == rec[name + "_sum"]
            assert torch.equal(x.flatten()[idx], rec[name + "_samples"])
            assert torch.equal(torch.from_numpy(np.packbits(m.numpy())), rec[name + "_mask_bits"])
        assert torch.equal(loc, rec["loc"])


def test_image_and_mask_pad_offsets_differ_by_design():
    """ImageOps.pad pastes at round-half-even(diff/2), pad_mask_to_max starts at floor(diff/2): diff = 3 -> 2 vs 1."""
    img = np.full((40, 43, 3), 200, np.uint8)
    t_img, t_mask, *_ = P.preprocess_sample(img, (0.0, 0.0, 43.0, 40.0))
    assert t_mask.any() and not t_mask.all()


# ---- GPU ------------------------------------------------------------------------------------------------------
def _edge_scenes():
    g = np.random.default_rng(7)
    def noise(h, w):
        return g.integers(0, 256, (h, w, 3), dtype=np.uint8)
    return [
        (noise(300, 400), (10.4, 20.6, 224.0, 224.0)),          # crop side exactly the output size: no resampling
        (noise(300, 400), (100.0, 50.0, 37.0, 19.0)),           # tiny crop: upscaling
        (noise(224, 224), (0.0, 0.0, 224.0, 224.0)),            # box = whole image; context fully blanked
        (noise(480, 640), (600.2, 400.7, 90.0, 200.0)),         # box clipped by the image border
        (noise(641, 127), (3.5, 2.5, 100.5, 600.5)),            # .5 coordinates: Python round() is half-to-even
        (noise(131, 640), (0.0, 0.0, 640.0, 1.4)),              # one-pixel-high strip
    ]


@pytest.mark.gpu
def test_cuda_preprocess_bit_exact_vs_oracle():
    from retr_b200.data_utils import preprocess_batch
    scenes = [P.synthetic_scene(s) for s in range(10, 22)] + _edge_scenes()
    images, boxes = [s[0] for s in scenes], [s[1] for s in scenes]
    out = preprocess_batch(images, boxes, device="cuda")
    t_img, t_mask, g_img, g_mask, loc = (o.cpu() for o in out)
    assert t_img.shape == (len(scenes), 3, 224, 224) and t_mask.dtype == torch.bool
    for b, (img, bb) in enumerate(scenes):
        want = P.preprocess_sample(img, bb)
        for name, got, ref in (("t_img", t_img[b], want[0]), ("t_mask", t_mask[b], want[1]), ("g_img", g_img[b], want[2]),
                               ("g_mask", g_mask[b], want[3]), ("loc", loc[b], want[4])):
            assert torch.equal(got, ref), (b, name, img.shape, bb,
                                           int((got != ref).sum()), float((got.float() - ref.float()).abs().max()))


@pytest.mark.gpu
def test_cuda_preprocess_feeds_the_model_and_rejects_bad_boxes():
    from retr_b200.data_utils import preprocess_batch
    img, bb = P.synthetic_scene(5)
    t_only = preprocess_batch([img], [bb], return_global_context=False, return_location_features=False)
    assert len(t_only) == 2
    with pytest.raises(ValueError):
        preprocess_batch([img], [(-3.0, 0.0, 10.0, 10.0)])
    with pytest.raises(ValueError):
        preprocess_batch([img], [(5000.0, 0.0, 10.0, 10.0)])
