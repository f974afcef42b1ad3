"""Pins oracle/aug_port.py (the restatement of the reference's input pipeline) against the real
third-party libraries the reference calls (Pillow, torchvision), both installed in this image: bit-exact for the uint8
paths, within documented rounding noise for the float convolution and the nearest-neighbour affine."""
import os
import random

import numpy as np
import pytest
import torch

PIL = pytest.importorskip("PIL")
tv = pytest.importorskip("torchvision")
from PIL import Image, ImageEnhance  # noqa: E402
import torchvision.transforms as T  # noqa: E402
import torchvision.transforms.functional as TF  # noqa: E402

from oracle import aug_port as A  # noqa: E402


def _img(seed, h=224, w=224, c=3):
    return np.random.default_rng(seed).integers(0, 256, (h, w, c) if c else (h, w), dtype=np.uint8)


@pytest.mark.parametrize("hw", [(480, 640), (1080, 1350), (100, 150), (224, 300), (531, 224), (224, 224)])
def test_pil_resize_bicubic_bit_exact(hw):
    """Image.open(p).resize((224, 224)) (dataset.py:38-41): down- and up-scaling, one axis unchanged, identity."""
    img = _img(hw[0], *hw)
    assert np.array_equal(np.asarray(Image.fromarray(img).resize((224, 224))), A.pil_resize_bicubic(img))
    g = _img(hw[1], *hw, c=0)
    assert np.array_equal(np.asarray(Image.fromarray(g).resize((224, 224))), A.pil_resize_bicubic(g))


def test_color_ops_bit_exact():
    img = _img(1)
    pil = Image.fromarray(img)
    assert np.array_equal(np.asarray(pil.convert("L")), A.to_l(img))
    for f in (0.6, 0.83, 1.0, 1.27, 1.4, 0.5, 1.5, 0.75, 1.25):
        assert np.array_equal(np.asarray(ImageEnhance.Brightness(pil).enhance(f)), A.adjust_brightness(img, f))
        assert np.array_equal(np.asarray(ImageEnhance.Contrast(pil).enhance(f)), A.adjust_contrast(img, f))
        assert np.array_equal(np.asarray(ImageEnhance.Color(pil).enhance(f)), A.adjust_saturation(img, f))
    for f in (-0.01, -0.004, 0.0, 0.0039, 0.0041, 0.01, 0.3, -0.5):
        assert np.array_equal(np.asarray(TF.adjust_hue(pil, f)), A.adjust_hue(img, f))


def test_hsv_conversions_over_the_colour_cube():
    """Pillow's rgb2hsv / hsv2rgb for every colour with even channel values plus all greys and extremes (2.1 M colours; set
    S4G_FULL_HSV=1 for all 2^24, which also passes)."""
    step = 1 if os.environ.get("S4G_FULL_HSV") else 2
    v = np.arange(0, 256, step, dtype=np.uint8)
    r, g, b = np.meshgrid(v, v, v, indexing="ij")
    cube = np.stack([r, g, b], -1).reshape(-1, len(v), 3)
    extra = np.stack([np.arange(256, dtype=np.uint8)] * 3, -1)[None]
    extra = np.concatenate([extra, 255 - extra, extra[:, ::-1] | 1], axis=0)
    for colours in (cube, extra):
        assert np.array_equal(np.asarray(Image.fromarray(colours).convert("HSV")), A.rgb_to_hsv(colours))
        assert np.array_equal(np.asarray(Image.fromarray(colours, mode="HSV").convert("RGB")), A.hsv_to_rgb(colours))


def test_color_jitter_composition():
    """ColorJitter.forward with the parameters ColorJitter.get_params drew, all 24 orders covered over the loop."""
    jit = T.ColorJitter(brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01)
    img = _img(3)
    pil = Image.fromarray(img)
    for seed in range(12):
        torch.manual_seed(seed)
        order, b, c, s, h = T.ColorJitter.get_params(jit.brightness, jit.contrast, jit.saturation, jit.hue)
        torch.manual_seed(seed)
        ref = np.asarray(jit(pil))
        assert np.array_equal(ref, A.color_jitter(img, order.tolist(), (b, c, s, h)))


def test_gaussian_blur_within_rounding_noise():
    smooth = np.asarray(Image.fromarray(_img(4, 28, 28)).resize((224, 224)))
    for img in (_img(5), smooth):
        for sigma in (0.001, 0.3, 0.8, 1.37, 2.0):
            ref = np.asarray(TF.gaussian_blur(Image.fromarray(img), [25, 25], [sigma, sigma])).astype(int)
            d = np.abs(ref - A.gaussian_blur(img, sigma))
            assert d.max() <= 1 and (d > 0).mean() < 5e-4, (sigma, d.max(), (d > 0).sum())


def test_tensor_conversions_exact():
    img = _img(6)
    ref = T.Normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])(T.ToTensor()(Image.fromarray(img))).numpy()
    assert np.array_equal(ref, A.to_tensor_normalize(img))
    assert np.array_equal(T.Grayscale()(T.ToTensor()(Image.fromarray(img))).numpy(), A.mask_to_tensor(img))
    g = _img(7, c=0)
    assert np.array_equal(T.Grayscale()(T.ToTensor()(Image.fromarray(g))).numpy(), A.mask_to_tensor(g))


def test_flips_and_affine_nearest():
    """TF.hflip / TF.vflip / TF.affine(nearest, fill) on tensors with the reference's parameter ranges (dataset.py:56-63):
    identical except for source coordinates within float32 noise of a half-integer (< 1e-5 of the pixels)."""
    x = torch.from_numpy(A.to_tensor_normalize(_img(8)))
    rnd = random.Random(0)
    bad = tot = 0
    for k in range(16):
        angle, tr = rnd.uniform(-180, 180), (rnd.uniform(-28, 28), rnd.uniform(-28, 28))
        sc, sh = rnd.uniform(0.5, 1.5), rnd.uniform(-22.5, 22.5)
        hf, vf = k & 1, k & 2
        y = TF.vflip(TF.hflip(x) if hf else x) if vf else (TF.hflip(x) if hf else x)
        ref = TF.affine(y, angle, tr, sc, sh, fill=-1.0).numpy()
        mine = A.geometric(x.numpy(), hf, vf, A.inverse_affine_matrix(angle, tr, sc, sh), -1.0)
        bad += int((ref != mine).any(0).sum())
        tot += 224 * 224
    assert bad <= 1e-5 * tot + 2, (bad, tot)
    assert np.array_equal(TF.hflip(x).numpy(), A.geometric(x.numpy(), True, False, None, 0.0))


def test_depth_dataset_helpers():
    """make_square and the I;16 resize of Depth_estimation/Data/dataset.py:9-16,46-50.  The reference pins Pillow 9.1.0, whose
    default filter for "I;16" is NEAREST; the installed Pillow defaults to BICUBIC for that mode, so NEAREST is requested here."""
    rng = np.random.default_rng(11)
    for h, w in ((1080, 1350), (480, 640), (224, 224), (100, 150), (720, 540)):
        d = rng.integers(0, 65536, (h, w), dtype=np.uint16)
        im = Image.fromarray(d)
        size = max(im.size)
        sq = Image.new("I;16", (size, size), 0)
        sq.paste(im, (int((size - im.size[0]) / 2), int((size - im.size[1]) / 2)))
        assert np.array_equal(np.array(sq), A.make_square(d))
        ref = np.array(sq.resize((224, 224), Image.NEAREST))
        assert np.array_equal(ref, A.pil_resize_nearest(A.make_square(d)))
        assert np.array_equal(torch.from_numpy(ref / 65535).float().numpy()[None], A.depth_target(d))
        rgb = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        sq3 = Image.new("RGB", (size, size), (0, 0, 0))
        sq3.paste(Image.fromarray(rgb), (int((size - w) / 2), int((size - h) / 2)))
        want = T.Normalize([0.485, 0.456, 0.406], [0.229, 0.224, 0.225])(T.ToTensor()(sq3.resize((224, 224)))).numpy()
        assert np.array_equal(want, A.depth_input(rgb))
