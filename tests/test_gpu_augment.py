"""GPU input pipeline (csrc/augment.cu) against the oracle restatement of the reference's Pillow / torchvision
transforms (oracle/aug_port.py, itself pinned against the real libraries in tests/test_aug_oracle.py)."""
import random

import numpy as np
import pytest
import torch

from oracle import aug_port as A  # checker only

pytestmark = pytest.mark.gpu

SIZES = [(480, 640), (1080, 1350), (100, 150), (224, 300), (531, 224), (224, 224), (720, 576), (333, 999)]


def _photo(seed, h, w, c=3):
    """Smooth structure + noise: something a blur and a resampler act on non-trivially."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    base = [127 + 100 * np.sin(xx / (11.0 + k) + seed) * np.cos(yy / (7.0 + 2 * k)) for k in range(c or 1)]
    img = np.stack(base, -1) + rng.normal(0, 12, (h, w, c or 1))
    img = np.clip(img, 0, 255).astype(np.uint8)
    return img if c else img[..., 0]


def _batch(mask_channels):
    imgs = [_photo(i, h, w) for i, (h, w) in enumerate(SIZES)]
    masks = [(_photo(100 + i, h, w, mask_channels) > 140).astype(np.uint8) * 255 for i, (h, w) in enumerate(SIZES)]
    return imgs, masks


@pytest.mark.parametrize("mask_channels", [0, 3])
def test_train_transform_matches_oracle(mask_channels):
    from ssl4gie_b200 import GpuInputPipeline
    imgs, masks = _batch(mask_channels)
    pipe = GpuInputPipeline(train=True)
    torch.manual_seed(3)
    random.seed(3)
    params = [pipe.draw() for _ in imgs]
    params[0].sigma = 0.001          # range ends
    params[1].sigma = 2.0
    params[2].affine = None          # flips only
    stages = {}
    x, y = pipe(imgs, masks, params=params, stages=stages)
    torch.cuda.synchronize()
    n_blur = n_geo = tot = 0
    for i, (img, mk, p) in enumerate(zip(imgs, masks, params)):
        r = A.pil_resize_bicubic(img)
        assert np.array_equal(stages["resized"][i].cpu().numpy(), r), ("resize", SIZES[i])          # Pillow resampler: bit-exact
        j = A.color_jitter(r, p.order, (p.brightness, p.contrast, p.saturation, p.hue))
        assert np.array_equal(stages["jittered"][i].cpu().numpy(), j), ("jitter", i, p.order)        # 8-bit ColorJitter: bit-exact
        ref = A.to_tensor_normalize(A.gaussian_blur(j, p.sigma))
        got = stages["blurred"][i].cpu().numpy()
        d = np.abs(got - ref)
        lsb = (1.0 / 255) / A.IMAGENET_STD.reshape(3, 1, 1)
        assert (d <= lsb * 1.001).all(), ("blur", i, d.max())                                        # at most one 8-bit step ...
        n_blur += int((d > 1e-6).sum())
        mref = A.mask_to_tensor(A.pil_resize_bicubic(mk))
        assert np.array_equal(stages["mask"][i].cpu().numpy(), mref), ("mask", i)
        m = A.inverse_affine_matrix(*p.affine) if p.affine is not None else None
        gx = A.geometric(got, p.hflip, p.vflip, m, -1.0)
        gy = A.geometric(mref, p.hflip, p.vflip, m, 0.0)
        n_geo += int((x[i].cpu().numpy() != gx).any(0).sum()) + int((y[i].cpu().numpy() != gy).any(0).sum())
        tot += 224 * 224
    print(f"[augment] blur: {n_blur} of {3 * tot} values off by one 8-bit step; geometric: {n_geo} of {2 * tot} pixels differ")
    assert n_blur <= 5e-4 * 3 * tot            # ... and only where the exact value sits on a rounding boundary
    assert n_geo <= 1e-5 * 2 * tot + 2


def test_eval_transform_is_exact():
    from ssl4gie_b200 import GpuInputPipeline
    imgs, _ = _batch(0)
    x, y = GpuInputPipeline(train=False)(imgs)
    assert y is None
    for i, img in enumerate(imgs):
        assert np.array_equal(x[i].cpu().numpy(), A.to_tensor_normalize(A.pil_resize_bicubic(img))), SIZES[i]


def test_rejects_bad_input():
    from ssl4gie_b200 import GpuInputPipeline
    pipe = GpuInputPipeline(train=False)
    with pytest.raises(ValueError):
        pipe([np.zeros((10, 10, 4), dtype=np.uint8)])
    with pytest.raises(RuntimeError, match="too long"):
        pipe([np.zeros((8, 6000, 3), dtype=np.uint8)])


def test_depth_pipeline_matches_oracle():
    """GpuInputPipeline(task="depth") (Depth_estimation/Data/dataset.py:43-70, Data/dataloaders.py:55-74): square padding + bicubic
    resize + ColorJitter + Normalize of the image and nearest resize / 65535 of the 16-bit depth map are all exact; flips exact."""
    from ssl4gie_b200 import GpuInputPipeline
    rng = np.random.default_rng(21)
    sizes = [(1080, 1350), (480, 640), (224, 224), (100, 150), (720, 540), (531, 531)]
    imgs = [_photo(30 + i, h, w) for i, (h, w) in enumerate(sizes)]
    depths = [rng.integers(0, 65536, (h, w), dtype=np.uint16) for h, w in sizes]
    pipe = GpuInputPipeline(train=True, task="depth")
    torch.manual_seed(5)
    random.seed(5)
    params = [pipe.draw() for _ in imgs]
    assert all(p.sigma is None and p.affine is None for p in params)
    params[0].hflip, params[0].vflip, params[1].hflip, params[1].vflip = True, True, False, True
    x, y = pipe(imgs, depths, params=params)
    for i, (img, d, p) in enumerate(zip(imgs, depths, params)):
        want_x = A.geometric(A.depth_input(img, p.order, (p.brightness, p.contrast, p.saturation, p.hue)), p.hflip, p.vflip, None, -1.0)
        want_y = A.geometric(A.depth_target(d), p.hflip, p.vflip, None, 0.0)
        assert np.array_equal(x[i].cpu().numpy(), want_x), ("image", sizes[i])
        assert np.array_equal(y[i].cpu().numpy(), want_y), ("depth", sizes[i])
    xe, ye = GpuInputPipeline(train=False, task="depth")(imgs, depths)
    for i, (img, d) in enumerate(zip(imgs, depths)):
        assert np.array_equal(xe[i].cpu().numpy(), A.depth_input(img)) and np.array_equal(ye[i].cpu().numpy(), A.depth_target(d))
