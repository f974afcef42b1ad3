"""ORACLE — test infrastructure, NOT product code.

numpy restatement of the reference's training-time input pipeline:
  Binary_segmentation/Data/dataset.py:34-64   PIL resize to 224, transforms, joint flips, joint affine (fill -1 / 0)
  Binary_segmentation/Data/dataloaders.py:62-76 ColorJitter(0.4, 0.5, 0.25, 0.01), GaussianBlur((25, 25), (0.001, 2.0)),
                                               ToTensor, Normalize(ImageNet)
The arithmetic lives in third-party dependencies that are not in /root/reference: Pillow (requirements.txt pins 9.x; the
8-bit resampler, ImageEnhance and the RGB<->HSV converters have not changed since) and torchvision 0.x
(transforms.functional: adjust_hue, gaussian_blur, affine).  Both are installed in this image, so tests/test_aug_oracle.py
pins every function below against the real libraries (bit-exact for the uint8 paths, exhaustively over all 2^24 colours for
the HSV round trip); nothing here is imported by the product package.
"""
import math

import numpy as np

# ----------------------------------------------------------------------------------------------- PIL Image.resize
PRECISION_BITS = 32 - 8 - 2  # Pillow src/libImaging/Resample.c


def _bicubic(x, a=-0.5):
    x = abs(x)
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def resample_coeffs(in_size, out_size, support=2.0, filt=_bicubic):
    """precompute_coeffs + normalize_coeffs_8bpc: per output index (xmin, integer weights)."""
    scale = filterscale = in_size / out_size
    if filterscale < 1.0:
        filterscale = 1.0
    support = support * filterscale
    ss = 1.0 / filterscale
    out = []
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        w = np.array([filt((x + xmin - center + 0.5) * ss) for x in range(xmax)], dtype=np.float64)
        tot = 0.0
        for v in w:      # sequential, as the C loop (numpy's sum is pairwise for longer runs)
            tot += float(v)
        if tot != 0.0:
            w = w / tot
        k = np.where(w < 0, (-0.5 + w * (1 << PRECISION_BITS)).astype(np.int64), (0.5 + w * (1 << PRECISION_BITS)).astype(np.int64))
        # C casts truncate toward zero: astype(int64) does the same
        out.append((xmin, k))
    return out


def _resample_axis0(img, out_size):
    """img [N, ...] uint8 -> [out_size, ...] uint8 along axis 0 with the 8-bit fixed-point accumulator."""
    coeffs = resample_coeffs(img.shape[0], out_size)
    res = np.empty((out_size,) + img.shape[1:], dtype=np.uint8)
    src = img.astype(np.int64)
    for i, (xmin, k) in enumerate(coeffs):
        acc = (1 << (PRECISION_BITS - 1)) + np.tensordot(k, src[xmin:xmin + len(k)], axes=(0, 0))
        res[i] = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return res


def pil_resize_bicubic(img, out_w=224, out_h=224):
    """Image.resize((out_w, out_h)) of an 8-bit image [H, W] or [H, W, C] (default resample BICUBIC, dataset.py:38-41):
    horizontal pass, then vertical pass, each rounding to uint8."""
    h = np.swapaxes(_resample_axis0(np.swapaxes(img, 0, 1), out_w), 0, 1) if img.shape[1] != out_w else img
    return _resample_axis0(h, out_h) if img.shape[0] != out_h else h


# ----------------------------------------------------------------------------------------------- ColorJitter on PIL images
def to_l(rgb):
    """Image.convert("L"): ITU-R 601-2 luma in 16.16 fixed point."""
    r, g, b = (rgb[..., i].astype(np.int64) for i in range(3))
    return ((r * 19595 + g * 38470 + b * 7471 + 0x8000) >> 16).astype(np.uint8)


def _blend(deg, img, f):
    """Image.blend(degenerate, image, factor) for 8-bit images: float32 interpolation, truncation (factor in [0, 1]) or
    clipping + truncation (extrapolation)."""
    a = deg.astype(np.float32)
    t = a + np.float32(f) * (img.astype(np.float32) - a)
    if 0.0 <= f <= 1.0:
        return t.astype(np.uint8)
    return np.where(t <= 0, 0, np.where(t >= 255, 255, t)).astype(np.uint8)


def adjust_brightness(rgb, f):
    """F.adjust_brightness on a PIL image = ImageEnhance.Brightness: blend with black."""
    return _blend(np.zeros_like(rgb), rgb, f)


def adjust_contrast(rgb, f):
    """ImageEnhance.Contrast: blend with the constant image of the rounded mean luma."""
    mean = int(to_l(rgb).astype(np.float64).mean() + 0.5)
    return _blend(np.full_like(rgb, mean), rgb, f)


def adjust_saturation(rgb, f):
    """ImageEnhance.Color: blend with the grey version of the image."""
    return _blend(np.repeat(to_l(rgb)[..., None], 3, axis=-1), rgb, f)


def rgb_to_hsv(rgb):
    """Pillow Convert.c rgb2hsv (8-bit H, S, V).  The C code mixes float variables with double literals; the promotions
    below are the ones that reproduce it for all 2^24 colours (tests/test_aug_oracle.py)."""
    f32, f64 = np.float32, np.float64
    r, g, b = (rgb[..., i].astype(np.int64) for i in range(3))
    maxc = np.maximum(r, np.maximum(g, b))
    minc = np.minimum(r, np.minimum(g, b))
    cr = (maxc - minc).astype(f32)
    safe = np.where(cr == 0, f32(1), cr).astype(f32)
    s = cr / np.where(maxc == 0, 1, maxc).astype(f32)
    rc, gc, bc = ((maxc - c).astype(f32) / safe for c in (r, g, b))
    h = np.where(r == maxc, (bc - gc).astype(f32),                                      # float - float
                 np.where(g == maxc, (2.0 + rc.astype(f64) - bc.astype(f64)).astype(f32),   # 2.0 + rc - bc in double -> float
                          (4.0 + gc.astype(f64) - rc.astype(f64)).astype(f32)))
    h = np.fmod(h.astype(f64) / 6.0 + 1.0, 1.0).astype(f32)                            # double, stored to the float h
    uh = np.clip((h.astype(f64) * 255.0).astype(np.int64), 0, 255)
    us = np.clip((s.astype(f64) * 255.0).astype(np.int64), 0, 255)
    grey = maxc == minc
    return np.stack([np.where(grey, 0, uh), np.where(grey, 0, us), maxc], axis=-1).astype(np.uint8)


def hsv_to_rgb(hsv):
    """Pillow Convert.c hsv2rgb: i / f from h * 6.0 / 255.0 in double (f, fs stored as float), C round() (half away from 0)."""
    f32, f64 = np.float32, np.float64
    h, s, v = (hsv[..., i].astype(np.int64) for i in range(3))
    hd = h.astype(f64) * 6.0 / 255.0
    i = np.floor(hd).astype(np.int64)
    f = (hd - i.astype(f64)).astype(f32)
    fs = (s.astype(f64) / 255.0).astype(f32)
    vf = v.astype(f64)

    def rnd(x):  # x >= 0 here
        return np.clip(np.floor(x + 0.5).astype(np.int64), 0, 255)

    p = rnd(vf * (1.0 - fs.astype(f64)))
    q = rnd(vf * (1.0 - (fs * f).astype(f64)))                     # fs * f is a float product
    t = rnd(vf * (1.0 - fs.astype(f64) * (1.0 - f.astype(f64))))
    i = i % 6
    r = np.choose(i, [v, q, p, p, t, v])
    g = np.choose(i, [t, v, v, q, p, p])
    b = np.choose(i, [p, p, t, v, v, q])
    grey = s == 0
    return np.stack([np.where(grey, v, r), np.where(grey, v, g), np.where(grey, v, b)], axis=-1).astype(np.uint8)


def adjust_hue(rgb, f):
    """torchvision F.adjust_hue on a PIL image (_functional_pil.adjust_hue): H += uint8(f * 255) with wrap-around."""
    hsv = rgb_to_hsv(rgb)
    hsv[..., 0] = (hsv[..., 0].astype(np.int64) + int(np.uint8(np.int64(f * 255)))) & 255
    return hsv_to_rgb(hsv)


COLOR_OPS = (adjust_brightness, adjust_contrast, adjust_saturation, adjust_hue)


def color_jitter(rgb, order, factors):
    """transforms.ColorJitter.forward (dataloaders.py:64-66): the four adjustments in the drawn order `order` (a permutation
    of 0..3 = brightness, contrast, saturation, hue) with the drawn factors[4]; every step re-quantises to uint8."""
    for k in order:
        rgb = COLOR_OPS[k](rgb, factors[k])
    return rgb


# ----------------------------------------------------------------------------------------------- GaussianBlur((25, 25), sigma)
def gaussian_kernel1d(ksize, sigma):
    """torchvision _get_gaussian_kernel1d in float32."""
    half = (ksize - 1) * 0.5
    x = np.linspace(-half, half, ksize, dtype=np.float32)
    pdf = np.exp(np.float32(-0.5) * (x / np.float32(sigma)) ** 2).astype(np.float32)
    return (pdf / pdf.sum(dtype=np.float32)).astype(np.float32)


def gaussian_blur(rgb, sigma, ksize=25):
    """F.gaussian_blur on an 8-bit image: float conv with the outer-product kernel on reflect-padded input, round half to
    even, back to uint8.  Accumulated here in float64 (torchvision: float32 conv2d, order unspecified): results agree except
    where the exact value sits within float32 noise of a .5 boundary (a handful of pixels per image, +-1)."""
    k = gaussian_kernel1d(ksize, sigma).astype(np.float64)
    pad = ksize // 2
    x = np.pad(rgb.astype(np.float64), ((pad, pad), (pad, pad), (0, 0)), mode="reflect")
    h = sum(k[i] * x[:, i:i + rgb.shape[1]] for i in range(ksize))
    v = sum(k[i] * h[i:i + rgb.shape[0]] for i in range(ksize))
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


IMAGENET_MEAN = np.array([0.485, 0.456, 0.406], dtype=np.float32)
IMAGENET_STD = np.array([0.229, 0.224, 0.225], dtype=np.float32)


def to_tensor_normalize(rgb):
    """ToTensor + Normalize (dataloaders.py:67-68): [H, W, 3] uint8 -> [3, H, W] float32."""
    x = rgb.astype(np.float32) / np.float32(255)
    return np.transpose((x - IMAGENET_MEAN) / IMAGENET_STD, (2, 0, 1))


def mask_to_tensor(rgb_or_l):
    """ToTensor + Grayscale (dataloaders.py:72-74): 3-channel masks through torchvision's rgb_to_grayscale
    (0.2989 r + 0.587 g + 0.114 b in float32), single-channel masks unchanged."""
    x = rgb_or_l.astype(np.float32) / np.float32(255)
    if x.ndim == 2:
        return x[None]
    return (np.float32(0.2989) * x[..., 0] + np.float32(0.587) * x[..., 1] + np.float32(0.114) * x[..., 2])[None]


# ----------------------------------------------------------------------------------------------- flips + affine (dataset.py:46-63)
def inverse_affine_matrix(angle, translate, scale, shear):
    """torchvision _get_inverse_affine_matrix with center (0, 0) (F.affine on tensors) and shear = (shear, 0)."""
    rot, sx, sy = math.radians(angle), math.radians(shear), 0.0
    tx, ty = translate
    a = math.cos(rot - sy) / math.cos(sy)
    b = -math.cos(rot - sy) * math.tan(sx) / math.cos(sy) - math.sin(rot)
    c = math.sin(rot - sy) / math.cos(sy)
    d = -math.sin(rot - sy) * math.tan(sx) / math.cos(sy) + math.cos(rot)
    m = [x / scale for x in (d, -b, 0.0, -c, a, 0.0)]
    m[2] += m[0] * (-tx) + m[1] * (-ty)
    m[5] += m[3] * (-tx) + m[4] * (-ty)
    return m


def affine_source_index(matrix, size=224):
    """Source pixel (iy, ix) of every output pixel for F.affine(..., interpolation=NEAREST) on a size x size tensor:
    _gen_affine_grid + grid_sample(nearest, zeros, align_corners=False) in float32; (-1, -1) where the source is outside."""
    f32 = np.float32
    theta = np.array(matrix, dtype=f32).reshape(2, 3)
    r = (theta.T / np.array([0.5 * size, 0.5 * size], dtype=f32)).astype(f32)   # [3, 2]
    xs = np.linspace(-size * 0.5 + 0.5, size * 0.5 + 0.5 - 1, size, dtype=f32)
    gx = (xs[None, :] * r[0, 0] + xs[:, None] * r[1, 0] + r[2, 0]).astype(f32)
    gy = (xs[None, :] * r[0, 1] + xs[:, None] * r[1, 1] + r[2, 1]).astype(f32)
    ix = np.rint((((gx + f32(1)) * f32(size) - f32(1)) / f32(2)).astype(f32)).astype(np.int64)
    iy = np.rint((((gy + f32(1)) * f32(size) - f32(1)) / f32(2)).astype(f32)).astype(np.int64)
    ok = (ix >= 0) & (ix < size) & (iy >= 0) & (iy < size)
    return np.where(ok, iy, -1), np.where(ok, ix, -1)


def geometric(x, hflip, vflip, matrix, fill):
    """dataset.py:46-63 on a [C, H, W] float tensor: TF.hflip, TF.vflip, then TF.affine(nearest, fill)."""
    if hflip:
        x = x[:, :, ::-1]
    if vflip:
        x = x[:, ::-1, :]
    if matrix is None:
        return np.ascontiguousarray(x)
    iy, ix = affine_source_index(matrix, x.shape[-1])
    out = x[:, np.clip(iy, 0, None), np.clip(ix, 0, None)]
    return np.where((iy < 0)[None], np.float32(fill), out).astype(np.float32)


# ----------------------------------------------------------------------------------------------- depth dataset (C3VD-style pairs)
def make_square(img):
    """Depth_estimation/Data/dataset.py:9-16: paste the image centred on a zero square of side max(h, w)
    (offsets int((size - w) / 2), int((size - h) / 2)).  img: [h, w, 3] uint8 or [h, w] uint16."""
    h, w = img.shape[:2]
    size = max(h, w)
    out = np.zeros((size, size) + img.shape[2:], dtype=img.dtype)
    oy, ox = int((size - h) / 2), int((size - w) / 2)
    out[oy:oy + h, ox:ox + w] = img
    return out


def pil_resize_nearest(img, out=224):
    """Image.resize((224, 224)) of an "I;16" image under the reference's pinned Pillow 9.1.0 (requirements.txt:3), whose default
    filter for the special modes is NEAREST (Image.py: `resample = NEAREST if ";" in mode else BICUBIC`; newer Pillow versions
    resample I;16 with BICUBIC): source index = floor((i + 0.5) * in / out) on both axes."""
    iy = np.floor((np.arange(out) + 0.5) * img.shape[0] / out).astype(np.int64)
    ix = np.floor((np.arange(out) + 0.5) * img.shape[1] / out).astype(np.int64)
    return img[np.ix_(iy, ix)]


def depth_target(depth_u16):
    """dataset.py:47-53: make_square(I;16) -> resize 224 -> np.array(y) / 65535 (float64) -> ToTensor -> .float(): [1, 224, 224]."""
    y = pil_resize_nearest(make_square(depth_u16)).astype(np.float64) / 65535
    return y.astype(np.float32)[None]


def depth_input(rgb, order=None, factors=None):
    """dataset.py:46 + Data/dataloaders.py:55-63: make_square -> bicubic resize 224 -> (training: ColorJitter) -> ToTensor ->
    Normalize.  No GaussianBlur in the depth transforms."""
    x = pil_resize_bicubic(make_square(rgb))
    if order is not None:
        x = color_jitter(x, order, factors)
    return to_tensor_normalize(x)
