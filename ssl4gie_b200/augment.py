"""GPU-side input pipeline: the per-sample transforms of the reference's DataLoader workers applied to a
whole batch of decoded 8-bit images on the device.

Mirrors ``Binary_segmentation/Data/dataloaders.py:62-76`` (ColorJitter(0.4, 0.5, 0.25, 0.01), GaussianBlur((25, 25),
(0.001, 2.0)), ToTensor, Normalize; ToTensor + Grayscale for masks) and ``Binary_segmentation/Data/dataset.py:34-64``
(PIL resize to 224 x 224, joint hflip / vflip with p = 0.5, joint affine with fill -1.0 / 0.0).  Random parameters are drawn on
the host exactly as the reference draws them (torchvision's get_params from torch's generator, then ``random.uniform``); the
pixel work runs in csrc/augment.cu through the C ABI (s4g_aug_*).  There is no CPU fallback."""
import ctypes
import math
import random

import numpy as np
import torch

from . import _C

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def inverse_affine_matrix(angle, translate, scale, shear):
    """torchvision.transforms.functional._get_inverse_affine_matrix for TF.affine on tensors: center (0, 0), shear (s, 0)."""
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


class SampleParams:
    """The random draws of one training sample, in the order the reference makes them."""
    __slots__ = ("order", "brightness", "contrast", "saturation", "hue", "sigma", "hflip", "vflip", "affine")


class GpuInputPipeline:
    """``pipe(images, masks) -> (x [B,3,224,224] float32, y [B,1,224,224] float32)`` on the GPU.

    images / masks: sequences of decoded 8-bit arrays ([h, w, 3] RGB; masks [h, w] or [h, w, 3]), any sizes.
    train=False is the evaluation transform (resize + ToTensor + Normalize; masks are returned resized only if given)."""

    def __init__(self, train=True, hflip=True, vflip=True, affine=True, brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01,
                 blur_sigma=(0.001, 2.0), device=None, task="seg"):
        """task="depth": the transforms of Depth_estimation/Data/dataset.py:43-70 + Data/dataloaders.py:55-74 instead -- sources
        padded to squares (make_square), no GaussianBlur, no affine, and `masks` are 16-bit depth maps ([h, w] uint16) resized with
        the nearest filter (Pillow 9.1.0's default for "I;16", the version the reference pins) and divided by 65535."""
        if task not in ("seg", "depth"):
            raise ValueError("task must be 'seg' or 'depth'")
        self.task = task
        if task == "depth":
            affine, blur_sigma = False, None
        self.train, self.hflip, self.vflip, self.affine = train, hflip, vflip, affine
        self.ranges = ((1 - brightness, 1 + brightness), (1 - contrast, 1 + contrast), (1 - saturation, 1 + saturation), (-hue, hue))
        self.blur_sigma = blur_sigma
        self.device = None if device is None else torch.device(device)   # None: the current CUDA device at call time
        self._mean = (ctypes.c_float * 3)(*IMAGENET_MEAN)
        self._std = (ctypes.c_float * 3)(*IMAGENET_STD)

    # ------------------------------------------------------------------------------------------ random parameters
    def draw(self):
        """One sample's parameters.  ColorJitter.get_params and GaussianBlur.get_params consume torch's global generator
        (randperm(4), four uniform_ draws, one uniform_ draw); the flips and the affine consume ``random`` (dataset.py:46-60)."""
        p = SampleParams()
        p.order = torch.randperm(4).tolist()
        p.brightness, p.contrast, p.saturation, p.hue = (float(torch.empty(1).uniform_(lo, hi)) for lo, hi in self.ranges)
        p.sigma = torch.empty(1).uniform_(self.blur_sigma[0], self.blur_sigma[1]).item() if self.blur_sigma is not None else None
        p.hflip = bool(self.hflip and random.uniform(0.0, 1.0) > 0.5)
        p.vflip = bool(self.vflip and random.uniform(0.0, 1.0) > 0.5)
        p.affine = None
        if self.affine:
            angle = random.uniform(-180.0, 180.0)
            h_trans = random.uniform(-224 / 8, 224 / 8)
            v_trans = random.uniform(-224 / 8, 224 / 8)
            scale = random.uniform(0.5, 1.5)
            shear = random.uniform(-22.5, 22.5)
            p.affine = (angle, (h_trans, v_trans), scale, shear)
        return p

    # ------------------------------------------------------------------------------------------ device stages
    def _resize(self, arrays, channels, square=False):
        return self._resize_dev(*self._pack(arrays, channels, square=square))

    def _depth_target(self, arrays):
        src, _, desc_d, B, _ = self._pack(arrays, 1, dtype=np.uint16, square=True)
        out = torch.empty((B, 1, 224, 224), dtype=torch.float32, device=self.device)
        _C.check(_C.lib().s4g_aug_depth_target(_C.ptr(src), _C.ptr(desc_d), _C.ptr(out), B, _C.stream_ptr()))
        return out

    def _pack(self, arrays, channels, dtype=np.uint8, square=False):
        """Host side: one pinned buffer with every image's bytes + the descriptor table, copied asynchronously.  square: every image
        is pasted centred on a zero square of side max(h, w) while it is packed (Depth_estimation/Data/dataset.py:9-16)."""
        B = len(arrays)
        item = np.dtype(dtype).itemsize
        desc = np.zeros((B, 5), dtype=np.int64)
        off = tmp = 0
        for i, a in enumerate(arrays):
            if a.dtype != dtype or a.ndim != (3 if channels == 3 else 2) or (channels == 3 and a.shape[2] != 3):
                raise ValueError("images must be uint8 [h, w, 3]; masks uint8 [h, w] or [h, w, 3] (depth maps uint16 [h, w]), alike within a batch")
            h, w = (max(a.shape[:2]),) * 2 if square else a.shape[:2]
            desc[i] = (off, h, w, channels, tmp)
            off += (h * w * channels * item + 15) // 16 * 16
            tmp += (h * 224 * channels + 15) // 16 * 16
        host = torch.empty(off, dtype=torch.uint8).pin_memory()
        hv = host.numpy()
        for i, a in enumerate(arrays):
            o, h, w = int(desc[i, 0]), int(desc[i, 1]), int(desc[i, 2])
            view = hv[o:o + h * w * channels * item].view(dtype).reshape((h, w, channels) if channels == 3 else (h, w))
            if square:
                view[...] = 0
                oy, ox = int((h - a.shape[0]) / 2), int((w - a.shape[1]) / 2)
                view[oy:oy + a.shape[0], ox:ox + a.shape[1]] = a
            else:
                view[...] = a
        src = host.to(self.device, non_blocking=True)
        desc_t = torch.from_numpy(desc)
        return src, desc_t, desc_t.to(self.device), B, channels

    def _resize_dev(self, src, desc_t, desc_d, B, channels):
        lib = _C.lib()
        need = lib.s4g_aug_resize_scratch_bytes(ctypes.c_void_p(desc_t.data_ptr()), B, channels)
        scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
        out = torch.empty((B, 224, 224, channels), dtype=torch.uint8, device=self.device)
        _C.check(lib.s4g_aug_resize(_C.ptr(src), _C.ptr(desc_d), ctypes.c_void_p(desc_t.data_ptr()), B, channels, _C.ptr(scratch), need,
                                    _C.ptr(out), _C.stream_ptr()))
        return out

    def _jitter(self, x8, params):
        rec = np.zeros((len(params), 8), dtype=np.int32)
        for i, p in enumerate(params):
            rec[i, :4] = p.order
            rec[i, 4] = int(np.uint8(np.int64(p.hue * 255)))                      # _functional_pil.adjust_hue
            rec[i, 5:8] = np.array([p.brightness, p.contrast, p.saturation], dtype=np.float32).view(np.int32)
        d = torch.from_numpy(rec).to(self.device)
        _C.check(_C.lib().s4g_aug_color_jitter(_C.ptr(x8), _C.ptr(d), len(params), _C.stream_ptr()))

    def _blur_normalize(self, x8, sigma):
        B = x8.shape[0]
        out = torch.empty((B, 3, 224, 224), dtype=torch.float32, device=self.device)
        sg = None if sigma is None else torch.tensor(sigma, dtype=torch.float32).to(self.device)
        _C.check(_C.lib().s4g_aug_blur_normalize(_C.ptr(x8), _C.ptr(sg), _C.ptr(out), self._mean, self._std, B, _C.stream_ptr()))
        return out

    def _mask_to_tensor(self, y8):
        B, C = y8.shape[0], y8.shape[3]
        out = torch.empty((B, 1, 224, 224), dtype=torch.float32, device=self.device)
        _C.check(_C.lib().s4g_aug_mask_to_tensor(_C.ptr(y8), _C.ptr(out), B, C, _C.stream_ptr()))
        return out

    def _geometric(self, x, y, params):
        rec = np.zeros((len(params), 9), dtype=np.int32)
        for i, p in enumerate(params):
            rec[i, 0], rec[i, 1], rec[i, 2] = int(p.hflip), int(p.vflip), int(p.affine is not None)
            if p.affine is not None:
                rec[i, 3:9] = np.array(inverse_affine_matrix(*p.affine), dtype=np.float32).view(np.int32)
        d = torch.from_numpy(rec).to(self.device)
        xo = torch.empty_like(x)
        yo = torch.empty_like(y) if y is not None else None
        _C.check(_C.lib().s4g_aug_geometric(_C.ptr(x), _C.ptr(y), _C.ptr(xo), _C.ptr(yo), _C.ptr(d), -1.0, 0.0, len(params), _C.stream_ptr()))
        return xo, yo

    # ------------------------------------------------------------------------------------------ the transform
    def __call__(self, images, masks=None, params=None, stages=None):
        """params: optional list of SampleParams (default: drawn here, one per sample, in batch order).  stages: optional
        dict that receives the intermediate tensors ("resized", "jittered", "blurred", "mask") for inspection."""
        if self.device is None:
            if not torch.cuda.is_available():
                raise RuntimeError("ssl4gie_b200: the input pipeline runs on a CUDA device — there is no CPU path")
            self.device = torch.device("cuda", torch.cuda.current_device())
        if self.device.type != "cuda":
            raise RuntimeError("ssl4gie_b200: the input pipeline runs on a CUDA device — there is no CPU path")
        _C.init(self.device.index or 0)
        B = len(images)
        with torch.cuda.device(self.device):
            depth = self.task == "depth"
            x8 = self._resize(list(images), 3, square=depth)
            y = None
            if masks is not None and depth:
                y = self._depth_target(list(masks))
            elif masks is not None:
                masks = list(masks)
                y8 = self._resize(masks, 3 if masks[0].ndim == 3 else 1)
                y = self._mask_to_tensor(y8)
            if stages is not None:
                stages["resized"] = x8.clone()
                stages["mask"] = y
            if not self.train:
                return self._blur_normalize(x8, None), y
            if params is None:
                params = [self.draw() for _ in range(B)]
            self._jitter(x8, params)
            if stages is not None:
                stages["jittered"] = x8.clone()
            x = self._blur_normalize(x8, [p.sigma for p in params] if self.blur_sigma is not None else None)
            if stages is not None:
                stages["blurred"] = x
            return self._geometric(x, y, params)
