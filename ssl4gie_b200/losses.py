"""Drop-in loss modules: same constructors / forward signatures as the reference, fused CUDA kernels underneath."""
import torch
import torch.nn as nn

from . import _C


class _DiceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, smooth):
        _C.require_cuda(logits, "logits")
        lg = logits.detach().float().contiguous()
        tg = targets.detach().float().contiguous()
        B = lg.shape[0]
        HW = lg.numel() // B
        scratch = torch.empty(_C.lib().s4g_dice_scratch_bytes(B), dtype=torch.uint8, device=lg.device)
        loss = torch.empty((), dtype=torch.float32, device=lg.device)
        _C.check(_C.lib().s4g_dice_fwd(_C.ptr(lg), _C.ptr(tg), _C.ptr(scratch), _C.ptr(loss), B, HW, smooth, _C.stream_ptr()))
        ctx.save_for_backward(lg, tg, scratch)
        ctx.smooth = smooth
        ctx.shape = logits.shape
        ctx.dtype = logits.dtype
        return loss

    @staticmethod
    def backward(ctx, gout):
        lg, tg, scratch = ctx.saved_tensors
        B = lg.shape[0]
        g = gout.detach().float().contiguous()
        d = torch.empty_like(lg)
        _C.check(_C.lib().s4g_dice_bwd(_C.ptr(lg), _C.ptr(tg), _C.ptr(scratch), _C.ptr(g), _C.ptr(d), B, lg.numel() // B,
                                       ctx.smooth, _C.stream_ptr()))
        return d.view(ctx.shape).to(ctx.dtype), None, None


class SoftDiceLoss(nn.Module):
    """reference: Binary_segmentation/Metrics/losses.py:5-24"""

    def __init__(self, smooth=1e-8):
        super().__init__()
        self.smooth = smooth

    def forward(self, logits, targets):
        return _DiceFn.apply(logits, targets, float(self.smooth))


class _SSIFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, prediction, target, alpha, scales, image_based, holder):
        _C.require_cuda(prediction, "prediction")
        p = prediction.detach().float().contiguous()
        t = target.detach().float().contiguous()
        B, H, W = p.shape[0], p.shape[-2], p.shape[-1]
        scratch = torch.empty(_C.lib().s4g_ssi_scratch_bytes(B), dtype=torch.uint8, device=p.device)
        loss = torch.empty((), dtype=torch.float32, device=p.device)
        ssi = torch.empty((B, H, W), dtype=torch.float32, device=p.device)
        _C.check(_C.lib().s4g_ssi_fwd(_C.ptr(p), _C.ptr(t), _C.ptr(scratch), _C.ptr(loss), _C.ptr(ssi), B, H, W, alpha, scales,
                                      int(image_based), _C.stream_ptr()))
        holder.append(ssi)
        ctx.save_for_backward(p, t, scratch)
        ctx.cfg = (alpha, scales, prediction.shape, prediction.dtype)
        return loss

    @staticmethod
    def backward(ctx, gout):
        p, t, scratch = ctx.saved_tensors
        alpha, scales, shape, dtype = ctx.cfg
        B, H, W = p.shape[0], p.shape[-2], p.shape[-1]
        g = gout.detach().float().contiguous()
        d = torch.empty_like(p)
        _C.check(_C.lib().s4g_ssi_bwd(_C.ptr(p), _C.ptr(t), _C.ptr(scratch), _C.ptr(g), _C.ptr(d), B, H, W, alpha, scales,
                                      _C.stream_ptr()))
        return d.view(shape).to(dtype), None, None, None, None, None


class ScaleAndShiftInvariantLoss(nn.Module):
    """reference: Depth_estimation/Metrics/losses.py:120-151.  ``reduction``: "batch-based" (:28-37, what the training
    scripts use) or anything else = image-based, exactly as the reference's constructors select it (:82-85, :99-102);
    ``scales`` up to 8 (224 px: the reference's ``[::2**k]`` sub-sampling is meaningful up to 7)."""

    def __init__(self, alpha=0.5, scales=4, reduction="batch-based"):
        super().__init__()
        if not 0 <= int(scales) <= 8:
            raise ValueError("ScaleAndShiftInvariantLoss: 0 <= scales <= 8")
        self.__alpha = alpha
        self.__scales = scales
        self.__image_based = reduction != "batch-based"
        self.__prediction_ssi = None

    def forward(self, prediction, target):
        holder = []
        loss = _SSIFn.apply(prediction, target, float(self.__alpha), int(self.__scales), self.__image_based, holder)
        self.__prediction_ssi = holder[0]
        return loss

    def __get_prediction_ssi(self):
        return self.__prediction_ssi

    prediction_ssi = property(__get_prediction_ssi)


class _CEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels, weight):
        _C.require_cuda(logits, "logits")
        lg = logits.detach().float().contiguous()
        B, N = lg.shape
        lab = labels.detach().to(device=lg.device, dtype=torch.int64).contiguous()
        w = weight.detach().to(device=lg.device, dtype=torch.float32).contiguous() if weight is not None else None
        loss = torch.empty((), dtype=torch.float32, device=lg.device)
        unit = torch.empty_like(lg)
        _C.check(_C.lib().s4g_ce_fwd(_C.ptr(lg), _C.ptr(lab), _C.ptr(w), _C.ptr(loss), _C.ptr(unit), B, N, _C.stream_ptr()))
        ctx.save_for_backward(unit)
        ctx.cfg = (logits.shape, logits.dtype)
        return loss

    @staticmethod
    def backward(ctx, gout):
        (unit,) = ctx.saved_tensors
        shape, dtype = ctx.cfg
        g = gout.detach().float().contiguous()
        d = torch.empty_like(unit)
        _C.check(_C.lib().s4g_ce_bwd(_C.ptr(unit), _C.ptr(g), _C.ptr(d), unit.shape[0], unit.shape[1], _C.stream_ptr()))
        return d.view(shape).to(dtype), None, None


class CrossEntropyLoss(nn.Module):
    """The loss of the classification fine-tune, nn.CrossEntropyLoss(torch.tensor(class_weights))
    (reference Classification/train_classification.py:157,278): weighted mean over the batch, integer class labels."""

    def __init__(self, weight=None):
        super().__init__()
        self.register_buffer("weight", None if weight is None else torch.as_tensor(weight, dtype=torch.float32))

    def forward(self, logits, labels):
        if logits.dim() != 2 or labels.dim() != 1:
            raise ValueError("CrossEntropyLoss: logits [B, num_classes] and integer labels [B] expected")
        return _CEFn.apply(logits, labels, self.weight)
