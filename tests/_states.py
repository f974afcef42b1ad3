"""Test helpers (checker side): synthetic targets that look like the real tasks' and a *well-conditioned* model state.

At random init the decoder's BatchNorm statistics are untrained, the seg logits sit on the flat part of the Dice loss and
the depth loss's 2x2 scale/shift solve is nearly singular: the reference's own fp32 gradients are then only 1e-1..1e-3
accurate (measured against fp64) and say nothing about a bf16 implementation.  A few dozen reference AdamW steps on a
fixed batch with blob-like targets move the model off that plateau; gradients there are well defined and the plain
north-star criterion (2e-2 relative, cosine 0.999) can be applied to every tensor.
"""
import torch

from oracle import ref_port  # checker only


def blob_masks(B, seed=0, size=224):
    """Union of 1-3 random discs per image (SURVEY.md §8d config 1 "blob variant", Kvasir-like masks)."""
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(size, dtype=torch.float32), torch.arange(size, dtype=torch.float32), indexing="ij")
    out = torch.zeros(B, 1, size, size)
    for b in range(B):
        for _ in range(int(torch.randint(1, 4, (1,), generator=g))):
            cx, cy = (torch.rand(2, generator=g) * 0.6 + 0.2) * size
            r = (torch.rand(1, generator=g) * 0.2 + 0.08) * size
            out[b, 0] = torch.maximum(out[b, 0], ((xx - cx) ** 2 + (yy - cy) ** 2 <= r * r).float())
    return out


def smooth_depth(B, seed=0, size=224):
    """Low-frequency depth maps in (0.05, 1] with a zero (invalid) border and ~3 % invalid specks, C3VD-like
    (Depth_estimation/Data/dataset.py:9-16 pads to a square with zeros)."""
    g = torch.Generator().manual_seed(seed)
    coarse = torch.rand(B, 1, 7, 7, generator=g)
    d = torch.nn.functional.interpolate(coarse, size=(size, size), mode="bicubic", align_corners=True).clamp(0.05, 1.0)
    d[torch.rand(B, 1, size, size, generator=g) < 0.03] = 0.0
    d[:, :, :12, :] = 0.0
    d[:, :, -12:, :] = 0.0
    return d


def structured_images(B, seed=0, size=224):
    """Images correlated with nothing in particular but not white noise: smooth colour fields plus noise."""
    g = torch.Generator().manual_seed(seed + 1000)
    low = torch.nn.functional.interpolate(torch.randn(B, 3, 14, 14, generator=g), size=(size, size), mode="bilinear", align_corners=False)
    return low + 0.3 * torch.randn(B, 3, size, size, generator=g)


def task_batch(dense, B, seed=0, size=224):
    """A learnable synthetic sample of the task: the target (blob mask / smooth depth) is visible in the image through
    a brightness cue under structured noise, so training has something to find and a fresh batch per step never gets
    memorised (the loss settles at a mid-range value with substantial gradients)."""
    x = structured_images(B, seed, size)
    if dense == "seg":
        t = blob_masks(B, seed, size)
        x = x + 1.2 * (t - 0.5)
    else:
        t = smooth_depth(B, seed, size)
        x = x + 1.5 * (t - 0.5) * (t > 0)
    return x, t


def warm_state(dense, sd, x, t, steps=50, lr=1e-4, alpha=0.1, fresh=None, device=None):
    """`steps` reference AdamW steps (fp32 oracle + torch.optim.AdamW, train_segmentation.py:204) on the fixed batch
    (x, t), or -- fresh = batch size -- on a new task_batch(dense, fresh, seed = 1000 + step) every step.
    Returns the state_dict (incl. BatchNorm running statistics) on the device."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    if device is None:
        device = x.device
    full = {k: v.to(device).clone() for k, v in sd.items()}
    params = {k: v.requires_grad_(True) for k, v in ref_port.trainable(full).items()}
    full.update(params)
    opt = torch.optim.AdamW(list(params.values()), lr=lr)
    bn_names = [k[:-len("running_mean")] for k in full if k.endswith("running_mean")]
    for k in range(steps):
        if fresh:
            x, t = (v.to(device) for v in task_batch(dense, fresh, 1000 + k))
        opt.zero_grad(set_to_none=True)
        stats = {}
        out = ref_port.forward(full, x, dense, True, None, stats)
        loss = ref_port.soft_dice_loss(out, t) if dense == "seg" else ref_port.ssi_loss(out, t, alpha)
        loss.backward()
        opt.step()
        with torch.no_grad():
            for p in bn_names:  # nn.BatchNorm2d momentum 0.1 running statistics, as the module would keep them
                if p in stats:
                    mean, var = stats[p]
                    full[p + "running_mean"].mul_(0.9).add_(0.1 * mean)
                    full[p + "running_var"].mul_(0.9).add_(0.1 * var)
                    full[p + "num_batches_tracked"] += 1
    return {k: v.detach().clone() for k, v in full.items()}


def warm_state_product(dense, sd, steps=30, lr=1e-4, alpha=0.1, batch=4):
    """The same warm-up run by the PRODUCT path (bf16 tensor-core schedule + FusedAdamW, fresh task_batch every step).  The
    product path is bit-reproducible, so every run of a test starts from the identical state (the oracle's cuDNN kernels
    are not); which arithmetic produced the state does not matter to a test that evaluates both sides AT that state."""
    from ssl4gie_b200 import FusedAdamW, ScaleAndShiftInvariantLoss, SoftDiceLoss, utils
    m = utils.get_ImageNet_or_random_ViT(False, 1 if dense == "seg" else None, False, dense, False, False)
    m.load_state_dict({k: v.cpu() for k, v in sd.items()}, strict=True)
    if dense == "seg":
        m.decoder.output_conv[3].p = 0.0
    m.precision, m.backend = "bf16", "tensor"
    m = m.cuda().train()
    opt = FusedAdamW(m, lr=lr)
    loss_fn = SoftDiceLoss() if dense == "seg" else ScaleAndShiftInvariantLoss(alpha=alpha)
    for k in range(steps):
        x, t = (v.cuda() for v in task_batch(dense, batch, 1000 + k))
        loss_fn(m(x), t).backward()
        opt.step()
    torch.cuda.synchronize()
    return {k: v.detach().clone() for k, v in m.state_dict().items()}
