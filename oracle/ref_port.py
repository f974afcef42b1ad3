"""ORACLE — test infrastructure, NOT product code.

A plain-PyTorch (ATen, CPU or CUDA) restatement of the SSL4GIE hot path:
ViT-B/16 encoder + DPT decoder forward, SoftDice and scale-and-shift-invariant
losses, and the reference training step.  Backward is torch autograd over this
restatement, exactly as in the reference.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import this file, and only as the checker / the
timed CPU baseline.  The product package ``ssl4gie_b200`` never imports it.

Pinning: the reference ships no tests or golden vectors, so
reference-held pins do not exist ("parity unpinned" in that sense).  Instead
``oracle/make_golden.py`` imports the UNMODIFIED reference modules from
/root/reference (through ``oracle/timm_shim``) in the build container and
(a) checks this restatement against them, (b) writes the committed fixtures in
``tests/golden/``.  ``tests/test_oracle.py`` re-checks this file against those
fixtures on any machine.

Third-party arithmetic not in /root/reference: ``timm==0.6.12``
(requirements.txt:7): PatchEmbed / Attention / Mlp / Block semantics are
restated from its published behaviour; anchored on the reference call sites
Models/models.py:326-338, :427-456, :545-558.

Every function cites the reference file:line it follows (paths relative to
/root/reference).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

TAPS = (2, 5, 8, 11)  # Models/models.py:335, :453, :555


# --------------------------------------------------------------------------- encoder
def patch_embed(sd, x):
    """timm PatchEmbed called at Models/models.py:326,428,546: Conv2d k16 s16, flatten(2).transpose(1,2)."""
    y = F.conv2d(x, sd["patch_embed.proj.weight"], sd["patch_embed.proj.bias"], stride=16)
    return y.flatten(2).transpose(1, 2)


def pos_embed(sd, x):
    """timm _pos_embed (Models/models.py:330,550) == explicit form at Models/models.py:445-448."""
    cls = sd["cls_token"].expand(x.shape[0], -1, -1)
    return torch.cat((cls, x), dim=1) + sd["pos_embed"]


def attention(sd, p, x, num_heads=12):
    """timm 0.6.12 Attention.forward (no SDPA; materialises the score matrix)."""
    B, N, C = x.shape
    qkv = F.linear(x, sd[p + "qkv.weight"], sd[p + "qkv.bias"])
    qkv = qkv.reshape(B, N, 3, num_heads, C // num_heads).permute(2, 0, 3, 1, 4)
    q, k, v = qkv.unbind(0)
    attn = (q @ k.transpose(-2, -1)) * ((C // num_heads) ** -0.5)
    attn = attn.softmax(dim=-1)
    y = (attn @ v).transpose(1, 2).reshape(B, N, C)
    return F.linear(y, sd[p + "proj.weight"], sd[p + "proj.bias"])


def block(sd, i, x, prefix="blocks.", num_heads=12):
    """timm 0.6.12 Block.forward: pre-LN residual x2, LN eps 1e-6 (Models/models.py:384,498), exact GELU."""
    p = f"{prefix}{i}."
    h = F.layer_norm(x, (x.shape[-1],), sd[p + "norm1.weight"], sd[p + "norm1.bias"], 1e-6)
    x = x + attention(sd, p + "attn.", h, num_heads)
    h = F.layer_norm(x, (x.shape[-1],), sd[p + "norm2.weight"], sd[p + "norm2.bias"], 1e-6)
    h = F.linear(h, sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"])
    h = F.gelu(h)
    h = F.linear(h, sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"])
    return x + h


def encoder(sd, x, dense=True, depth=12):
    """forward_features, Models/models.py:325-338 (== :427-456, :545-558)."""
    x = pos_embed(sd, patch_embed(sd, x))
    z = []
    for i in range(depth):
        x = block(sd, i, x)
        if dense and i in TAPS:
            z.append(x)
    if dense:
        return z
    return F.layer_norm(x, (x.shape[-1],), sd["norm.weight"], sd["norm.bias"], 1e-6)


# --------------------------------------------------------------------------- decoder
def _bn(sd, p, x, training, stats_out=None):
    """nn.BatchNorm2d(256), eps 1e-5, momentum 0.1 (Models/DPT_decoder.py:204-206,492)."""
    if training:
        y = F.batch_norm(x, None, None, sd[p + "weight"], sd[p + "bias"], True, 0.1, 1e-5)
        if stats_out is not None:
            n = x.numel() // x.shape[1]
            mean = x.float().mean((0, 2, 3))
            var = x.float().var((0, 2, 3), unbiased=False)
            stats_out[p] = (mean.detach(), (var * n / max(n - 1, 1)).detach())
        return y
    return F.batch_norm(x, sd[p + "running_mean"], sd[p + "running_var"], sd[p + "weight"], sd[p + "bias"],
                        False, 0.1, 1e-5)


def rcu(sd, p, x, bn, training, stats_out=None):
    """ResidualConvUnit_custom.forward, Models/DPT_decoder.py:212-233."""
    out = F.relu(x)
    out = F.conv2d(out, sd[p + "conv1.weight"], sd.get(p + "conv1.bias"), padding=1)
    if bn:
        out = _bn(sd, p + "bn1.", out, training, stats_out)
    out = F.relu(out)
    out = F.conv2d(out, sd[p + "conv2.weight"], sd.get(p + "conv2.bias"), padding=1)
    if bn:
        out = _bn(sd, p + "bn2.", out, training, stats_out)
    return out + x


def fusion(sd, p, xs, bn, training, stats_out=None):
    """FeatureFusionBlock_custom.forward, Models/DPT_decoder.py:281-301."""
    out = xs[0]
    if len(xs) == 2:
        out = out + rcu(sd, p + "resConfUnit1.", xs[1], bn, training, stats_out)
    out = rcu(sd, p + "resConfUnit2.", out, bn, training, stats_out)
    out = F.interpolate(out, scale_factor=2, mode="bilinear", align_corners=True)
    return F.conv2d(out, sd[p + "out_conv.weight"], sd[p + "out_conv.bias"])


def decoder(sd, z, dense="seg", training=True, dropout_mask=None, stats_out=None, p="decoder."):
    """DPT_decoder.forward_skip + forward, Models/DPT_decoder.py:501-539.

    ``dropout_mask`` ([B,256,112,112] of 0/1) replaces the RNG of nn.Dropout(0.1) (:494); None => p=0.
    """
    B = z[0].shape[0]
    maps = [t[:, 1:].transpose(1, 2).unflatten(2, (14, 14)).contiguous() for t in z]  # :503-515
    l1 = F.conv2d(maps[0], sd[p + "act_postprocess12.0.weight"], sd[p + "act_postprocess12.0.bias"])
    l1 = F.conv_transpose2d(l1, sd[p + "act_postprocess12.1.weight"], sd[p + "act_postprocess12.1.bias"], stride=4)
    l2 = F.conv2d(maps[1], sd[p + "act_postprocess22.0.weight"], sd[p + "act_postprocess22.0.bias"])
    l2 = F.conv_transpose2d(l2, sd[p + "act_postprocess22.1.weight"], sd[p + "act_postprocess22.1.bias"], stride=2)
    l3 = F.conv2d(maps[2], sd[p + "act_postprocess32.0.weight"], sd[p + "act_postprocess32.0.bias"])
    l4 = F.conv2d(maps[3], sd[p + "act_postprocess42.0.weight"], sd[p + "act_postprocess42.0.bias"])
    l4 = F.conv2d(l4, sd[p + "act_postprocess42.1.weight"], sd[p + "act_postprocess42.1.bias"], stride=2, padding=1)
    l1 = F.conv2d(l1, sd[p + "layer1_rn.weight"], None, padding=1)  # :522-525
    l2 = F.conv2d(l2, sd[p + "layer2_rn.weight"], None, padding=1)
    l3 = F.conv2d(l3, sd[p + "layer3_rn.weight"], None, padding=1)
    l4 = F.conv2d(l4, sd[p + "layer4_rn.weight"], None, padding=1)
    bn = dense == "seg"  # :461
    path4 = fusion(sd, p + "refinenet4.", [l4], bn, training, stats_out)  # :532
    path3 = fusion(sd, p + "refinenet3.", [path4, l3], bn, training, stats_out)
    path2 = fusion(sd, p + "refinenet2.", [path3, l2], bn, training, stats_out)
    path1 = fusion(sd, p + "refinenet1.", [path2, l1], bn, training, stats_out)
    if dense == "depth":  # :469-482
        o = F.conv2d(path1, sd[p + "output_conv.0.weight"], sd[p + "output_conv.0.bias"], padding=1)
        o = F.interpolate(o, scale_factor=2, mode="bilinear", align_corners=True)
        o = F.conv2d(o, sd[p + "output_conv.2.weight"], sd[p + "output_conv.2.bias"], padding=1)
        o = F.relu(o)
        o = F.conv2d(o, sd[p + "output_conv.4.weight"], sd[p + "output_conv.4.bias"])
        return torch.sigmoid(o)
    o = F.conv2d(path1, sd[p + "output_conv.0.weight"], None, padding=1)  # :484-497
    o = _bn(sd, p + "output_conv.1.", o, training, stats_out)
    o = F.relu(o)
    if training and dropout_mask is not None:
        o = o * dropout_mask.to(o.dtype) * (1.0 / 0.9)
    o = F.conv2d(o, sd[p + "output_conv.4.weight"], sd[p + "output_conv.4.bias"])
    return F.interpolate(o, scale_factor=2, mode="bilinear", align_corners=True)


def forward(sd, x, dense="seg", training=True, dropout_mask=None, stats_out=None, out_token="cls"):
    """VisionTransformer_from_Any.forward, Models/models.py:340-357."""
    if dense:
        return decoder(sd, encoder(sd, x, True), dense, training, dropout_mask, stats_out)
    t = encoder(sd, x, False)
    t = t[:, 0] if out_token == "cls" else t[:, 1:].mean(1)
    if "lin_head.weight" in sd:
        t = F.linear(t, sd["lin_head.weight"], sd["lin_head.bias"])
    return t


# --------------------------------------------------------------------------- losses
def soft_dice_loss(logits, targets, smooth=1e-8):
    """SoftDiceLoss.forward, Binary_segmentation/Metrics/losses.py:10-24."""
    num = targets.size(0)
    m1 = torch.sigmoid(logits).view(num, -1)
    m2 = targets.view(num, -1)
    score = 2.0 * ((m1 * m2).sum(1) + smooth) / ((m1 * m1).sum(1) + (m2 * m2).sum(1) + smooth)
    return 1 - score.sum() / num


def compute_scale_and_shift(prediction, target, mask):
    """Depth_estimation/Metrics/losses.py:5-25."""
    a_00 = torch.sum(mask * prediction * prediction, (1, 2))
    a_01 = torch.sum(mask * prediction, (1, 2))
    a_11 = torch.sum(mask, (1, 2))
    b_0 = torch.sum(mask * prediction * target, (1, 2))
    b_1 = torch.sum(mask * target, (1, 2))
    x_0 = torch.zeros_like(b_0)
    x_1 = torch.zeros_like(b_1)
    det = a_00 * a_11 - a_01 * a_01
    valid = det.nonzero()
    x_0[valid] = (a_11[valid] * b_0[valid] - a_01[valid] * b_1[valid]) / det[valid]
    x_1[valid] = (-a_01[valid] * b_0[valid] + a_00[valid] * b_1[valid]) / det[valid]
    return x_0, x_1


def _reduction_batch_based(image_loss, M):
    """Depth_estimation/Metrics/losses.py:28-37."""
    divisor = torch.sum(M)
    if divisor == 0:
        return 0
    return torch.sum(image_loss) / divisor


def _reduction_image_based(image_loss, M):
    """Depth_estimation/Metrics/losses.py:40-48: mean over images of (image term / image mask count)."""
    valid = M.nonzero()
    image_loss = image_loss.clone()
    image_loss[valid] = image_loss[valid] / M[valid]
    return torch.mean(image_loss)


def _gradient_loss(prediction, target, mask, reduction=_reduction_batch_based):
    """Depth_estimation/Metrics/losses.py:60-77."""
    M = torch.sum(mask, (1, 2))
    diff = torch.mul(mask, prediction - target)
    grad_x = torch.abs(diff[:, :, 1:] - diff[:, :, :-1])
    grad_x = torch.mul(torch.mul(mask[:, :, 1:], mask[:, :, :-1]), grad_x)
    grad_y = torch.abs(diff[:, 1:, :] - diff[:, :-1, :])
    grad_y = torch.mul(torch.mul(mask[:, 1:, :], mask[:, :-1, :]), grad_y)
    image_loss = torch.sum(grad_x, (1, 2)) + torch.sum(grad_y, (1, 2))
    return reduction(image_loss, M)


def ssi_loss(prediction, target, alpha=0.5, scales=4, return_ssi=False, reduction="batch-based"):
    """ScaleAndShiftInvariantLoss.forward, Depth_estimation/Metrics/losses.py:130-146; `reduction` selects the
    reduction function the way MSELoss / GradientLoss do (:82-85, :99-102)."""
    red = _reduction_batch_based if reduction == "batch-based" else _reduction_image_based
    prediction = prediction.squeeze(1)
    target = target.squeeze(1)
    mask = target > 0
    scale, shift = compute_scale_and_shift(prediction, target, mask)
    ssi = scale.view(-1, 1, 1) * prediction + shift.view(-1, 1, 1)
    M = torch.sum(mask, (1, 2))
    res = ssi - target
    total = red(torch.sum(mask * res * res, (1, 2)), 2 * M)  # mse_loss :51-57
    if alpha > 0:
        reg = 0
        for s in range(scales):  # GradientLoss.forward :104-117
            step = 2 ** s
            reg = reg + _gradient_loss(ssi[:, ::step, ::step], target[:, ::step, ::step], mask[:, ::step, ::step], red)
        total = total + alpha * reg
    return (total, ssi) if return_ssi else total


# --------------------------------------------------------------------------- training step
def make_state(dense="seg", seed=42, head_classes=None, device="cpu"):
    """Random-init state_dict with the reference's names/shapes (SURVEY.md §8b manifest).

    Initialisers follow torch/timm defaults closely enough for a *workload*; parity tests never
    depend on the init law because both sides consume the same tensors.
    """
    g = torch.Generator().manual_seed(seed)

    def tn(*shape, std=0.02):
        return torch.nn.init.trunc_normal_(torch.empty(*shape), std=std, generator=g)

    def ku(*shape):  # kaiming-uniform(a=sqrt(5)) as nn.Conv2d/Linear default
        fan_in = shape[1] * (shape[2] * shape[3] if len(shape) == 4 else 1)
        b = 1.0 / math.sqrt(fan_in)
        return (torch.rand(*shape, generator=g) * 2 - 1) * b

    sd = {"cls_token": torch.randn(1, 1, 768, generator=g) * 1e-6, "pos_embed": tn(1, 197, 768)}
    sd["patch_embed.proj.weight"] = ku(768, 3, 16, 16)
    sd["patch_embed.proj.bias"] = ku(768, 768, 1, 1)[:, 0, 0, 0].clone()
    for i in range(12):
        p = f"blocks.{i}."
        for n in ("norm1", "norm2"):
            sd[p + n + ".weight"] = torch.ones(768)
            sd[p + n + ".bias"] = torch.zeros(768)
        for n, (o, c) in {"attn.qkv": (2304, 768), "attn.proj": (768, 768), "mlp.fc1": (3072, 768),
                          "mlp.fc2": (768, 3072)}.items():
            sd[p + n + ".weight"] = tn(o, c)
            sd[p + n + ".bias"] = torch.zeros(o)
    sd["norm.weight"] = torch.ones(768)
    sd["norm.bias"] = torch.zeros(768)
    if head_classes:
        sd["lin_head.weight"] = ku(head_classes, 768)
        sd["lin_head.bias"] = torch.zeros(head_classes)
    if dense:
        d = "decoder."
        feats = [96, 192, 384, 768]

        def conv(name, o, c, k, bias=True, transpose=False):
            sd[d + name + ".weight"] = ku(c, o, k, k) if transpose else ku(o, c, k, k)
            if bias:
                sd[d + name + ".bias"] = (torch.rand(o, generator=g) * 2 - 1) / math.sqrt(c * k * k)

        conv("act_postprocess12.0", 96, 768, 1)
        conv("act_postprocess12.1", 96, 96, 4, transpose=True)
        conv("act_postprocess22.0", 192, 768, 1)
        conv("act_postprocess22.1", 192, 192, 2, transpose=True)
        conv("act_postprocess32.0", 384, 768, 1)
        conv("act_postprocess42.0", 768, 768, 1)
        conv("act_postprocess42.1", 768, 768, 3)
        for k, c in enumerate(feats):
            conv(f"layer{k + 1}_rn", 256, c, 3, bias=False)
        bn = dense == "seg"

        def bnp(name):
            sd[d + name + ".weight"] = torch.ones(256)
            sd[d + name + ".bias"] = torch.zeros(256)
            sd[d + name + ".running_mean"] = torch.zeros(256)
            sd[d + name + ".running_var"] = torch.ones(256)
            sd[d + name + ".num_batches_tracked"] = torch.zeros((), dtype=torch.int64)

        for r in (1, 2, 3, 4):
            conv(f"refinenet{r}.out_conv", 256, 256, 1)
            for u in (1, 2):
                for c in (1, 2):
                    conv(f"refinenet{r}.resConfUnit{u}.conv{c}", 256, 256, 3, bias=not bn)
                if bn:
                    bnp(f"refinenet{r}.resConfUnit{u}.bn1")
                    bnp(f"refinenet{r}.resConfUnit{u}.bn2")
        if bn:
            conv("output_conv.0", 256, 256, 3, bias=False)
            bnp("output_conv.1")
            conv("output_conv.4", 1, 256, 1)
        else:
            conv("output_conv.0", 128, 256, 3)
            conv("output_conv.2", 32, 128, 3)
            conv("output_conv.4", 1, 32, 1)
    return {k: v.to(device) for k, v in sd.items()}


NON_TRAINABLE_SUFFIXES = ("running_mean", "running_var", "num_batches_tracked")


def trainable(sd):
    return {k: v for k, v in sd.items() if not k.endswith(NON_TRAINABLE_SUFFIXES)}


def weighted_ce_loss(logits, labels, class_weights=None):
    """nn.CrossEntropyLoss(torch.tensor(class_weights)) as built at Classification/train_classification.py:278:
    sum_i w[y_i] * (-log softmax(x_i)[y_i]) / sum_i w[y_i]."""
    logp = torch.log_softmax(logits.float(), dim=1)
    nll = -logp.gather(1, labels.view(-1, 1)).squeeze(1)
    if class_weights is None:
        return nll.mean()
    w = class_weights.to(nll.dtype)[labels]
    return (w * nll).sum() / w.sum()


def loss_and_grads(sd, x, target, dense="seg", dropout_mask=None, alpha=0.1, stats_out=None, out_token="cls",
                   class_weights=None):
    """One reference fwd+loss+bwd (Binary_segmentation/train_segmentation.py:44-47 without AMP scaling; dense=None:
    Classification/train_classification.py:44-47 with integer class labels as target)."""
    params = {k: v.detach().clone().requires_grad_(True) for k, v in trainable(sd).items()}
    full = dict(sd)
    full.update(params)
    out = forward(full, x, dense, True, dropout_mask, stats_out, out_token)
    if not dense:
        loss = weighted_ce_loss(out, target, class_weights)
        loss.backward()
        return out.detach(), loss.detach(), {k: v.grad for k, v in params.items()}
    loss = soft_dice_loss(out, target) if dense == "seg" else ssi_loss(out, target, alpha)
    loss.backward()
    grads = {k: v.grad for k, v in params.items()}
    return out.detach(), loss.detach(), grads


def adamw_step(p, g, m, v, step, lr=1e-4, b1=0.9, b2=0.999, eps=1e-8, wd=0.01):
    """torch.optim.AdamW single-tensor math (train_segmentation.py:204 defaults)."""
    p.mul_(1 - lr * wd)
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


# --------------------------------------------------------------------------- MAE pretraining step (BASELINE config 5)
def sincos_2d(embed_dim, grid_size=14):
    """Models/mae/util/pos_embed.py:20-69 (get_2d_sincos_pos_embed, cls_token=True): w-coordinate first, then h."""
    import numpy as np

    def one(d, pos):
        omega = 1.0 / 10000 ** (np.arange(d // 2, dtype=float) / (d / 2.0))
        out = np.einsum("m,d->md", pos.reshape(-1), omega)
        return np.concatenate([np.sin(out), np.cos(out)], axis=1)

    gh, gw = np.arange(grid_size, dtype=np.float32), np.arange(grid_size, dtype=np.float32)
    grid = np.stack(np.meshgrid(gw, gh), axis=0).reshape(2, 1, grid_size, grid_size)
    emb = np.concatenate([one(embed_dim // 2, grid[0]), one(embed_dim // 2, grid[1])], axis=1)
    return torch.from_numpy(np.concatenate([np.zeros([1, embed_dim]), emb], axis=0)).float().unsqueeze(0)


def mae_make_state(seed=42, device="cpu"):
    """state_dict of mae_vit_base_patch16 (Models/mae/models_mae.py:26-59,223-228): names / shapes as the module's,
    initial values by the laws of initialize_weights (:63-93) from a seeded generator (not the module's RNG stream)."""
    g = torch.Generator().manual_seed(seed)

    def xavier(o, c):
        b = math.sqrt(6.0 / (o + c))
        return (torch.rand(o, c, generator=g) * 2 - 1) * b

    sd = {"cls_token": torch.randn(1, 1, 768, generator=g) * 0.02, "pos_embed": sincos_2d(768),
          "mask_token": torch.randn(1, 1, 512, generator=g) * 0.02, "decoder_pos_embed": sincos_2d(512)}
    sd["patch_embed.proj.weight"] = xavier(768, 768).reshape(768, 3, 16, 16)
    sd["patch_embed.proj.bias"] = (torch.rand(768, generator=g) * 2 - 1) / math.sqrt(768)

    def blocks(prefix, depth, C):
        for i in range(depth):
            p = f"{prefix}{i}."
            for n in ("norm1", "norm2"):
                sd[p + n + ".weight"] = torch.ones(C)
                sd[p + n + ".bias"] = torch.zeros(C)
            for n, (o, c) in {"attn.qkv": (3 * C, C), "attn.proj": (C, C), "mlp.fc1": (4 * C, C), "mlp.fc2": (C, 4 * C)}.items():
                sd[p + n + ".weight"] = xavier(o, c)
                sd[p + n + ".bias"] = torch.zeros(o)

    blocks("blocks.", 12, 768)
    sd["norm.weight"], sd["norm.bias"] = torch.ones(768), torch.zeros(768)
    sd["decoder_embed.weight"], sd["decoder_embed.bias"] = xavier(512, 768), torch.zeros(512)
    blocks("decoder_blocks.", 8, 512)
    sd["decoder_norm.weight"], sd["decoder_norm.bias"] = torch.ones(512), torch.zeros(512)
    sd["decoder_pred.weight"], sd["decoder_pred.bias"] = xavier(768, 512), torch.zeros(768)
    return {k: v.to(device) for k, v in sd.items()}


MAE_FROZEN = ("pos_embed", "decoder_pos_embed")  # requires_grad=False, models_mae.py:37,50


def mae_random_masking(x, noise, mask_ratio=0.75):
    """models_mae.py:123-148 with the uniform noise passed in (the reference draws torch.rand(N, L) at :132)."""
    N, L, D = x.shape
    len_keep = int(L * (1 - mask_ratio))
    ids_shuffle = torch.argsort(noise, dim=1)
    ids_restore = torch.argsort(ids_shuffle, dim=1)
    ids_keep = ids_shuffle[:, :len_keep]
    x_masked = torch.gather(x, 1, ids_keep.unsqueeze(-1).repeat(1, 1, D))
    mask = torch.ones([N, L], device=x.device, dtype=x.dtype)
    mask[:, :len_keep] = 0
    return x_masked, torch.gather(mask, 1, ids_restore), ids_restore


def mae_patchify(imgs, p=16):
    """models_mae.py:95-107."""
    h = w = imgs.shape[2] // p
    x = imgs.reshape(imgs.shape[0], 3, h, p, w, p)
    return torch.einsum("nchpwq->nhwpqc", x).reshape(imgs.shape[0], h * w, p * p * 3)


def mae_forward(sd, imgs, noise, mask_ratio=0.75, norm_pix_loss=True):
    """MaskedAutoencoderViT.forward (models_mae.py:216-220): forward_encoder :150-170, forward_decoder :172-196,
    forward_loss :198-214.  Returns (loss, pred [N, 196, 768], mask [N, 196])."""
    x = patch_embed(sd, imgs) + sd["pos_embed"][:, 1:, :]
    x, mask, ids_restore = mae_random_masking(x, noise, mask_ratio)
    cls = (sd["cls_token"] + sd["pos_embed"][:, :1, :]).expand(x.shape[0], -1, -1)
    x = torch.cat((cls, x), dim=1)
    for i in range(12):
        x = block(sd, i, x)
    x = F.layer_norm(x, (768,), sd["norm.weight"], sd["norm.bias"], 1e-6)
    x = F.linear(x, sd["decoder_embed.weight"], sd["decoder_embed.bias"])
    mask_tokens = sd["mask_token"].repeat(x.shape[0], ids_restore.shape[1] + 1 - x.shape[1], 1)
    x_ = torch.cat([x[:, 1:, :], mask_tokens], dim=1)
    x_ = torch.gather(x_, 1, ids_restore.unsqueeze(-1).repeat(1, 1, x.shape[2]))
    x = torch.cat([x[:, :1, :], x_], dim=1) + sd["decoder_pos_embed"]
    for i in range(8):
        x = block(sd, i, x, "decoder_blocks.", 16)
    x = F.layer_norm(x, (512,), sd["decoder_norm.weight"], sd["decoder_norm.bias"], 1e-6)
    pred = F.linear(x, sd["decoder_pred.weight"], sd["decoder_pred.bias"])[:, 1:, :]
    target = mae_patchify(imgs)
    if norm_pix_loss:
        mean = target.mean(dim=-1, keepdim=True)
        var = target.var(dim=-1, keepdim=True)
        target = (target - mean) / (var + 1.0e-6) ** 0.5
    loss = ((pred - target) ** 2).mean(dim=-1)
    return (loss * mask).sum() / mask.sum(), pred, mask


def mae_loss_and_grads(sd, imgs, noise, mask_ratio=0.75, norm_pix_loss=True):
    """One MAE step's loss and parameter gradients (engine_pretrain.py:47-58 without the scaler)."""
    params = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items() if k not in MAE_FROZEN}
    full = dict(sd)
    full.update(params)
    loss, pred, mask = mae_forward(full, imgs, noise, mask_ratio, norm_pix_loss)
    loss.backward()
    return loss.detach(), pred.detach(), mask, {k: v.grad for k, v in params.items()}


def mae_param_groups(sd, weight_decay=0.05):
    """timm optim_factory.add_weight_decay as called at Models/mae/main_pretrain.py:179: no decay for 1-D tensors and
    biases (cls_token / mask_token are 3-D and DO decay); frozen tensors are skipped."""
    decay, no_decay = [], []
    for k, v in sd.items():
        if k in MAE_FROZEN:
            continue
        (no_decay if v.dim() == 1 or k.endswith(".bias") else decay).append(k)
    return [{"names": no_decay, "weight_decay": 0.0}, {"names": decay, "weight_decay": weight_decay}]


def mae_lr(it, it_per_epoch, base_lr, warmup_epochs=40, epochs=800, min_lr=0.0):
    """Models/mae/util/lr_sched.py:9-21 called per iteration with epoch = it / len(loader) + epoch (engine_pretrain.py:42-43)."""
    e = it / it_per_epoch
    if e < warmup_epochs:
        return base_lr * e / warmup_epochs
    return min_lr + (base_lr - min_lr) * 0.5 * (1.0 + math.cos(math.pi * (e - warmup_epochs) / (epochs - warmup_epochs)))
