"""MAE pretraining (BASELINE config 5): MaskedAutoencoderViT with the reference's constructor, parameter names and
forward signature (reference Models/mae/models_mae.py), the optimizer groups of Models/mae/main_pretrain.py:179-180 and the
per-iteration schedule of Models/mae/util/lr_sched.py.  forward() runs the fused CUDA schedule (engine.run_mae) and the
masked norm-pix loss kernel; there is no PyTorch-op fallback."""
import math

import torch
import torch.nn as nn

from . import _C
from .engine import run_mae
from .models import _Block, _PatchEmbed, _mae_sincos_2d


class _MaeLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, imgs, mask, norm_pix):
        _C.require_cuda(pred, "pred")
        p = pred.detach().float().contiguous()
        x = imgs.detach().float().contiguous()
        mk = mask.detach().float().contiguous()
        B = p.shape[0]
        scratch = torch.empty(_C.lib().s4g_mae_loss_scratch_bytes(B), dtype=torch.uint8, device=p.device)
        loss = torch.empty((), dtype=torch.float32, device=p.device)
        _C.check(_C.lib().s4g_mae_loss_fwd(_C.ptr(p), _C.ptr(x), _C.ptr(mk), _C.ptr(scratch), _C.ptr(loss), B, int(norm_pix),
                                           _C.stream_ptr()))
        ctx.save_for_backward(p, x, mk, scratch)
        ctx.norm_pix = int(norm_pix)
        return loss

    @staticmethod
    def backward(ctx, gout):
        p, x, mk, scratch = ctx.saved_tensors
        g = gout.detach().float().contiguous()
        d = torch.empty_like(p)
        _C.check(_C.lib().s4g_mae_loss_bwd(_C.ptr(p), _C.ptr(x), _C.ptr(mk), _C.ptr(scratch), _C.ptr(g), _C.ptr(d), p.shape[0],
                                           ctx.norm_pix, _C.stream_ptr()))
        return d, None, None, None


class MaskedAutoencoderViT(nn.Module):
    """reference Models/mae/models_mae.py:22-220.  The B200 schedule is built for mae_vit_base_patch16 (768/12/12 encoder,
    512/8/16 decoder, 224 px, patch 16) and mask_ratio 0.75 (49 visible patches)."""

    dense = "mae"  # selects S4G_DENSE_MAE in the engine

    def __init__(self, img_size=224, patch_size=16, in_chans=3, embed_dim=1024, depth=24, num_heads=16,
                 decoder_embed_dim=512, decoder_depth=8, decoder_num_heads=16, mlp_ratio=4., norm_layer=nn.LayerNorm,
                 norm_pix_loss=False):
        super().__init__()
        cfg = (img_size, patch_size, in_chans, embed_dim, depth, num_heads, decoder_embed_dim, decoder_depth, decoder_num_heads, mlp_ratio)
        if cfg != (224, 16, 3, 768, 12, 12, 512, 8, 16, 4):
            raise NotImplementedError("the B200 schedule is specialised for mae_vit_base_patch16 (dec512d8b)")
        self.patch_embed = _PatchEmbed(img_size, patch_size, in_chans, embed_dim)
        num_patches = self.patch_embed.num_patches
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, num_patches + 1, embed_dim), requires_grad=False)
        self.blocks = nn.ModuleList([_Block(embed_dim, num_heads, mlp_ratio) for _ in range(depth)])
        self.norm = nn.LayerNorm(embed_dim, eps=1e-6)
        self.decoder_embed = nn.Linear(embed_dim, decoder_embed_dim, bias=True)
        self.mask_token = nn.Parameter(torch.zeros(1, 1, decoder_embed_dim))
        self.decoder_pos_embed = nn.Parameter(torch.zeros(1, num_patches + 1, decoder_embed_dim), requires_grad=False)
        self.decoder_blocks = nn.ModuleList([_Block(decoder_embed_dim, decoder_num_heads, mlp_ratio) for _ in range(decoder_depth)])
        self.decoder_norm = nn.LayerNorm(decoder_embed_dim, eps=1e-6)
        self.decoder_pred = nn.Linear(decoder_embed_dim, patch_size ** 2 * in_chans, bias=True)
        self.norm_pix_loss = norm_pix_loss
        self.precision = "auto"  # "auto": bf16 tensor-core path under torch.autocast, fp32 validation path otherwise
        self.backend = "auto"
        self.initialize_weights()

    def initialize_weights(self):
        """models_mae.py:63-93: frozen sin-cos tables, xavier-uniform linears (patch embed as a linear), N(0, .02) tokens."""
        self.pos_embed.data.copy_(torch.from_numpy(_mae_sincos_2d(self.pos_embed.shape[-1], 14)).float().unsqueeze(0))
        self.decoder_pos_embed.data.copy_(torch.from_numpy(_mae_sincos_2d(self.decoder_pos_embed.shape[-1], 14)).float().unsqueeze(0))
        w = self.patch_embed.proj.weight.data
        torch.nn.init.xavier_uniform_(w.view([w.shape[0], -1]))
        torch.nn.init.normal_(self.cls_token, std=.02)
        torch.nn.init.normal_(self.mask_token, std=.02)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                torch.nn.init.xavier_uniform_(m.weight)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
            elif isinstance(m, nn.LayerNorm):
                nn.init.constant_(m.bias, 0)
                nn.init.constant_(m.weight, 1.0)

    def patchify(self, imgs):
        """[N, 3, H, W] -> [N, L, p*p*3], element order (row in patch, column in patch, channel) as models_mae.py:95-107.
        Host-side convenience only: the loss kernel builds the target from the image itself."""
        p = self.patch_embed.patch_size[0]
        n, c, hh, ww = imgs.shape
        gh, gw = hh // p, ww // p
        tiles = imgs.view(n, c, gh, p, gw, p).permute(0, 2, 4, 3, 5, 1)        # n, gh, gw, row, col, channel
        return tiles.reshape(n, gh * gw, p * p * c)

    def unpatchify(self, x):
        """Inverse of patchify (models_mae.py:109-121): [N, L, p*p*3] -> [N, 3, H, W]."""
        p = self.patch_embed.patch_size[0]
        n, L, _ = x.shape
        g = int(round(L ** 0.5))
        tiles = x.view(n, g, g, p, p, 3).permute(0, 5, 1, 3, 2, 4)             # n, channel, gh, row, gw, col
        return tiles.reshape(n, 3, g * p, g * p)

    def forward_loss(self, imgs, pred, mask):
        """models_mae.py:198-214."""
        return _MaeLossFn.apply(pred, imgs, mask, self.norm_pix_loss)

    def forward(self, imgs, mask_ratio=0.75, noise=None):
        """models_mae.py:216-220: (loss, pred [N, L, p*p*3], mask [N, L]).  `noise` ([N, L], optional) replaces the uniform noise
        random_masking draws (:132); by default it is drawn the same way, from torch's generator of the input's device."""
        if int(196 * (1 - mask_ratio)) != 49:
            raise NotImplementedError("the B200 schedule is built for mask_ratio 0.75 (49 visible patches)")
        if noise is None:
            noise = torch.rand(imgs.shape[0], 196, device=imgs.device)
        pred, mask, self.last_ids_restore = run_mae(self, imgs, noise)
        return self.forward_loss(imgs, pred, mask), pred, mask


def mae_vit_base_patch16_dec512d8b(**kwargs):
    """models_mae.py:223-228."""
    return MaskedAutoencoderViT(patch_size=16, embed_dim=768, depth=12, num_heads=12, decoder_embed_dim=512, decoder_depth=8,
                                decoder_num_heads=16, mlp_ratio=4, **kwargs)


mae_vit_base_patch16 = mae_vit_base_patch16_dec512d8b


def add_weight_decay(model, weight_decay=1e-5, skip_list=()):
    """timm optim_factory.add_weight_decay as called at Models/mae/main_pretrain.py:179: no decay for 1-D tensors, biases and
    names in skip_list; frozen tensors are left out."""
    decay, no_decay = [], []
    for name, param in model.named_parameters():
        if not param.requires_grad:
            continue
        if len(param.shape) == 1 or name.endswith(".bias") or name in skip_list:
            no_decay.append(param)
        else:
            decay.append(param)
    return [{"params": no_decay, "weight_decay": 0.}, {"params": decay, "weight_decay": weight_decay}]


def adjust_learning_rate(optimizer, epoch, lr, warmup_epochs=40, epochs=800, min_lr=0.0):
    """Models/mae/util/lr_sched.py:9-21 (half-cycle cosine after a linear warm-up); `epoch` is fractional
    (engine_pretrain.py:42-43: data_iter_step / len(data_loader) + epoch)."""
    if epoch < warmup_epochs:
        cur = lr * epoch / warmup_epochs
    else:
        cur = min_lr + (lr - min_lr) * 0.5 * (1. + math.cos(math.pi * (epoch - warmup_epochs) / (epochs - warmup_epochs)))
    for group in optimizer.param_groups:
        group["lr"] = cur * group["lr_scale"] if "lr_scale" in group else cur
    return cur


def train_one_epoch(model, data_loader, optimizer, device, epoch, loss_scaler=None, log_writer=None, args=None, grad_sync=None):
    """Drop-in for Models/mae/engine_pretrain.py:20-81 without its two host synchronisations per step (`loss.item()` at :50,
    `torch.cuda.synchronize()` at :62): the per-iteration LR schedule (:42-43), the bf16 forward (`torch.cuda.amp.autocast`
    at :47), backward and the optimizer step are enqueued back to back, the loss is accumulated on the device and read once at
    the end of the epoch (a non-finite loss is reported then, not at the step that produced it).

    args needs .lr, .warmup_epochs, .epochs, .min_lr, .mask_ratio and .accum_iter == 1 (gradient accumulation would need the
    flat gradient arena to add instead of overwrite).  loss_scaler is accepted for signature compatibility and unused: bf16
    needs no loss scaling.  grad_sync: a ddp.GradientAllReduce when training data-parallel.  Returns {"loss": mean, "lr": last}."""
    if getattr(args, "accum_iter", 1) != 1:
        raise NotImplementedError("ssl4gie_b200 MAE pretraining: accum_iter must be 1")
    model.train(True)
    core = model.module if hasattr(model, "module") else model
    n = len(data_loader)
    total = torch.zeros((), dtype=torch.float32, device=device)
    lr = 0.0
    for it, (samples, _) in enumerate(data_loader):
        lr = adjust_learning_rate(optimizer, it / n + epoch, args.lr, args.warmup_epochs, args.epochs, getattr(args, "min_lr", 0.0))
        samples = samples.to(device, non_blocking=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss, _, _ = core(samples, mask_ratio=args.mask_ratio)
        loss.backward()
        if grad_sync is not None:
            grad_sync.reduce()
        optimizer.step()
        optimizer.zero_grad()
        total += loss.detach()
        if log_writer is not None and (it + 1) % 20 == 0:
            log_writer.add_scalar("lr", lr, int((it / n + epoch) * 1000))
    mean = (total / max(n, 1)).item()
    if not math.isfinite(mean):
        raise FloatingPointError("Loss is {}, stopping training".format(mean))
    print("Averaged stats: lr: {:.6f}  loss: {:.4f}".format(lr, mean))
    return {"loss": mean, "lr": lr}
