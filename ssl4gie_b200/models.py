"""ViT-B/16 wrappers with the reference's constructors, attributes and state_dict keys
(reference Models/models.py:262-578).  forward() enters the fused CUDA schedule (engine.run_model)."""
import math
from functools import partial

import numpy as np
import torch
import torch.nn as nn

from .dpt_decoder import DPT_decoder
from .engine import run_model


class _PatchEmbed(nn.Module):
    """timm 0.6.12 PatchEmbed surface: .proj Conv2d(3, D, 16, 16), grid/patch metadata."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3, embed_dim=768):
        super().__init__()
        self.img_size = (img_size, img_size)
        self.patch_size = (patch_size, patch_size)
        self.grid_size = (img_size // patch_size, img_size // patch_size)
        self.num_patches = self.grid_size[0] * self.grid_size[1]
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_size, stride=patch_size)


class _Attention(nn.Module):
    def __init__(self, dim, num_heads):
        super().__init__()
        self.num_heads = num_heads
        self.scale = (dim // num_heads) ** -0.5
        self.qkv = nn.Linear(dim, dim * 3, bias=True)
        self.proj = nn.Linear(dim, dim)


class _Mlp(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.fc2 = nn.Linear(hidden, dim)


class _Block(nn.Module):
    """timm 0.6.12 Block parameter layout: norm1, attn.{qkv,proj}, norm2, mlp.{fc1,fc2}; LayerNorm eps 1e-6."""

    def __init__(self, dim, num_heads, mlp_ratio=4.0):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=1e-6)
        self.attn = _Attention(dim, num_heads)
        self.norm2 = nn.LayerNorm(dim, eps=1e-6)
        self.mlp = _Mlp(dim, int(dim * mlp_ratio))


def _check_supported(embed_dim, depth, num_heads, det):
    if det:
        raise NotImplementedError("det=True (ViTDet windowed-attention detection backbone) is outside the B200 hot path")
    if (embed_dim, depth, num_heads) != (768, 12, 12):
        raise NotImplementedError("the B200 schedule is specialised for ViT-B/16 (768, 12, 12)")


class _Backbone(nn.Module):
    """Shared tail of the three wrappers' constructors (reference Models/models.py:303-323, :401-415, :513-529)."""

    def _finish(self, head, num_classes, frozen, dense, det, embed_dim, out_token):
        self.head_bool = head
        self.num_classes_head = num_classes if head else 0
        if head:
            self.lin_head = nn.Linear(embed_dim, num_classes)
        self.frozen = frozen
        self.dense = dense
        self.det = det
        if dense:
            self.decoder = DPT_decoder(num_classes=num_classes, dense=dense)
        self.out_token = out_token
        self.precision = "auto"  # "auto": bf16 tensor-core path under torch.autocast, fp32 validation path otherwise
        self.backend = "auto"

    def forward(self, x):
        """dense: [B,1,224,224]; otherwise final norm + cls / mean-token pooling (+ lin_head): [B,num_classes] or [B,768]
        (reference Models/models.py:340-357)."""
        return run_model(self, x)


class VisionTransformer_from_Any(_Backbone):
    """reference Models/models.py:262-357 (timm VisionTransformer base, random or ImageNet init)."""

    def __init__(self, head, num_classes, frozen, dense, det, fixed_size, embed_dim, depth, num_heads, out_token,
                 ImageNet_weights=False):
        super().__init__()
        _check_supported(embed_dim, depth, num_heads, det)
        self.num_features = self.embed_dim = embed_dim
        self.patch_embed = _PatchEmbed(224, 16, 3, embed_dim)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.randn(1, 197, embed_dim) * 0.02)
        self.blocks = nn.Sequential(*[_Block(embed_dim, num_heads) for _ in range(depth)])
        self.norm = nn.LayerNorm(embed_dim, eps=1e-6)
        nn.init.trunc_normal_(self.pos_embed, std=0.02)  # timm init_weights
        nn.init.normal_(self.cls_token, std=1e-6)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                nn.init.zeros_(m.bias)
        if ImageNet_weights:
            # reference Models/models.py:286-290: timm's download_cached_file + load_pretrained of the AugReg B/16 .npz.
            # There is no network here, so the file must already be where timm would have cached it (or be named by
            # ImageNet_weights=<path> / SSL4GIE_B200_IMAGENET_NPZ); the .npz -> state_dict key map is timm's.
            self.load_pretrained(_find_imagenet_npz(ImageNet_weights))
        self.head = nn.Identity()
        self._finish(head, num_classes, frozen, dense, det, embed_dim, out_token)

    @torch.no_grad()
    def load_pretrained(self, checkpoint_path, prefix=""):
        """timm 0.6.12 VisionTransformer.load_pretrained -> _load_weights for a Flax/JAX ViT ``.npz`` (the call at
        reference Models/models.py:290).  Kernels are stored input-major (HWIO / [in, heads, head_dim]); the classifier
        ("head/*", 1000 ImageNet classes) is dropped, as the reference replaces it with Identity right after loading."""
        w = np.load(checkpoint_path)
        sd = npz_to_state_dict(w, prefix, depth=len(self.blocks))
        own = self.state_dict()
        for k, v in sd.items():
            if tuple(own[k].shape) != tuple(v.shape):
                raise ValueError(f"{checkpoint_path}: {k} has shape {tuple(v.shape)}, expected {tuple(own[k].shape)}")
            own[k].copy_(v)


IMAGENET_NPZ = ("B_16-i21k-300ep-lr_0.001-aug_medium1-wd_0.1-do_0.0-sd_0.0--imagenet2012-steps_20k-lr_0.01-res_224.npz")


def _find_imagenet_npz(arg):
    import os
    cands = []
    if isinstance(arg, (str, os.PathLike)):
        cands.append(os.fspath(arg))
    if os.environ.get("SSL4GIE_B200_IMAGENET_NPZ"):
        cands.append(os.environ["SSL4GIE_B200_IMAGENET_NPZ"])
    cands.append(os.path.join(torch.hub.get_dir(), "checkpoints", IMAGENET_NPZ))  # timm download_cached_file's location
    for c in cands:
        if os.path.isfile(c):
            return c
    raise RuntimeError("ImageNet_weights=True needs the AugReg ViT-B/16 checkpoint " + IMAGENET_NPZ + " (reference "
                       "Models/models.py:286-290 downloads it; this build never touches the network): place it at " +
                       cands[-1] + ", or pass its path as ImageNet_weights / SSL4GIE_B200_IMAGENET_NPZ")


def npz_to_state_dict(w, prefix="", depth=12):
    """Flax ViT checkpoint arrays -> timm VisionTransformer state_dict entries (timm 0.6.12 _load_weights)."""
    def t(a):
        return torch.from_numpy(np.asarray(a, dtype=np.float32))

    sd = {}
    k = t(w[prefix + "embedding/kernel"])                       # [16, 16, 3, D] HWIO
    sd["patch_embed.proj.weight"] = k.permute(3, 2, 0, 1).contiguous()
    sd["patch_embed.proj.bias"] = t(w[prefix + "embedding/bias"])
    sd["cls_token"] = t(w[prefix + "cls"]).reshape(1, 1, -1)
    sd["pos_embed"] = t(w[prefix + "Transformer/posembed_input/pos_embedding"]).reshape(1, -1, k.shape[-1])
    sd["norm.weight"] = t(w[prefix + "Transformer/encoder_norm/scale"])
    sd["norm.bias"] = t(w[prefix + "Transformer/encoder_norm/bias"])
    for i in range(depth):
        b = f"{prefix}Transformer/encoderblock_{i}/"
        mha = b + "MultiHeadDotProductAttention_1/"
        o = f"blocks.{i}."
        sd[o + "norm1.weight"] = t(w[b + "LayerNorm_0/scale"])
        sd[o + "norm1.bias"] = t(w[b + "LayerNorm_0/bias"])
        sd[o + "attn.qkv.weight"] = torch.cat([t(w[mha + n + "/kernel"]).flatten(1).T for n in ("query", "key", "value")])
        sd[o + "attn.qkv.bias"] = torch.cat([t(w[mha + n + "/bias"]).reshape(-1) for n in ("query", "key", "value")])
        sd[o + "attn.proj.weight"] = t(w[mha + "out/kernel"]).permute(2, 0, 1).flatten(1).contiguous()   # [H, hd, D] -> [D, H*hd]
        sd[o + "attn.proj.bias"] = t(w[mha + "out/bias"])
        for r in range(2):
            sd[o + f"mlp.fc{r + 1}.weight"] = t(w[b + f"MlpBlock_3/Dense_{r}/kernel"]).T.contiguous()
            sd[o + f"mlp.fc{r + 1}.bias"] = t(w[b + f"MlpBlock_3/Dense_{r}/bias"])
        sd[o + "norm2.weight"] = t(w[b + "LayerNorm_2/scale"])
        sd[o + "norm2.bias"] = t(w[b + "LayerNorm_2/bias"])
    return sd


def _sincos_1d(embed_dim, pos):
    omega = np.arange(embed_dim // 2, dtype=float)
    omega /= embed_dim / 2.0
    omega = 1.0 / 10000 ** omega
    out = np.einsum("m,d->md", pos.reshape(-1), omega)
    return np.concatenate([np.sin(out), np.cos(out)], axis=1)


def _mae_sincos_2d(embed_dim, grid_size):
    """reference Models/mae/util/pos_embed.py:20-66 (cls_token=True)"""
    gh = np.arange(grid_size, dtype=np.float32)
    gw = np.arange(grid_size, dtype=np.float32)
    grid = np.stack(np.meshgrid(gw, gh), axis=0).reshape([2, 1, grid_size, grid_size])
    emb = np.concatenate([_sincos_1d(embed_dim // 2, grid[0]), _sincos_1d(embed_dim // 2, grid[1])], axis=1)
    return np.concatenate([np.zeros([1, embed_dim]), emb], axis=0)


class ViT_from_MAE(_Backbone):
    """reference Models/models.py:360-475 (MaskedAutoencoderViT base with the MAE decoder deleted)."""

    def __init__(self, weight_path, head, num_classes, frozen, dense, det, fixed_size, embed_dim, depth, num_heads, out_token):
        super().__init__()
        _check_supported(embed_dim, depth, num_heads, det)
        self.patch_embed = _PatchEmbed(224, 16, 3, embed_dim)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, 197, embed_dim), requires_grad=False)
        self.blocks = nn.ModuleList([_Block(embed_dim, num_heads) for _ in range(depth)])
        self.norm = nn.LayerNorm(embed_dim, eps=1e-6)
        # survives the `del`s at Models/models.py:395-399, so it is part of the state_dict contract
        self.decoder_pos_embed = nn.Parameter(torch.zeros(1, 197, 512), requires_grad=False)
        self.norm_pix_loss = False
        # MaskedAutoencoderViT.initialize_weights (Models/mae/models_mae.py:64-92)
        self.pos_embed.data.copy_(torch.from_numpy(_mae_sincos_2d(embed_dim, 14)).float().unsqueeze(0))
        self.decoder_pos_embed.data.copy_(torch.from_numpy(_mae_sincos_2d(512, 14)).float().unsqueeze(0))
        w = self.patch_embed.proj.weight.data
        nn.init.xavier_uniform_(w.view([w.shape[0], -1]))
        nn.init.normal_(self.cls_token, std=0.02)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.xavier_uniform_(m.weight)
                nn.init.constant_(m.bias, 0)
        if weight_path is not None:
            weights = torch.load(weight_path, map_location="cpu")["model"]
            self.load_my_state_dict(weights)
        self.head = head
        self._finish(head, num_classes, frozen, dense, det, embed_dim, out_token)

    def load_my_state_dict(self, state_dict):
        """reference Models/models.py:417-425: copy entries whose names exist here, skip the rest silently."""
        i = 0
        own_state = self.state_dict()
        for name, param in state_dict.items():
            if name not in own_state:
                continue
            own_state[name].copy_(param)
            i += 1
        print(f"Successfully loaded params for {i} items")


class ViT_from_MoCoV3(_Backbone):
    """reference Models/models.py:478-578 (VisionTransformerMoCo base, Models/moco_v3/vits.py:25-69)."""

    def __init__(self, weight_path, head, num_classes, frozen, dense, det, fixed_size, embed_dim, out_token):
        super().__init__()
        _check_supported(embed_dim, 12, 12, det)
        self.num_features = self.embed_dim = embed_dim
        self.patch_embed = _PatchEmbed(224, 16, 3, embed_dim)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, embed_dim))
        self.blocks = nn.Sequential(*[_Block(embed_dim, 12) for _ in range(12)])
        self.norm = nn.LayerNorm(embed_dim, eps=1e-6)
        # build_2d_sincos_position_embedding (vits.py:53-69): frozen
        gw, gh = torch.meshgrid(torch.arange(14, dtype=torch.float32), torch.arange(14, dtype=torch.float32), indexing="ij")
        pos_dim = embed_dim // 4
        omega = 1.0 / (10000.0 ** (torch.arange(pos_dim, dtype=torch.float32) / pos_dim))
        out_w = torch.einsum("m,d->md", [gw.flatten(), omega])
        out_h = torch.einsum("m,d->md", [gh.flatten(), omega])
        pos = torch.cat([torch.sin(out_w), torch.cos(out_w), torch.sin(out_h), torch.cos(out_h)], dim=1)[None]
        self.pos_embed = nn.Parameter(torch.cat([torch.zeros(1, 1, embed_dim), pos], dim=1), requires_grad=False)
        for name, m in self.named_modules():  # vits.py:32-41
            if isinstance(m, nn.Linear):
                if "qkv" in name:
                    val = math.sqrt(6.0 / float(m.weight.shape[0] // 3 + m.weight.shape[1]))
                    nn.init.uniform_(m.weight, -val, val)
                else:
                    nn.init.xavier_uniform_(m.weight)
                nn.init.zeros_(m.bias)
        nn.init.normal_(self.cls_token, std=1e-6)
        val = math.sqrt(6.0 / float(3 * 16 * 16 + embed_dim))
        nn.init.uniform_(self.patch_embed.proj.weight, -val, val)
        nn.init.zeros_(self.patch_embed.proj.bias)
        self.head = nn.Identity()
        if weight_path is not None:
            weights = torch.load(weight_path, map_location="cpu")
            self.load_state_dict(weights)  # strict, before decoder / lin_head exist (Models/models.py:506-509)
        self.patch_embed.proj.weight.requires_grad = True
        self.patch_embed.proj.bias.requires_grad = True
        self._finish(head, num_classes, frozen, dense, det, embed_dim, out_token)
