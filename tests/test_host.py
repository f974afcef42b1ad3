"""CPU checks of the host side: the drop-in modules own exactly the reference's state_dict (golden manifests made from
the unmodified reference), factories keep the reference signatures, the arena / bucket plan is consistent, and the
bucketed gradient all-reduce works across 2 ranks (gloo)."""
import ctypes
import inspect
import json
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _manifest(model):
    return {k: [list(v.shape), str(v.dtype)] for k, v in model.state_dict().items()}


@pytest.mark.parametrize("dense", ["seg", "depth"])
def test_state_dict_manifest_matches_reference(dense, golden_dir):
    from ssl4gie_b200 import utils
    with open(os.path.join(golden_dir, f"manifest_{dense}.json")) as f:
        ref = json.load(f)
    m = utils.get_ImageNet_or_random_ViT(False, 1 if dense == "seg" else None, False, dense, False, False)
    assert _manifest(m) == ref


def test_variant_manifests_and_frozen_flags(golden_dir):
    """MAE / MoCo v3 wrappers: keys, shapes, non-trainable sets and the fixed sin-cos position embeddings."""
    from ssl4gie_b200 import utils
    with open(os.path.join(golden_dir, "manifest_variants.json")) as f:
        ref = json.load(f)
    mae = utils.get_MAE_backbone(None, False, 1, False, "seg", False)
    assert _manifest(mae) == ref["mae_seg"]
    assert sorted(n for n, p in mae.named_parameters() if not p.requires_grad) == ref["mae_seg_nograd"]
    head = np.load(os.path.join(golden_dir, "mae_pos_embed_head.npy"))
    assert np.allclose(mae.pos_embed[0, :4, :16].numpy(), head, atol=1e-6)
    moco = utils.get_MoCoV3_backbone(None, "vit_b", False, None, False, "depth", False)
    assert _manifest(moco) == ref["moco_depth"]
    assert sorted(n for n, p in moco.named_parameters() if not p.requires_grad) == ref["moco_depth_nograd"]
    head = np.load(os.path.join(golden_dir, "moco_pos_embed_head.npy"))
    assert np.allclose(moco.pos_embed[0, :4, :16].numpy(), head, atol=1e-6)
    cls = utils.get_ImageNet_or_random_ViT(True, 23, False, None, False, False)
    assert _manifest(cls) == ref["any_cls23"]


def test_imagenet_npz_loader(tmp_path, monkeypatch):
    """ImageNet_weights=True (reference Models/models.py:286-290 -> timm load_pretrained of a Flax .npz): the key map
    and layout transposes, on a synthetic checkpoint written in the Flax layout; the classifier arrays are ignored."""
    import numpy as np
    from ssl4gie_b200 import utils
    g = torch.Generator().manual_seed(5)
    want = {}
    arrays = {"head/kernel": np.zeros((768, 1000), np.float32), "head/bias": np.zeros(1000, np.float32)}

    def rnd(*shape):
        return torch.randn(*shape, generator=g) * 0.05

    want["patch_embed.proj.weight"] = rnd(768, 3, 16, 16)
    arrays["embedding/kernel"] = want["patch_embed.proj.weight"].permute(2, 3, 1, 0).numpy()          # HWIO
    want["patch_embed.proj.bias"] = rnd(768); arrays["embedding/bias"] = want["patch_embed.proj.bias"].numpy()
    want["cls_token"] = rnd(1, 1, 768); arrays["cls"] = want["cls_token"].numpy()
    want["pos_embed"] = rnd(1, 197, 768); arrays["Transformer/posembed_input/pos_embedding"] = want["pos_embed"].numpy()
    want["norm.weight"] = rnd(768); arrays["Transformer/encoder_norm/scale"] = want["norm.weight"].numpy()
    want["norm.bias"] = rnd(768); arrays["Transformer/encoder_norm/bias"] = want["norm.bias"].numpy()
    for i in range(12):
        b, o = f"Transformer/encoderblock_{i}/", f"blocks.{i}."
        for ln, fl in (("norm1", "LayerNorm_0"), ("norm2", "LayerNorm_2")):
            want[o + ln + ".weight"] = rnd(768); arrays[b + fl + "/scale"] = want[o + ln + ".weight"].numpy()
            want[o + ln + ".bias"] = rnd(768); arrays[b + fl + "/bias"] = want[o + ln + ".bias"].numpy()
        qkv_w, qkv_b = rnd(2304, 768), rnd(2304)
        want[o + "attn.qkv.weight"], want[o + "attn.qkv.bias"] = qkv_w, qkv_b
        for j, n in enumerate(("query", "key", "value")):
            # Flax: y[h, d] = sum_in x[in] * kernel[in, h, d]
            arrays[b + f"MultiHeadDotProductAttention_1/{n}/kernel"] = qkv_w[j * 768:(j + 1) * 768].T.reshape(768, 12, 64).numpy()
            arrays[b + f"MultiHeadDotProductAttention_1/{n}/bias"] = qkv_b[j * 768:(j + 1) * 768].reshape(12, 64).numpy()
        want[o + "attn.proj.weight"] = rnd(768, 768)
        arrays[b + "MultiHeadDotProductAttention_1/out/kernel"] = want[o + "attn.proj.weight"].T.reshape(12, 64, 768).numpy()
        want[o + "attn.proj.bias"] = rnd(768); arrays[b + "MultiHeadDotProductAttention_1/out/bias"] = want[o + "attn.proj.bias"].numpy()
        for r, (n_out, n_in) in enumerate(((3072, 768), (768, 3072))):
            want[o + f"mlp.fc{r + 1}.weight"] = rnd(n_out, n_in)
            arrays[b + f"MlpBlock_3/Dense_{r}/kernel"] = want[o + f"mlp.fc{r + 1}.weight"].T.numpy()
            want[o + f"mlp.fc{r + 1}.bias"] = rnd(n_out); arrays[b + f"MlpBlock_3/Dense_{r}/bias"] = want[o + f"mlp.fc{r + 1}.bias"].numpy()
    path = str(tmp_path / "vit_b16.npz")
    np.savez(path, **arrays)
    m = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, path)
    sd = m.state_dict()
    for k, v in want.items():
        assert torch.equal(sd[k], v), k
    # the attention layout means what timm's Attention computes: q of head h = x @ kernel[:, h, :] + bias[h]
    x = torch.randn(3, 768, generator=g)
    q_flax = torch.einsum("ni,ihd->nhd", x, torch.from_numpy(arrays["Transformer/encoderblock_3/MultiHeadDotProductAttention_1/query/kernel"])) + \
        torch.from_numpy(arrays["Transformer/encoderblock_3/MultiHeadDotProductAttention_1/query/bias"])
    q_timm = torch.nn.functional.linear(x, sd["blocks.3.attn.qkv.weight"], sd["blocks.3.attn.qkv.bias"])[:, :768].reshape(3, 12, 64)
    assert torch.allclose(q_flax, q_timm, atol=1e-5)
    # ImageNet_weights=True with no file anywhere: a clear error, never a download
    monkeypatch.setenv("TORCH_HOME", str(tmp_path / "empty_hub"))
    monkeypatch.delenv("SSL4GIE_B200_IMAGENET_NPZ", raising=False)
    with pytest.raises(RuntimeError, match="never touches the network"):
        utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, True)
    monkeypatch.setenv("SSL4GIE_B200_IMAGENET_NPZ", path)
    m2 = utils.get_ImageNet_or_random_ViT(True, 23, False, None, False, True)
    assert torch.equal(m2.state_dict()["blocks.11.mlp.fc2.weight"], want["blocks.11.mlp.fc2.weight"])


def test_factory_signatures_match_reference():
    """reference utils.py:8-23, :26-50, :59-82 (callers pass positionally)."""
    from ssl4gie_b200 import utils
    assert list(inspect.signature(utils.get_MAE_backbone).parameters) == \
        ["weight_path", "head", "num_classes", "frozen", "dense", "det", "fixed_size", "out_token"]
    assert list(inspect.signature(utils.get_MoCoV3_backbone).parameters) == \
        ["weight_path", "arch", "head", "num_classes", "frozen", "dense", "det", "fixed_size", "out_token"]
    assert list(inspect.signature(utils.get_ImageNet_or_random_ViT).parameters) == \
        ["head", "num_classes", "frozen", "dense", "det", "ImageNet_weights", "fixed_size", "out_token"]


def test_load_state_dict_roundtrip_and_cpu_refusal():
    from oracle import ref_port  # checker only: supplies reference-named tensors
    from ssl4gie_b200 import utils
    sd = ref_port.make_state("seg", seed=3)
    m = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
    missing, unexpected = m.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    for k, v in m.state_dict().items():
        assert torch.equal(v, sd[k]), k
    with pytest.raises(RuntimeError, match="no CPU"):
        m(torch.zeros(1, 3, 224, 224))


def _bucket_plan(model, n_buckets):
    """Bucket membership straight from the C library (needs no GPU: binding only records pointers)."""
    from ssl4gie_b200 import _C
    from ssl4gie_b200.engine import arena_layout
    lib = _C.lib()
    offsets, n_active, total = arena_layout(model)
    desc = _C.S4GModelDesc(batch=2, dense=_C.DENSE[model.dense], dtype=_C.BF16, backend=_C.TENSOR, frozen=0, world_size=1,
                           num_classes=0, out_token=0, dropout_p=0.1)
    h = ctypes.c_void_p()
    _C.check(lib.s4g_model_create(ctypes.byref(desc), ctypes.byref(h)))
    grads = torch.zeros(n_active)
    for n, p in model.named_parameters():
        off = offsets[n][0]
        g = ctypes.c_void_p(grads.data_ptr() + 4 * off) if off < n_active else None
        _C.check(lib.s4g_model_bind(h, n.encode(), _C.ptr(p.data), g, p.numel()))
    plan = [[s for s in lib.s4g_model_bucket_params(h, i, n_buckets).decode().split("\n") if s] for i in range(n_buckets)]
    lib.s4g_model_destroy(h)
    return plan, offsets, n_active


def test_arena_layout_and_bucket_plan():
    from ssl4gie_b200 import utils
    from ssl4gie_b200.ddp import bucket_slices
    m = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
    plan, offsets, n_active = _bucket_plan(m, 5)
    names = [n for b in plan for n in b]
    with_grad = [n for n in offsets if offsets[n][0] < n_active]
    assert sorted(names) == sorted(with_grad) and len(names) == len(set(names))
    # the statically-unused parameters are outside the gradient arena (reference: grad stays None)
    for n in offsets:
        unused = n.startswith("norm.") or n.startswith("decoder.refinenet4.resConfUnit1.")
        assert (offsets[n][0] >= n_active) == unused, n
        assert offsets[n][0] % 64 == 0
    # bucket 0 = decoder (finishes first in backward), last bucket holds block 0 and the patch embedding
    assert all(n.startswith("decoder.") for n in plan[0])
    assert "blocks.11.mlp.fc2.weight" in plan[1] and "blocks.0.attn.qkv.weight" in plan[-1] and "patch_embed.proj.weight" in plan[-1]
    sl = bucket_slices(plan, offsets)
    assert all(len(r) == 1 and 0 <= r[0][0] < r[0][1] <= n_active for r in sl)   # one contiguous range per bucket
    # one bucket = the whole gradient arena (its event is recorded at the END of backward, csrc/engine.cu)
    plan1, _, _ = _bucket_plan(m, 1)
    assert sorted(plan1[0]) == sorted(with_grad)
    assert bucket_slices(plan1, offsets)[0][0][0] == 0
    # two buckets: decoder | whole encoder
    plan2, _, _ = _bucket_plan(m, 2)
    assert all(n.startswith("decoder.") for n in plan2[0]) and sorted(plan2[0] + plan2[1]) == sorted(with_grad)


def _gloo_worker(rank, world, port, plan, offsets, n_active, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ssl4gie_b200.ddp import allreduce_buckets, bucket_slices
    g = torch.Generator().manual_seed(100 + rank)
    flat = torch.randn(n_active, generator=g)
    allreduce_buckets(flat, bucket_slices(plan, offsets))
    if rank == 0:
        torch.save(flat, out)
    dist.destroy_process_group()


@pytest.mark.parametrize("which", ["depth", "mae"])
def test_bucketed_allreduce_two_ranks_gloo(tmp_path, which):
    """world_size 2 on CPU: the bucket ranges cover every gradient exactly once, so the bucketed exchange equals one flat
    all-reduce -- also for MaskedAutoencoderViT, whose root-level parameters (cls_token, mask_token) sit at the front of the
    arena although they belong to the last / first bucket."""
    import ssl4gie_b200
    from ssl4gie_b200 import utils
    from ssl4gie_b200.ddp import bucket_slices
    if which == "mae":
        m = ssl4gie_b200.mae_vit_base_patch16(norm_pix_loss=True)
    else:
        m = utils.get_ImageNet_or_random_ViT(False, None, False, "depth", False, False)
    plan, offsets, n_active = _bucket_plan(m, 4)
    ranges = sorted(r for b in bucket_slices(plan, offsets) for r in b)
    assert all(a[1] <= b[0] for a, b in zip(ranges, ranges[1:])), "bucket ranges overlap"
    if which == "mae":
        assert "mask_token" in plan[0] and "decoder_pred.weight" in plan[0] and "norm.weight" in plan[0]
        assert "cls_token" in plan[-1] and "patch_embed.proj.weight" in plan[-1] and "pos_embed" not in sum(plan, [])
    out = str(tmp_path / "r0.pt")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_gloo_worker, args=(2, port, plan, offsets, n_active, out), nprocs=2, join=True)
    got = torch.load(out)
    want = sum(torch.randn(n_active, generator=torch.Generator().manual_seed(100 + r)) for r in range(2))
    # padding gaps between parameters may be reduced or not; every parameter range must be
    for n, (off, numel) in offsets.items():
        if off < n_active:
            assert torch.allclose(got[off:off + numel], want[off:off + numel]), n


def test_mae_module_manifest_groups_and_schedule(golden_dir):
    """MaskedAutoencoderViT (ssl4gie_b200/mae.py) has the reference module's state_dict (fixture written from the unmodified
    Models/mae/models_mae.py), the same frozen tensors, timm's add_weight_decay grouping and lr_sched's schedule."""
    import json
    import os

    import ssl4gie_b200
    from oracle import ref_port
    m = ssl4gie_b200.mae_vit_base_patch16(norm_pix_loss=True)
    with open(os.path.join(golden_dir, "manifest_mae.json")) as f:
        man = json.load(f)
    sd = m.state_dict()
    assert sorted(sd) == sorted(man)
    for k, (shape, dtype) in man.items():
        assert list(sd[k].shape) == shape and str(sd[k].dtype) == dtype, k
    assert sorted(n for n, p in m.named_parameters() if not p.requires_grad) == sorted(ref_port.MAE_FROZEN)
    assert torch.equal(sd["pos_embed"], ref_port.sincos_2d(768)) and torch.equal(sd["decoder_pos_embed"], ref_port.sincos_2d(512))
    groups = ssl4gie_b200.add_weight_decay(m, 0.05)
    names = {id(p): n for n, p in m.named_parameters()}
    ref_groups = ref_port.mae_param_groups({k: v for k, v in sd.items()}, 0.05)
    for g, r in zip(groups, ref_groups):
        assert g["weight_decay"] == r["weight_decay"] and sorted(names[id(p)] for p in g["params"]) == sorted(r["names"])

    class _Opt:
        param_groups = [{"lr": 0.0}, {"lr": 0.0, "lr_scale": 0.5}]

    for it in (0, 7, 400, 4000, 50000):
        lr = ssl4gie_b200.adjust_learning_rate(_Opt, it / 100, 1.5e-4, warmup_epochs=40, epochs=800)
        assert abs(lr - ref_port.mae_lr(it, 100, 1.5e-4)) < 1e-15
        assert _Opt.param_groups[0]["lr"] == lr and _Opt.param_groups[1]["lr"] == lr * 0.5
    with pytest.raises(NotImplementedError):
        ssl4gie_b200.MaskedAutoencoderViT()  # mae_vit_large defaults: outside the B200 schedule


def test_input_pipeline_draws_like_the_reference():
    """GpuInputPipeline.draw consumes torch's generator and `random` exactly as one sample of the reference does:
    ColorJitter.get_params, GaussianBlur.get_params (dataloaders.py:64-66), then dataset.py:46-60."""
    import random

    T = pytest.importorskip("torchvision.transforms")
    from ssl4gie_b200 import GpuInputPipeline
    from ssl4gie_b200.augment import inverse_affine_matrix
    from torchvision.transforms.functional import _get_inverse_affine_matrix
    pipe = GpuInputPipeline(train=True)
    jit = T.ColorJitter(brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01)
    for seed in range(5):
        torch.manual_seed(seed)
        random.seed(seed)
        p, p2 = pipe.draw(), pipe.draw()
        torch.manual_seed(seed)
        random.seed(seed)
        for q in (p, p2):
            order, b, c, s, h = T.ColorJitter.get_params(jit.brightness, jit.contrast, jit.saturation, jit.hue)
            sigma = T.GaussianBlur.get_params(0.001, 2.0)
            assert (q.order, q.brightness, q.contrast, q.saturation, q.hue, q.sigma) == (order.tolist(), b, c, s, h, sigma)
            assert q.hflip == (random.uniform(0.0, 1.0) > 0.5) and q.vflip == (random.uniform(0.0, 1.0) > 0.5)
            ref = (random.uniform(-180.0, 180.0), (random.uniform(-28.0, 28.0), random.uniform(-28.0, 28.0)),
                   random.uniform(0.5, 1.5), random.uniform(-22.5, 22.5))
            assert q.affine == ref
            assert inverse_affine_matrix(*ref) == _get_inverse_affine_matrix([0.0, 0.0], ref[0], list(ref[1]), ref[2], [ref[3], 0.0])


def test_depth_input_pipeline_draws():
    """task="depth" (Depth_estimation/Data/dataloaders.py:55-63, dataset.py:55-63): ColorJitter.get_params, then the two flips;
    no blur, no affine draw."""
    import random

    T = pytest.importorskip("torchvision.transforms")
    from ssl4gie_b200 import GpuInputPipeline
    pipe = GpuInputPipeline(train=True, task="depth")
    jit = T.ColorJitter(brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01)
    torch.manual_seed(3)
    random.seed(3)
    ps = [pipe.draw() for _ in range(3)]
    torch.manual_seed(3)
    random.seed(3)
    for q in ps:
        order, b, c, s_, h = T.ColorJitter.get_params(jit.brightness, jit.contrast, jit.saturation, jit.hue)
        assert (q.order, q.brightness, q.contrast, q.saturation, q.hue) == (order.tolist(), b, c, s_, h)
        assert q.hflip == (random.uniform(0.0, 1.0) > 0.5) and q.vflip == (random.uniform(0.0, 1.0) > 0.5)
        assert q.sigma is None and q.affine is None


def test_checkpoint_scaler_state_loads_into_an_enabled_gradscaler():
    """save_checkpoint(scaler=None) stores the state of a fresh, enabled GradScaler: the reference's build() loads
    "scaler_state_dict" unconditionally (train_segmentation.py:205-207) and torch refuses an empty dict there."""
    from ssl4gie_b200 import train
    sc = torch.amp.GradScaler("cpu")
    assert sc.is_enabled() and sc.state_dict() == train._default_scaler_state()
    sc.load_state_dict(train._default_scaler_state())
    assert sc.get_scale() == 65536.0
