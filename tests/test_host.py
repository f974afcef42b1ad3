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
    assert all(s is not None and 0 <= s[0] < s[1] <= n_active for s in sl)


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


def test_bucketed_allreduce_two_ranks_gloo(tmp_path):
    """world_size 2 on CPU: the union of the bucket slices covers every gradient exactly once, so the bucketed exchange
    equals one flat all-reduce."""
    from ssl4gie_b200 import utils
    m = utils.get_ImageNet_or_random_ViT(False, None, False, "depth", False, False)
    plan, offsets, n_active = _bucket_plan(m, 4)
    out = str(tmp_path / "r0.pt")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_gloo_worker, args=(2, port, plan, offsets, n_active, out), nprocs=2, join=True)
    got = torch.load(out)
    want = sum(torch.randn(n_active, generator=torch.Generator().manual_seed(100 + r)) for r in range(2))
    # padding gaps between parameters may be reduced or not; every parameter range must be
    for n, (off, numel) in offsets.items():
        if off < n_active:
            assert torch.allclose(got[off:off + numel], want[off:off + numel]), n
