"""ORACLE tooling — test infrastructure, NOT product code.

Runs in the BUILD container only (needs /root/reference).  Imports the
UNMODIFIED reference modules (Models/models.py, Models/DPT_decoder.py, both
Metrics/losses.py, utils.py) through ``oracle/timm_shim`` and

  1. strict-loads ``ref_port.make_state`` tensors into the reference modules
     (=> proves the state_dict manifest: names, shapes, dtypes),
  2. runs one fwd + loss + bwd at B=2 (BASELINE.json configs[0]) for seg and depth,
  3. asserts ``oracle/ref_port.py`` reproduces logits / loss / every gradient,
  4. writes small fixtures to tests/golden/ (sub-sampled logits, loss, per-parameter
     gradient norms + leading elements, BN running stats, None-grad list, loss KATs).

Usage:  python oracle/make_golden.py
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("SSL4GIE_REF", "/root/reference")
sys.path.insert(0, os.path.join(HERE, "timm_shim"))
sys.path.insert(0, REF)
sys.path.insert(0, os.path.dirname(HERE))
np.float = float  # Models/mae/util/pos_embed.py:56 uses the removed alias

import importlib.util  # noqa: E402

import utils as ref_utils  # noqa: E402  (/root/reference/utils.py)

from oracle import ref_port  # noqa: E402


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


seg_losses = _load(os.path.join(REF, "Binary_segmentation/Metrics/losses.py"), "ref_seg_losses")
depth_losses = _load(os.path.join(REF, "Depth_estimation/Metrics/losses.py"), "ref_depth_losses")

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def inputs(dense, B=2):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(B, 3, 224, 224, generator=g)
    if dense == "seg":
        t = (torch.rand(B, 1, 224, 224, generator=g) > 0.5).float()
    else:
        t = torch.rand(B, 1, 224, 224, generator=g)
        t[torch.rand(B, 1, 224, 224, generator=g) < 0.1] = 0.0
        t[:, :, :8, :] = 0.0
    return x, t


def run(dense):
    torch.manual_seed(42)
    model = ref_utils.get_ImageNet_or_random_ViT(False, 1 if dense == "seg" else None, False, dense, False, False)
    sd = ref_port.make_state(dense, seed=42)
    # the reference module also owns ViT `head`-less keys only; strict load proves the manifest
    missing, unexpected = model.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    manifest = {k: [list(v.shape), str(v.dtype)] for k, v in model.state_dict().items()}
    model.train()
    if dense == "seg":
        model.decoder.output_conv[3].p = 0.0  # Dropout(0.1) (DPT_decoder.py:494): RNG-free fixture
    x, t = inputs(dense)
    out = model(x)
    loss_fn = seg_losses.SoftDiceLoss() if dense == "seg" else depth_losses.ScaleAndShiftInvariantLoss(alpha=0.1)
    loss = loss_fn(out, t)
    loss.backward()
    none_grads = sorted(n for n, p in model.named_parameters() if p.requires_grad and p.grad is None)
    grads = {n: p.grad for n, p in model.named_parameters() if p.grad is not None}

    # ---- check the restatement against the real thing
    o2, l2, g2 = ref_port.loss_and_grads(sd, x, t, dense, None, 0.1)
    err = (o2 - out.detach()).abs().max().item() / out.detach().abs().max().item()
    assert err < 1e-5, err
    assert abs(l2.item() - loss.item()) < 1e-6 * max(1, abs(loss.item())), (l2.item(), loss.item())
    worst = 0.0
    for n, g in grads.items():
        e = (g2[n] - g).abs().max().item() / (g.abs().max().item() + 1e-20)
        worst = max(worst, e)
    assert worst < 1e-4, worst
    assert sorted(n for n, g in g2.items() if g is None) == none_grads
    print(f"[{dense}] ref_port vs reference: out {err:.2e} loss {abs(l2.item()-loss.item()):.2e} grads {worst:.2e}")

    fix = {
        "loss": np.float64(loss.item()),
        "out_sub": out.detach()[:, :, ::8, ::8].numpy(),
        "out_absmax": np.float64(out.detach().abs().max().item()),
    }
    names = sorted(grads)
    fix["grad_names"] = np.array(names)
    fix["grad_norm"] = np.array([grads[n].double().norm().item() for n in names])
    fix["grad_head"] = np.stack([
        np.pad(grads[n].flatten()[:8].numpy(), (0, max(0, 8 - grads[n].numel()))) for n in names])
    fix["none_grads"] = np.array(none_grads)
    if dense == "seg":
        msd = model.state_dict()
        bn_names = sorted(k for k in msd if k.endswith("running_mean"))
        fix["bn_names"] = np.array(bn_names)
        fix["bn_running_mean"] = np.stack([msd[k].numpy() for k in bn_names])
        fix["bn_running_var"] = np.stack([msd[k.replace("running_mean", "running_var")].numpy() for k in bn_names])
        fix["bn_tracked"] = np.array([int(msd[k.replace("running_mean", "num_batches_tracked")]) for k in bn_names])
    np.savez_compressed(os.path.join(OUT, f"step_{dense}_b2.npz"), **fix)
    with open(os.path.join(OUT, f"manifest_{dense}.json"), "w") as f:
        json.dump(manifest, f, indent=0, sort_keys=True)
    return model


def run_cls(out_token):
    """Classification branch (Models/models.py:338,348-354) with the 23-class weighted CE of
    Classification/train_classification.py:157,278, through the UNMODIFIED reference module."""
    torch.manual_seed(42)
    model = ref_utils.get_ImageNet_or_random_ViT(True, 23, False, None, False, False, out_token=out_token)
    sd = ref_port.make_state(None, seed=42, head_classes=23)
    missing, unexpected = model.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    model.train()
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 3, 224, 224, generator=g)
    y = torch.tensor([3, 17])
    cw = torch.linspace(0.5, 2.0, 23)
    out = model(x)
    loss = torch.nn.CrossEntropyLoss(cw)(out, y)
    loss.backward()
    grads = {n: p.grad for n, p in model.named_parameters() if p.grad is not None}
    o2, l2, g2 = ref_port.loss_and_grads(sd, x, y, None, out_token=out_token, class_weights=cw)
    err = (o2 - out.detach()).abs().max().item() / out.detach().abs().max().item()
    assert err < 1e-5, err
    assert abs(l2.item() - loss.item()) < 1e-6 * max(1, abs(loss.item()))
    worst = max((g2[n] - g).abs().max().item() / (g.abs().max().item() + 1e-20) for n, g in grads.items())
    assert worst < 1e-4, worst
    assert sorted(n for n, gg in g2.items() if gg is None) == sorted(n for n, p in model.named_parameters() if p.grad is None)
    print(f"[cls/{out_token}] ref_port vs reference: out {err:.2e} loss {abs(l2.item()-loss.item()):.2e} grads {worst:.2e}")
    names = sorted(grads)
    fix = {"loss": np.float64(loss.item()), "logits": out.detach().numpy(), "labels": y.numpy(), "class_weights": cw.numpy(),
           "grad_names": np.array(names), "grad_norm": np.array([grads[n].double().norm().item() for n in names]),
           "grad_head": np.stack([np.pad(grads[n].flatten()[:8].numpy(), (0, max(0, 8 - grads[n].numel()))) for n in names])}
    np.savez_compressed(os.path.join(OUT, f"step_cls_{out_token}_b2.npz"), **fix)


def loss_kats():
    """RNG-free known-answer tests straight from the reference loss code."""
    kat = {}
    logits = torch.linspace(-3, 3, 128, dtype=torch.float64).view(2, 1, 8, 8).float().requires_grad_(True)
    targets = (torch.arange(128) % 3 == 0).float().view(2, 1, 8, 8)
    loss = seg_losses.SoftDiceLoss()(logits, targets)
    loss.backward()
    kat["dice_loss"] = loss.item()
    kat["dice_grad"] = logits.grad.flatten().tolist()
    kat["dice_zeros_ones"] = seg_losses.SoftDiceLoss()(torch.zeros(2, 1, 4, 4), torch.ones(2, 1, 4, 4)).item()
    i = torch.arange(512, dtype=torch.float64).view(2, 1, 16, 16)
    pred = (0.5 + 0.4 * torch.sin(0.37 * i)).float()
    targ = (0.5 + 0.45 * torch.cos(0.11 * i)).float()
    targ[(torch.arange(512) % 7 == 0).view(2, 1, 16, 16)] = 0
    for alpha in (0.1, 0.0):
        p = pred.clone().requires_grad_(True)
        fn = depth_losses.ScaleAndShiftInvariantLoss(alpha=alpha)
        loss = fn(p, targ)
        loss.backward()
        kat[f"ssi_loss_a{alpha}"] = loss.item()
        kat[f"ssi_grad_a{alpha}"] = p.grad.flatten().tolist()
    s, sh = depth_losses.compute_scale_and_shift(pred.squeeze(1), targ.squeeze(1), targ.squeeze(1) > 0)
    kat["ssi_scale"] = s.tolist()
    kat["ssi_shift"] = sh.tolist()
    # bigger random case incl. an all-masked image (empty-mask edge case, losses.py:20,34)
    g = torch.Generator().manual_seed(3)
    p = torch.rand(3, 1, 32, 32, generator=g)
    t = torch.rand(3, 1, 32, 32, generator=g)
    t[torch.rand(3, 1, 32, 32, generator=g) < 0.2] = 0
    t[2] = 0
    pp = p.clone().requires_grad_(True)
    loss = depth_losses.ScaleAndShiftInvariantLoss(alpha=0.1)(pp, t)
    loss.backward()
    kat["ssi_rand_loss"] = loss.item()
    kat["ssi_rand_grad_norm"] = pp.grad.double().norm().item()
    kat["ssi_rand_grad_head"] = pp.grad.flatten()[:16].tolist()
    # reduction="image-based" (losses.py:40-48) and more than four scales, same random case and a 64 x 64 one
    for tag, pq, tq, kw in (("img", p, t, dict(alpha=0.1, reduction="image-based")),
                            ("img_a0", p, t, dict(alpha=0.0, reduction="image-based"))):
        pp = pq.clone().requires_grad_(True)
        loss = depth_losses.ScaleAndShiftInvariantLoss(**kw)(pp, tq)
        loss.backward()
        kat[f"ssi_{tag}_loss"] = loss.item()
        kat[f"ssi_{tag}_grad_norm"] = pp.grad.double().norm().item()
        kat[f"ssi_{tag}_grad_head"] = pp.grad.flatten()[:16].tolist()
    g = torch.Generator().manual_seed(4)
    p6 = torch.rand(2, 1, 64, 64, generator=g)
    t6 = torch.rand(2, 1, 64, 64, generator=g)
    t6[torch.rand(2, 1, 64, 64, generator=g) < 0.15] = 0
    for tag, kw in (("s6", dict(alpha=0.25, scales=6)), ("s5_img", dict(alpha=0.25, scales=5, reduction="image-based"))):
        pp = p6.clone().requires_grad_(True)
        loss = depth_losses.ScaleAndShiftInvariantLoss(**kw)(pp, t6)
        loss.backward()
        kat[f"ssi_{tag}_loss"] = loss.item()
        kat[f"ssi_{tag}_grad_norm"] = pp.grad.double().norm().item()
        kat[f"ssi_{tag}_grad_head"] = pp.grad.flatten()[:16].tolist()
        l4 = ref_port.ssi_loss(p6, t6, kw["alpha"], kw["scales"], reduction=kw.get("reduction", "batch-based"))
        assert abs(l4.item() - loss.item()) < 1e-7, (tag, l4.item(), loss.item())
    with open(os.path.join(OUT, "loss_kats.json"), "w") as f:
        json.dump(kat, f)
    # restatement agrees
    l2 = ref_port.soft_dice_loss(logits.detach(), targets)
    assert abs(l2.item() - kat["dice_loss"]) < 1e-7
    l3 = ref_port.ssi_loss(pred, targ, 0.1)
    assert abs(l3.item() - kat["ssi_loss_a0.1"]) < 1e-7
    print("loss KATs:", kat["dice_loss"], kat["ssi_loss_a0.1"], kat["ssi_loss_a0.0"], kat["ssi_scale"], kat["ssi_shift"])


def mae_inputs(B=2):
    g = torch.Generator().manual_seed(5)
    return torch.randn(B, 3, 224, 224, generator=g), torch.rand(B, 196, generator=g)


def run_mae():
    """One MAE pretraining step (BASELINE config 5) through the UNMODIFIED Models/mae/models_mae.py at B=2, with the
    masking noise of random_masking (:132) supplied by patching torch.rand for the duration of the call."""
    from Models.mae import models_mae

    torch.manual_seed(42)
    model = models_mae.mae_vit_base_patch16(norm_pix_loss=True)
    sd = ref_port.mae_make_state(seed=42)
    ref_pos = model.pos_embed.detach().clone(), model.decoder_pos_embed.detach().clone()
    assert torch.equal(ref_pos[0], sd["pos_embed"]) and torch.equal(ref_pos[1], sd["decoder_pos_embed"]), "sin-cos tables"
    missing, unexpected = model.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    model.train()
    x, noise = mae_inputs()
    real_rand = torch.rand
    torch.rand = lambda *a, **k: noise.clone()
    try:
        loss, pred, mask = model(x, mask_ratio=0.75)
    finally:
        torch.rand = real_rand
    loss.backward()
    none_grads = sorted(n for n, p in model.named_parameters() if p.requires_grad and p.grad is None)
    frozen = sorted(n for n, p in model.named_parameters() if not p.requires_grad)
    assert frozen == sorted(ref_port.MAE_FROZEN) and not none_grads, (frozen, none_grads)
    grads = {n: p.grad for n, p in model.named_parameters() if p.grad is not None}
    l2, p2, m2, g2 = ref_port.mae_loss_and_grads(sd, x, noise)
    assert torch.equal(m2, mask)
    err = (p2 - pred.detach()).abs().max().item() / pred.detach().abs().max().item()
    worst = max((g2[n] - g).abs().max().item() / (g.abs().max().item() + 1e-20) for n, g in grads.items())
    assert err < 1e-5 and abs(l2.item() - loss.item()) < 1e-6 and worst < 1e-4, (err, l2.item(), loss.item(), worst)
    print(f"[mae] ref_port vs reference: pred {err:.2e} loss {abs(l2.item()-loss.item()):.2e} grads {worst:.2e}")
    names = sorted(grads)
    fix = {"loss": np.float64(loss.item()), "pred_sub": pred.detach()[:, ::7, ::16].numpy(), "mask": mask.numpy(),
           "pred_absmax": np.float64(pred.detach().abs().max().item()), "grad_names": np.array(names),
           "grad_norm": np.array([grads[n].double().norm().item() for n in names]),
           "grad_head": np.stack([np.pad(grads[n].flatten()[:8].numpy(), (0, max(0, 8 - grads[n].numel()))) for n in names])}
    np.savez_compressed(os.path.join(OUT, "step_mae_b2.npz"), **fix)
    with open(os.path.join(OUT, "manifest_mae.json"), "w") as f:
        json.dump({k: [list(v.shape), str(v.dtype)] for k, v in model.state_dict().items()}, f, indent=0, sort_keys=True)


def variants_manifest():
    """Key/shape manifests for the MAE- and MoCo-based wrappers (Models/models.py:360-578)."""
    from Models import models as ref_models

    out = {}
    torch.manual_seed(42)
    m = ref_models.ViT_from_MAE(None, False, 1, False, "seg", False, None, 768, 12, 12, "cls")
    out["mae_seg"] = {k: [list(v.shape), str(v.dtype)] for k, v in m.state_dict().items()}
    out["mae_seg_nograd"] = sorted(n for n, p in m.named_parameters() if not p.requires_grad)
    np.save(os.path.join(OUT, "mae_pos_embed_head.npy"), m.pos_embed[0, :4, :16].detach().numpy())
    # MoCo wrapper strict-loads encoder-only weights before adding the decoder (models.py:506-509)
    from Models.moco_v3 import vits

    base = vits.VisionTransformerMoCo(patch_size=16, embed_dim=768, depth=12, num_heads=12, mlp_ratio=4,
                                      qkv_bias=True, num_classes=4096)
    base.head = torch.nn.Identity()
    path = "/tmp/_moco_enc.pth"
    torch.save(base.state_dict(), path)
    m2 = ref_utils.get_MoCoV3_backbone(path, "vit_b", False, None, False, "depth", False)
    out["moco_depth"] = {k: [list(v.shape), str(v.dtype)] for k, v in m2.state_dict().items()}
    out["moco_depth_nograd"] = sorted(n for n, p in m2.named_parameters() if not p.requires_grad)
    np.save(os.path.join(OUT, "moco_pos_embed_head.npy"), m2.pos_embed[0, :4, :16].detach().numpy())
    m3 = ref_utils.get_ImageNet_or_random_ViT(True, 23, False, None, False, False)
    out["any_cls23"] = {k: [list(v.shape), str(v.dtype)] for k, v in m3.state_dict().items()}
    with open(os.path.join(OUT, "manifest_variants.json"), "w") as f:
        json.dump(out, f, indent=0, sort_keys=True)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    loss_kats()
    if "--kats-only" in sys.argv:
        sys.exit(0)
    if "--mae-only" in sys.argv:
        run_mae()
        sys.exit(0)
    if "--cls-only" not in sys.argv:
        run("seg")
        run("depth")
    run_cls("cls")
    run_cls("spatial")
    run_mae()
    if "--cls-only" in sys.argv:
        sys.exit(0)
    try:
        variants_manifest()
    except Exception as e:  # variants are informational; the step fixtures above are the pins
        print("variants manifest failed:", repr(e))
        raise
