"""CPU checks of the ORACLE (oracle/ref_port.py) against the committed golden fixtures that
oracle/make_golden.py produced from the UNMODIFIED reference modules (tests/golden/*).  No GPU, no product code."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import ref_port


def _inputs(dense, B=2):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(B, 3, 224, 224, generator=g)
    if dense == "seg":
        t = (torch.rand(B, 1, 224, 224, generator=g) > 0.5).float()
    else:
        t = torch.rand(B, 1, 224, 224, generator=g)
        t[torch.rand(B, 1, 224, 224, generator=g) < 0.1] = 0.0
        t[:, :, :8, :] = 0.0
    return x, t


@pytest.fixture(scope="module")
def kats(golden_dir):
    with open(os.path.join(golden_dir, "loss_kats.json")) as f:
        return json.load(f)


def _kat_dice():
    logits = torch.linspace(-3, 3, 128, dtype=torch.float64).view(2, 1, 8, 8).float()
    targets = (torch.arange(128) % 3 == 0).float().view(2, 1, 8, 8)
    return logits, targets


def _kat_ssi():
    i = torch.arange(512, dtype=torch.float64).view(2, 1, 16, 16)
    pred = (0.5 + 0.4 * torch.sin(0.37 * i)).float()
    targ = (0.5 + 0.45 * torch.cos(0.11 * i)).float()
    targ[(torch.arange(512) % 7 == 0).view(2, 1, 16, 16)] = 0
    return pred, targ


def test_dice_kat(kats):
    """SoftDiceLoss KAT from the reference's own loss code (Binary_segmentation/Metrics/losses.py:10-24)."""
    logits, targets = _kat_dice()
    x = logits.clone().requires_grad_(True)
    loss = ref_port.soft_dice_loss(x, targets)
    loss.backward()
    assert abs(loss.item() - kats["dice_loss"]) < 1e-7
    assert abs(loss.item() - 0.549143434) < 1e-6  # SURVEY.md §8c.5 value
    assert torch.allclose(x.grad.flatten(), torch.tensor(kats["dice_grad"]), rtol=1e-5, atol=1e-9)
    z = ref_port.soft_dice_loss(torch.zeros(2, 1, 4, 4), torch.ones(2, 1, 4, 4)).item()
    assert abs(z - kats["dice_zeros_ones"]) < 1e-7 and abs(z - 0.2) < 1e-6


@pytest.mark.parametrize("alpha", [0.1, 0.0])
def test_ssi_kat(kats, alpha):
    """ScaleAndShiftInvariantLoss KATs (Depth_estimation/Metrics/losses.py:130-146), incl. the gradient through
    compute_scale_and_shift."""
    pred, targ = _kat_ssi()
    p = pred.clone().requires_grad_(True)
    loss = ref_port.ssi_loss(p, targ, alpha)
    loss.backward()
    assert abs(loss.item() - kats[f"ssi_loss_a{alpha}"]) < 1e-7
    assert torch.allclose(p.grad.flatten(), torch.tensor(kats[f"ssi_grad_a{alpha}"]), rtol=1e-4, atol=1e-9)
    s, sh = ref_port.compute_scale_and_shift(pred.squeeze(1), targ.squeeze(1), targ.squeeze(1) > 0)
    assert torch.allclose(s, torch.tensor(kats["ssi_scale"]), rtol=1e-5)
    assert torch.allclose(sh, torch.tensor(kats["ssi_shift"]), rtol=1e-5)
    assert ref_port.ssi_loss(targ.clamp_min(1e-3), targ.clamp_min(1e-3), alpha).item() < 1e-10  # loss(pred = target) = 0


def test_ssi_random_with_empty_mask_image(kats):
    g = torch.Generator().manual_seed(3)
    p = torch.rand(3, 1, 32, 32, generator=g)
    t = torch.rand(3, 1, 32, 32, generator=g)
    t[torch.rand(3, 1, 32, 32, generator=g) < 0.2] = 0
    t[2] = 0
    pp = p.clone().requires_grad_(True)
    loss = ref_port.ssi_loss(pp, t, 0.1)
    loss.backward()
    assert abs(loss.item() - kats["ssi_rand_loss"]) < 1e-7
    assert abs(pp.grad.double().norm().item() - kats["ssi_rand_grad_norm"]) < 1e-5 * kats["ssi_rand_grad_norm"]
    assert torch.allclose(pp.grad.flatten()[:16], torch.tensor(kats["ssi_rand_grad_head"]), rtol=1e-4, atol=1e-10)
    # all-masked batch: reduction_batch_based returns python 0 (losses.py:34-35)
    assert ref_port.ssi_loss(p, torch.zeros_like(t), 0.0) == 0


def _ssi_ext_cases():
    g = torch.Generator().manual_seed(3)
    p = torch.rand(3, 1, 32, 32, generator=g)
    t = torch.rand(3, 1, 32, 32, generator=g)
    t[torch.rand(3, 1, 32, 32, generator=g) < 0.2] = 0
    t[2] = 0
    g = torch.Generator().manual_seed(4)
    p6 = torch.rand(2, 1, 64, 64, generator=g)
    t6 = torch.rand(2, 1, 64, 64, generator=g)
    t6[torch.rand(2, 1, 64, 64, generator=g) < 0.15] = 0
    return {"img": (p, t, dict(alpha=0.1, reduction="image-based")), "img_a0": (p, t, dict(alpha=0.0, reduction="image-based")),
            "s6": (p6, t6, dict(alpha=0.25, scales=6)), "s5_img": (p6, t6, dict(alpha=0.25, scales=5, reduction="image-based"))}


@pytest.mark.parametrize("tag", ["img", "img_a0", "s6", "s5_img"])
def test_ssi_image_based_and_more_scales(kats, tag):
    """reduction="image-based" (Depth_estimation/Metrics/losses.py:40-48) and scales > 4 (:104-117): known answers from
    the reference's own loss code (oracle/make_golden.py)."""
    p, t, kw = _ssi_ext_cases()[tag]
    pp = p.clone().requires_grad_(True)
    loss = ref_port.ssi_loss(pp, t, **kw)
    loss.backward()
    assert abs(loss.item() - kats[f"ssi_{tag}_loss"]) < 1e-7
    assert abs(pp.grad.double().norm().item() - kats[f"ssi_{tag}_grad_norm"]) < 1e-5 * kats[f"ssi_{tag}_grad_norm"]
    assert torch.allclose(pp.grad.flatten()[:16], torch.tensor(kats[f"ssi_{tag}_grad_head"]), rtol=1e-4, atol=1e-10)


@pytest.mark.parametrize("dense", ["seg", "depth"])
def test_step_matches_reference_fixture(dense, golden_dir):
    """One fwd + loss + bwd at B=2 fp32 (BASELINE.json configs[0]) reproduces the reference's logits, loss, gradient
    norms / leading elements and the None-gradient set."""
    torch.set_num_threads(os.cpu_count() or 1)
    fix = np.load(os.path.join(golden_dir, f"step_{dense}_b2.npz"))
    sd = ref_port.make_state(dense, seed=42)
    x, t = _inputs(dense)
    stats = {}
    out, loss, grads = ref_port.loss_and_grads(sd, x, t, dense, None, 0.1, stats)
    ref_sub = torch.from_numpy(fix["out_sub"])
    assert (out[:, :, ::8, ::8] - ref_sub).abs().max().item() < 1e-4 * float(fix["out_absmax"])
    assert abs(loss.item() - float(fix["loss"])) < 1e-5 * abs(float(fix["loss"]))
    assert sorted(n for n, g in grads.items() if g is None) == sorted(fix["none_grads"].tolist())
    floor = 1e-3 * float(np.median(fix["grad_norm"]))
    for i, n in enumerate(fix["grad_names"].tolist()):
        g = grads[n]
        rn = float(fix["grad_norm"][i])
        assert abs(g.double().norm().item() - rn) < 2e-3 * max(rn, floor), n
        k = min(8, g.numel())
        head = torch.from_numpy(fix["grad_head"][i][:k])
        assert (g.flatten()[:k] - head).abs().max().item() < 2e-3 * max(g.abs().max().item(), floor), n
    if dense == "seg":
        # running stats after one training forward: 0.9 * init + 0.1 * batch stat (unbiased var)
        for i, k in enumerate(fix["bn_names"].tolist()):
            p = k[:-len("running_mean")]
            if p not in stats:
                assert int(fix["bn_tracked"][i]) == 0  # refinenet4.resConfUnit1 never runs (DPT_decoder.py:288)
                continue
            mean, var = stats[p]
            assert torch.allclose(0.1 * mean, torch.from_numpy(fix["bn_running_mean"][i]), rtol=1e-3, atol=1e-6), k
            assert torch.allclose(0.9 + 0.1 * var, torch.from_numpy(fix["bn_running_var"][i]), rtol=1e-4), k
            assert int(fix["bn_tracked"][i]) == 1


@pytest.mark.parametrize("dense", ["seg", "depth"])
def test_state_manifest(dense, golden_dir):
    """make_state() has exactly the reference module's state_dict keys / shapes / dtypes."""
    with open(os.path.join(golden_dir, f"manifest_{dense}.json")) as f:
        man = json.load(f)
    sd = ref_port.make_state(dense, seed=1)
    assert sorted(sd) == sorted(man)
    for k, (shape, dtype) in man.items():
        assert list(sd[k].shape) == shape and str(sd[k].dtype) == dtype, k


def test_adamw_restatement_matches_torch():
    g = torch.Generator().manual_seed(5)
    p0 = torch.randn(1000, generator=g)
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([p], lr=1e-4)
    q, m, v = p0.clone(), torch.zeros(1000), torch.zeros(1000)
    for step in range(1, 4):
        grad = torch.randn(1000, generator=g)
        p.grad = grad.clone()
        opt.step()
        ref_port.adamw_step(q, grad, m, v, step)
    assert torch.allclose(q, p.detach(), rtol=1e-6, atol=1e-8)


@pytest.mark.parametrize("out_token", ["cls", "spatial"])
def test_cls_step_matches_reference_fixture(out_token, golden_dir):
    """Classification branch (Models/models.py:338,348-354) + the weighted CE of train_classification.py:278 at B=2:
    the restatement reproduces the unmodified reference's logits, loss and every gradient."""
    torch.set_num_threads(os.cpu_count() or 1)
    fix = np.load(os.path.join(golden_dir, f"step_cls_{out_token}_b2.npz"))
    sd = ref_port.make_state(None, seed=42, head_classes=23)
    x = torch.randn(2, 3, 224, 224, generator=torch.Generator().manual_seed(0))
    y, cw = torch.from_numpy(fix["labels"]), torch.from_numpy(fix["class_weights"])
    out, loss, grads = ref_port.loss_and_grads(sd, x, y, None, out_token=out_token, class_weights=cw)
    assert (out - torch.from_numpy(fix["logits"])).abs().max().item() < 1e-5 * float(np.abs(fix["logits"]).max())
    assert abs(loss.item() - float(fix["loss"])) < 1e-5 * abs(float(fix["loss"]))
    assert all(g is not None for g in grads.values())  # the final norm and the head are live in this branch
    floor = 1e-3 * float(np.median(fix["grad_norm"]))
    for i, n in enumerate(fix["grad_names"].tolist()):
        rn = float(fix["grad_norm"][i])
        assert abs(grads[n].double().norm().item() - rn) < 2e-3 * max(rn, floor), n


def test_weighted_ce_restatement_matches_torch():
    g = torch.Generator().manual_seed(2)
    x = torch.randn(16, 23, generator=g) * 2
    y = torch.randint(0, 23, (16,), generator=g)
    w = torch.rand(23, generator=g) + 0.5
    assert torch.allclose(ref_port.weighted_ce_loss(x, y, w), torch.nn.CrossEntropyLoss(w)(x, y), rtol=1e-6)
    assert torch.allclose(ref_port.weighted_ce_loss(x, y), torch.nn.CrossEntropyLoss()(x, y), rtol=1e-6)


def mae_inputs(B=2):
    """The inputs of oracle/make_golden.py:mae_inputs (image batch, masking noise)."""
    g = torch.Generator().manual_seed(5)
    return torch.randn(B, 3, 224, 224, generator=g), torch.rand(B, 196, generator=g)


def test_mae_step_matches_reference_fixture(golden_dir):
    """One MAE pretraining step at B=2 fp32 (BASELINE.json config 5: mae_vit_base_patch16, norm_pix_loss, mask 0.75)
    reproduces the unmodified Models/mae/models_mae.py: mask, prediction, loss, every gradient's norm and head."""
    torch.set_num_threads(os.cpu_count() or 1)
    fix = np.load(os.path.join(golden_dir, "step_mae_b2.npz"))
    sd = ref_port.mae_make_state(seed=42)
    with open(os.path.join(golden_dir, "manifest_mae.json")) as f:
        man = json.load(f)
    assert sorted(sd) == sorted(man)
    for k, (shape, dtype) in man.items():
        assert list(sd[k].shape) == shape and str(sd[k].dtype) == dtype, k
    x, noise = mae_inputs()
    loss, pred, mask, grads = ref_port.mae_loss_and_grads(sd, x, noise)
    assert torch.equal(mask, torch.from_numpy(fix["mask"]))
    assert (pred[:, ::7, ::16] - torch.from_numpy(fix["pred_sub"])).abs().max().item() < 1e-4 * float(fix["pred_absmax"])
    assert abs(loss.item() - float(fix["loss"])) < 1e-5 * abs(float(fix["loss"]))
    assert sorted(grads) == sorted(fix["grad_names"].tolist())
    floor = 1e-3 * float(np.median(fix["grad_norm"]))
    for i, n in enumerate(fix["grad_names"].tolist()):
        g, rn = grads[n], float(fix["grad_norm"][i])
        assert abs(g.double().norm().item() - rn) < 2e-3 * max(rn, floor), n
        k = min(8, g.numel())
        assert (g.flatten()[:k] - torch.from_numpy(fix["grad_head"][i][:k])).abs().max().item() < 2e-3 * max(g.abs().max().item(), floor), n


def test_mae_param_groups_and_lr():
    """timm add_weight_decay rule (main_pretrain.py:179) and the per-iteration half-cosine (lr_sched.py:9-21)."""
    sd = ref_port.mae_make_state(seed=1)
    nd, d = ref_port.mae_param_groups(sd, 0.05)
    assert nd["weight_decay"] == 0.0 and d["weight_decay"] == 0.05
    assert "cls_token" in d["names"] and "mask_token" in d["names"] and "blocks.0.attn.qkv.weight" in d["names"]
    assert "blocks.0.norm1.weight" in nd["names"] and "decoder_pred.bias" in nd["names"]
    assert "pos_embed" not in nd["names"] + d["names"]
    assert len(nd["names"]) + len(d["names"]) == len(sd) - 2
    assert ref_port.mae_lr(0, 100, 1e-3) == 0.0
    assert abs(ref_port.mae_lr(2000, 100, 1e-3) - 0.5e-3) < 1e-12         # half-way through warm-up (epoch 20 of 40)
    assert abs(ref_port.mae_lr(4000, 100, 1e-3) - 1e-3) < 1e-12           # end of warm-up
    assert abs(ref_port.mae_lr(80000, 100, 1e-3)) < 1e-12                 # end of the schedule
