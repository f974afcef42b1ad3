"""MAE pretraining step (BASELINE.json config 5) on the GPU against the oracle restatement of
Models/mae/models_mae.py and the fixture written from the unmodified reference (tests/golden/step_mae_b2.npz)."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_port  # checker only

pytestmark = pytest.mark.gpu


def _inputs(B=2, seed=5):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, 3, 224, 224, generator=g), torch.rand(B, 196, generator=g)


def _model(sd, precision, backend):
    import ssl4gie_b200
    m = ssl4gie_b200.mae_vit_base_patch16(norm_pix_loss=True)
    m.load_state_dict({k: v.cpu() for k, v in sd.items()}, strict=True)
    m.precision, m.backend = precision, backend
    return m.cuda().train()


def _l2(a, b):
    return ((a.double() - b.double()).norm() / (b.double().norm() + 1e-300)).item()


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


def _oracle(sd, x, noise, dtype=torch.float32):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sdd = {k: v.cuda().to(dtype) for k, v in sd.items()}
    return ref_port.mae_loss_and_grads(sdd, x.cuda().to(dtype), noise.cuda().to(dtype))


def test_masking_matches_argsort():
    """random_masking (models_mae.py:123-148): ids_restore / mask equal torch.argsort's on the same noise."""
    sd = ref_port.mae_make_state(seed=3)
    m = _model(sd, "fp32", "simt")
    x, noise = _inputs(4, seed=11)
    with torch.no_grad():
        _, _, mask = m(x.cuda(), noise=noise.cuda())
    ids_shuffle = torch.argsort(noise, dim=1)
    ids_restore = torch.argsort(ids_shuffle, dim=1)
    assert torch.equal(m.last_ids_restore.cpu().long(), ids_restore)
    ref_mask = torch.ones(4, 196)
    ref_mask[:, :49] = 0
    assert torch.equal(mask.cpu(), torch.gather(ref_mask, 1, ids_restore))


def test_loss_kernel_matches_oracle():
    """forward_loss (models_mae.py:198-214) and its gradient on random predictions, norm_pix on and off."""
    import ssl4gie_b200
    g = torch.Generator().manual_seed(9)
    B = 3
    imgs = torch.randn(B, 3, 224, 224, generator=g).cuda()
    mask = (torch.rand(B, 196, generator=g) < 0.75).float().cuda()
    for norm_pix in (True, False):
        m = ssl4gie_b200.mae_vit_base_patch16(norm_pix_loss=norm_pix)
        pred = torch.randn(B, 196, 768, generator=g).cuda().requires_grad_(True)
        loss = m.forward_loss(imgs, pred, mask)
        loss.backward()
        p2 = pred.detach().double().requires_grad_(True)
        target = ref_port.mae_patchify(imgs.double())
        if norm_pix:
            target = (target - target.mean(-1, keepdim=True)) / (target.var(-1, keepdim=True) + 1e-6) ** 0.5
        ref = ((((p2 - target) ** 2).mean(-1)) * mask.double()).sum() / mask.double().sum()
        ref.backward()
        assert abs(loss.item() - ref.item()) < 1e-5 * abs(ref.item())
        assert _l2(pred.grad, p2.grad) < 1e-5


def test_fp32_step_matches_reference_fixture(golden_dir):
    """fp32 validation mode at B=2: mask exact, pred to 1e-4, loss to 1e-5 against the fixture of the unmodified reference;
    every gradient against the oracle in fp64 (relative L2 <= 1e-3) and the fixture's norms / leading elements."""
    fix = np.load(os.path.join(golden_dir, "step_mae_b2.npz"))
    sd = ref_port.mae_make_state(seed=42)
    m = _model(sd, "fp32", "simt")
    x, noise = _inputs()
    loss, pred, mask = m(x.cuda(), noise=noise.cuda())
    loss.backward()
    assert torch.equal(mask.cpu(), torch.from_numpy(fix["mask"]))
    e_pred = (pred.detach().cpu()[:, ::7, ::16] - torch.from_numpy(fix["pred_sub"])).abs().max().item() / float(fix["pred_absmax"])
    e_loss = abs(loss.item() - float(fix["loss"])) / abs(float(fix["loss"]))
    l64, p64, _, g64 = _oracle(sd, x, noise, torch.float64)
    worst = max((_l2(p.grad, g64[n]), n) for n, p in m.named_parameters() if p.requires_grad)
    print(f"[mae fp32 B=2] pred {e_pred:.2e} loss {e_loss:.2e} (vs fp64 oracle {abs(loss.item() - l64.item()):.2e}) worst gradient l2 {worst[0]:.2e} ({worst[1]})")
    assert e_pred < 1e-4 and e_loss < 1e-5
    assert m.pos_embed.grad is None and m.decoder_pos_embed.grad is None
    assert worst[0] < 1e-3, worst
    floor = 1e-3 * float(np.median(fix["grad_norm"]))
    for i, n in enumerate(fix["grad_names"].tolist()):
        g, rn = m.get_parameter(n).grad, float(fix["grad_norm"][i])
        assert abs(g.double().norm().item() - rn) < 2e-3 * max(rn, floor), n
        k = min(8, g.numel())
        assert (g.flatten()[:k].cpu() - torch.from_numpy(fix["grad_head"][i][:k])).abs().max().item() < 2e-3 * max(g.abs().max().item(), floor), n


@pytest.mark.parametrize("B", [2, 16, 64])
def test_bf16_step_against_oracle(B):
    """bf16 tensor-core path (tcgen05 GEMMs, tensor-core attention incl. the padded 32-dim decoder heads): pred 2e-2 of its
    range, loss 1e-2, the whole gradient AND every single tensor within relative L2 2e-2 / cosine 0.999 of the fp32 oracle
    (measured: 0.7e-2 whole, 1.1e-2 worst tensor); a second run is bit-identical."""
    sd = ref_port.mae_make_state(seed=42)
    m = _model(sd, "bf16", "tensor")
    x, noise = _inputs(B, seed=5 if B == 2 else 21 + B)
    loss, pred, mask = m(x.cuda(), noise=noise.cuda())
    loss.backward()
    grads = {n: p.grad.clone() for n, p in m.named_parameters() if p.requires_grad}
    l_ref, p_ref, m_ref, g_ref = _oracle(sd, x, noise)
    assert torch.equal(mask, m_ref)
    e_pred = (pred.detach() - p_ref).abs().max().item() / p_ref.abs().max().item()
    e_loss = abs(loss.item() - l_ref.item()) / abs(l_ref.item())
    names = sorted(grads)
    ga = torch.cat([grads[n].flatten().double() for n in names])
    gr = torch.cat([g_ref[n].flatten().double() for n in names])
    per = sorted((( _l2(grads[n], g_ref[n]), _cos(grads[n], g_ref[n]), n) for n in names), reverse=True)
    strict = sum(1 for r in per if r[0] <= 2e-2 and r[1] >= 0.999)
    print(f"[mae bf16 B={B}] pred {e_pred:.2e} loss {e_loss:.2e} whole gradient l2 {_l2(ga, gr):.3e} cos {_cos(ga, gr):.6f}; "
          f"per tensor max l2 {per[0][0]:.3e} ({per[0][2]}) min cos {min(r[1] for r in per):.6f} strict_pass {strict}/{len(per)}")
    assert e_pred < 2e-2 and e_loss < 1e-2
    assert _l2(ga, gr) <= 2e-2 and _cos(ga, gr) >= 0.999
    assert strict == len(per), per[:4]   # the plain criterion (2e-2 relative L2, cosine 0.999) for EVERY tensor
    m.zero_grad(set_to_none=True)
    loss2, pred2, _ = m(x.cuda(), noise=noise.cuda())
    loss2.backward()
    assert torch.equal(pred2, pred) and loss2.item() == loss.item()
    assert all(torch.equal(m.get_parameter(n).grad, grads[n]) for n in names)


def test_pretrain_steps_follow_the_reference_optimizer():
    """20 steps of the reference recipe (AdamW betas (0.9, 0.95), weight decay 0.05 on >= 2-D tensors only, per-iteration
    warm-up + half-cosine LR; main_pretrain.py:179-180, engine_pretrain.py:42-43) in fp32 against the oracle with
    torch.optim.AdamW: the loss of every step within 1e-3, final parameters within 1e-3 of the update they received."""
    from ssl4gie_b200 import FusedAdamW, add_weight_decay, adjust_learning_rate
    sd = ref_port.mae_make_state(seed=7)
    m = _model(sd, "fp32", "simt")
    opt = FusedAdamW(m, lr=1e-3, betas=(0.9, 0.95), param_groups=add_weight_decay(m, 0.05))
    assert [len(g["params"]) for g in opt.param_groups] == [len(g["names"]) for g in ref_port.mae_param_groups(sd)]
    torch.backends.cuda.matmul.allow_tf32 = False
    full = {k: v.cuda().clone() for k, v in sd.items()}
    params = {k: v.requires_grad_(True) for k, v in full.items() if k not in ref_port.MAE_FROZEN}
    full.update(params)
    ropt = torch.optim.AdamW([{"params": [params[n] for n in g["names"]], "weight_decay": g["weight_decay"]}
                              for g in ref_port.mae_param_groups(sd)], lr=1e-3, betas=(0.9, 0.95))
    steps, worst = 20, 0.0
    for it in range(steps):
        x, noise = (t.cuda() for t in _inputs(2, seed=100 + it))
        lr = adjust_learning_rate(opt, it / 10, 1e-3, warmup_epochs=1, epochs=4)
        assert abs(lr - ref_port.mae_lr(it, 10, 1e-3, 1, 4)) < 1e-12
        for g in ropt.param_groups:
            g["lr"] = lr
        loss, _, _ = m(x, noise=noise)
        loss.backward()
        opt.step()
        ropt.zero_grad(set_to_none=True)
        l_ref, _, _ = ref_port.mae_forward(full, x, noise)
        l_ref.backward()
        ropt.step()
        worst = max(worst, abs(loss.item() - l_ref.item()) / abs(l_ref.item()))
    moved = max(_l2(p.detach() - sd[n].cuda(), params[n].detach() - sd[n].cuda()) for n, p in m.named_parameters() if p.requires_grad)
    print(f"[mae fp32 {steps} optimizer steps] worst loss deviation {worst:.2e}, worst relative error of a parameter's total update {moved:.2e}")
    assert worst < 1e-3 and moved < 2e-2


def test_train_one_epoch_loop():
    """ssl4gie_b200.mae.train_one_epoch (engine_pretrain.py:20-81 without its per-step host syncs): bf16 under autocast, the
    per-iteration schedule, a falling loss on a repeated batch, and the same result from a second identical run."""
    import types

    from ssl4gie_b200 import FusedAdamW, add_weight_decay
    from ssl4gie_b200.mae import train_one_epoch
    args = types.SimpleNamespace(lr=3e-4, warmup_epochs=1, epochs=10, min_lr=0.0, mask_ratio=0.75, accum_iter=1)
    x, _ = _inputs(8, seed=31)
    loader = [(x, None)] * 6
    out = []
    for _ in range(2):
        torch.manual_seed(0)            # the masking noise is drawn from torch's CUDA generator, as in the reference
        m = _model(ref_port.mae_make_state(seed=42), "auto", "auto")
        opt = FusedAdamW(m, lr=args.lr, betas=(0.9, 0.95), param_groups=add_weight_decay(m, 0.05))
        stats = [train_one_epoch(m, loader, opt, torch.device("cuda"), ep, None, None, args) for ep in range(3)]
        out.append([s["loss"] for s in stats])
        assert abs(stats[-1]["lr"] - ref_port.mae_lr(2 * 6 + 5, 6, args.lr, 1, 10)) < 1e-12
    assert out[0][2] < out[0][0] and out[0] == out[1], out
