"""Classification branch (reference Models/models.py:338,348-354 + the weighted CE of
Classification/train_classification.py:278) through the drop-in nn.Module API, against the committed golden fixtures
generated from the unmodified reference and against the oracle run live."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_port  # noqa: E402  (checker only)


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30)).item()


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


def _build(out_token, precision, frozen=False, head=True, seed=42):
    from ssl4gie_b200 import utils
    sd = ref_port.make_state(None, seed=seed, head_classes=23 if head else None)
    m = utils.get_ImageNet_or_random_ViT(head, 23, frozen, None, False, False, out_token=out_token)
    m.load_state_dict(sd, strict=True)
    m.precision = precision
    return m.cuda().train(), sd


def _inputs(B=2):
    g = torch.Generator().manual_seed(0)
    return torch.randn(B, 3, 224, 224, generator=g), torch.tensor([3, 17] * (B // 2)), torch.linspace(0.5, 2.0, 23)


@pytest.mark.parametrize("out_token", ["cls", "spatial"])
def test_cls_fp32_step_matches_golden(out_token, golden_dir):
    """fp32 validation mode, B=2, tolerance 1e-4 relative on logits / loss; every gradient against the fixture."""
    from ssl4gie_b200 import CrossEntropyLoss
    fix = np.load(os.path.join(golden_dir, f"step_cls_{out_token}_b2.npz"))
    m, sd = _build(out_token, "fp32")
    x, y, cw = _inputs()
    out = m(x.cuda())
    assert out.shape == (2, 23) and out.dtype == torch.float32
    loss = CrossEntropyLoss(cw).cuda()(out, y.cuda())
    loss.backward()
    assert _rel(out.detach().cpu(), torch.from_numpy(fix["logits"])) < 1e-4
    assert abs(loss.item() - float(fix["loss"])) < 1e-4 * abs(float(fix["loss"]))
    grads = {n: p.grad for n, p in m.named_parameters()}
    floor = 1e-3 * float(np.median(fix["grad_norm"]))
    worst = ("", 0.0)
    for i, n in enumerate(fix["grad_names"].tolist()):
        g = grads[n]
        assert g is not None, n
        rn = float(fix["grad_norm"][i])
        e = abs(g.double().norm().item() - rn) / max(rn, floor)
        k = min(8, g.numel())
        e2 = (g.flatten()[:k].cpu() - torch.from_numpy(fix["grad_head"][i][:k])).abs().max().item() / max(g.abs().max().item(), floor)
        if max(e, e2) > worst[1]:
            worst = (n, max(e, e2))
    assert worst[1] < 5e-3, worst
    assert sorted(n for n, p in m.named_parameters() if p.requires_grad and p.grad is None) == []


@pytest.mark.parametrize("out_token", ["cls", "spatial"])
def test_cls_fp32_step_matches_fp64_oracle(out_token):
    from ssl4gie_b200 import CrossEntropyLoss
    m, sd = _build(out_token, "fp32", seed=7)
    x, y, cw = _inputs(4)
    out = m(x.cuda())
    loss = CrossEntropyLoss(cw).cuda()(out, y.cuda())
    loss.backward()
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
    o2, l2, g2 = ref_port.loss_and_grads(sd64, x.double(), y, None, out_token=out_token, class_weights=cw.double())
    assert _rel(out.detach().cpu(), o2) < 1e-4
    assert abs(loss.item() - l2.item()) < 1e-4 * abs(l2.item())
    for n, p in m.named_parameters():
        assert _rel(p.grad.cpu(), g2[n]) < 2e-3 and _cos(p.grad.cpu(), g2[n]) > 0.9999, n


def test_cls_bf16_step_tracks_oracle():
    """bf16 tensor-core mode: 2e-2 relative on logits, cosine >= 0.999 on gradients (north_star tolerances)."""
    from ssl4gie_b200 import CrossEntropyLoss
    m, sd = _build("cls", "bf16")
    x, y, cw = _inputs(4)
    out = m(x.cuda())
    loss = CrossEntropyLoss(cw).cuda()(out, y.cuda())
    loss.backward()
    o2, l2, g2 = ref_port.loss_and_grads({k: v.cuda() for k, v in sd.items()}, x.cuda(), y.cuda(), None, out_token="cls",
                                         class_weights=cw.cuda())
    assert _rel(out.detach(), o2) < 2e-2
    assert abs(loss.item() - l2.item()) < 2e-2 * abs(l2.item())
    for n in ("lin_head.weight", "norm.weight", "blocks.11.mlp.fc2.weight", "blocks.0.attn.qkv.weight", "patch_embed.proj.weight"):
        assert _cos(m.get_parameter(n).grad, g2[n]) > 0.999, n


def test_cls_frozen_trains_only_the_head():
    """frozen=True: the encoder runs under no_grad (Models/models.py:341-343), only lin_head receives gradients."""
    from ssl4gie_b200 import CrossEntropyLoss
    m, sd = _build("cls", "fp32", frozen=True)
    x, y, cw = _inputs()
    loss = CrossEntropyLoss(cw).cuda()(m(x.cuda()), y.cuda())
    loss.backward()
    o2, l2, g2 = ref_port.loss_and_grads(sd, x, y, None, out_token="cls", class_weights=cw)
    assert abs(loss.item() - l2.item()) < 1e-4 * abs(l2.item())
    for n, p in m.named_parameters():
        if n.startswith("lin_head."):
            assert _rel(p.grad.cpu(), g2[n]) < 1e-3, n
        else:
            assert p.grad is None, n


def test_cls_headless_returns_pooled_features():
    m, sd = _build("spatial", "fp32", head=False)
    x, _, _ = _inputs()
    with torch.no_grad():
        out = m.eval()(x.cuda())
    ref = ref_port.forward(sd, x, None, False, out_token="spatial")
    assert out.shape == (2, 768) and _rel(out.cpu(), ref) < 1e-4


def test_cross_entropy_matches_torch():
    from ssl4gie_b200 import CrossEntropyLoss
    g = torch.Generator().manual_seed(1)
    for B, N, weighted in ((128, 23, True), (5, 23, False), (64, 1000, True)):
        x = (torch.randn(B, N, generator=g) * 3).cuda().requires_grad_(True)
        y = torch.randint(0, N, (B,), generator=g).cuda()
        w = (torch.rand(N, generator=g) + 0.5).cuda() if weighted else None
        loss = CrossEntropyLoss(w).cuda()(x, y)
        (loss * 2.5).backward()
        x2 = x.detach().double().requires_grad_(True)
        l2 = torch.nn.CrossEntropyLoss(w.double() if weighted else None)(x2, y)
        (l2 * 2.5).backward()
        assert abs(loss.item() - l2.item()) < 1e-5 * abs(l2.item())
        assert _rel(x.grad, x2.grad) < 1e-5
