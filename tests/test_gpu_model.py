"""Whole-model parity through the drop-in nn.Module API: logits, loss, every parameter gradient, BN running
statistics and the None-gradient set, against (a) the committed golden fixtures generated from the unmodified
reference and (b) the oracle restatement run live on the same inputs."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_port  # noqa: E402  (checker only)


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


def _build(dense, precision, backend="auto"):
    from ssl4gie_b200 import utils
    sd = ref_port.make_state(dense, seed=42)
    m = utils.get_ImageNet_or_random_ViT(False, 1 if dense == "seg" else None, False, dense, False, False)
    m.load_state_dict(sd, strict=True)
    m.precision = precision
    m.backend = backend
    if dense == "seg":
        m.decoder.output_conv[3].p = 0.0
    return m.cuda().train(), sd


def _loss(dense):
    from ssl4gie_b200 import ScaleAndShiftInvariantLoss, SoftDiceLoss
    return SoftDiceLoss() if dense == "seg" else ScaleAndShiftInvariantLoss(alpha=0.1)


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30)).item()


def _l2(a, b):
    return ((a.double() - b.double()).norm() / (b.double().norm() + 1e-300)).item()


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


@pytest.mark.parametrize("dense", ["seg", "depth"])
def test_fp32_step_matches_golden(dense, golden_dir):
    """BASELINE.json configs[0]: B=2, fp32 validation mode, tolerance 1e-4 relative (north_star)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    fix = np.load(os.path.join(golden_dir, f"step_{dense}_b2.npz"))
    m, sd = _build(dense, "fp32")
    x, t = _inputs(dense)
    out = m(x.cuda())
    loss = _loss(dense)(out, t.cuda())
    loss.backward()
    assert out.shape == (2, 1, 224, 224) and out.dtype == torch.float32
    ref_sub = torch.from_numpy(fix["out_sub"])
    assert _rel(out.detach().cpu()[:, :, ::8, ::8], ref_sub) < 1e-4
    assert abs(loss.item() - float(fix["loss"])) < 1e-4 * abs(float(fix["loss"]))
    grads = {n: p.grad for n, p in m.named_parameters()}
    none = sorted(n for n, p in m.named_parameters() if p.requires_grad and p.grad is None)
    assert none == sorted(fix["none_grads"].tolist())
    worst = ("", 0.0)
    for i, n in enumerate(fix["grad_names"].tolist()):
        g = grads[n]
        assert g is not None, n
        gn = g.double().norm().item()
        rn = float(fix["grad_norm"][i])
        if dense == "depth" and n.startswith("decoder.output_conv.4."):
            # sums of ~1e5 cancelling terms (the SSI loss is shift invariant and sigmoid' ~ const at init): for the bias
            # the fp64 oracle gives -4.2e-8 where the reference's own fp32 run gives -6.8e-6, and its fp32 weight
            # gradient is 70% off the fp64 one — these two fixture entries are rounding noise, so only the magnitude
            # is checked here (the tight check is against the fp64 oracle in the next test)
            assert gn < 5 * max(rn, 1e-5), (n, gn, rn)
            continue
        # gradients that are analytically ~0 (the last bias under the shift-invariant depth loss) are rounding noise
        # on both sides: measure against a floor tied to the typical gradient scale
        floor = 1e-3 * float(np.median(fix["grad_norm"]))
        e = abs(gn - rn) / max(rn, floor)
        k = min(8, g.numel())
        head = torch.from_numpy(fix["grad_head"][i][:k])
        e2 = (g.flatten()[:k].cpu() - head).abs().max().item() / max(g.abs().max().item(), floor)
        if max(e, e2) > worst[1]:
            worst = (n, max(e, e2))
    # per-tensor gradients of the CPU-fp32 fixture are themselves only ~1e-3 accurate for the cancellation-heavy
    # BatchNorm/LayerNorm affine gradients; the tight bound (vs an fp64 run) is test_fp32_step_matches_oracle_*.
    # depth: the 2x2 scale/shift solve at random init is so ill-conditioned that the reference's own fp32 gradients
    # are 36% (relative L2) off its fp64 ones for most tensors (measured with oracle/ref_port.py on the CPU) and
    # the fp32 summation order decides the value: the fixture pins logits, loss and the None-gradient set, the
    # gradients only to that noise level
    assert worst[1] < (5e-3 if dense == "seg" else 1.0), worst
    if dense == "seg":
        msd = m.state_dict()
        for i, k in enumerate(fix["bn_names"].tolist()):
            assert _rel(msd[k].cpu(), torch.from_numpy(fix["bn_running_mean"][i])) < 1e-4 or \
                (msd[k].cpu() - torch.from_numpy(fix["bn_running_mean"][i])).abs().max() < 1e-6, k
            kv = k.replace("running_mean", "running_var")
            assert _rel(msd[kv].cpu(), torch.from_numpy(fix["bn_running_var"][i])) < 1e-4, kv
            assert int(msd[k.replace("running_mean", "num_batches_tracked")]) == int(fix["bn_tracked"][i])


@pytest.mark.parametrize("dense", ["seg", "depth"])
def test_fp32_step_matches_oracle_with_dropout_mask(dense):
    """Same step against the live oracle, B=3 (odd batch), seg with an explicit Dropout(0.1) keep-mask."""
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    m, sd = _build(dense, "fp32")
    x, t = _inputs(dense, 3)
    mask_nchw = None
    if dense == "seg":
        m.decoder.output_conv[3].p = 0.1
        g = torch.Generator().manual_seed(9)
        mask_nchw = (torch.rand(3, 256, 112, 112, generator=g) >= 0.1)
        m._parity_dropout_mask = mask_nchw.permute(0, 2, 3, 1).contiguous().to(torch.uint8).cuda()
    out = m(x.cuda())
    loss = _loss(dense)(out, t.cuda())
    loss.backward()
    sdg = {k: v.cuda() for k, v in sd.items()}
    mk = mask_nchw.cuda() if mask_nchw is not None else None
    o32, l32, g32 = ref_port.loss_and_grads(sdg, x.cuda(), t.cuda(), dense, mk, 0.1)
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sdg.items()}
    o64, l64, g64 = ref_port.loss_and_grads(sd64, x.cuda().double(), t.cuda().double(), dense, mk, 0.1)
    # The bar.  Logits and loss: 1e-4 relative to the exact (fp64) reference result.  Gradients: train-mode
    # BatchNorm (seg) and the 2x2 scale/shift solve (depth) make several gradients ill-conditioned — the
    # reference's OWN fp32 evaluation is only 1e-3..1e-1 accurate on them (measured against fp64) — so a tensor
    # passes at 1e-4 relative L2, or when it is no further from the fp64 truth than 3x the reference-fp32 run.
    assert _rel(out.detach(), o64) < max(1e-4, 3 * _rel(o32, o64))
    assert abs(loss.item() - l64.item()) < 1e-4 * abs(l64.item())
    bad = []
    for n, p in m.named_parameters():
        if g64.get(n) is None:
            assert p.grad is None, n
            continue
        e, e_ref = _l2(p.grad, g64[n]), _l2(g32[n], g64[n])
        if e > max(1e-4, 3 * e_ref):
            bad.append((n, e, e_ref))
    assert not bad, bad[:10]


@pytest.mark.parametrize("backend", ["simt", "tensor"])
@pytest.mark.parametrize("dense", ["seg", "depth"])
def test_bf16_step_matches_oracle(dense, backend):
    """bf16 mode: 2e-2 relative on logits and gradients (north_star), cosine >= 0.999 on gradients."""
    m, sd = _build(dense, "bf16", backend)
    x, t = _inputs(dense, 4)
    out = m(x.cuda())
    loss = _loss(dense)(out, t.cuda())
    loss.backward()
    sdg = {k: v.cuda() for k, v in sd.items()}
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sdg.items()}
    o64, l64, g64 = ref_port.loss_and_grads(sd64, x.cuda().double(), t.cuda().double(), dense, None, 0.1)
    # the reference run the way it trains: the same modules under bf16 autocast on this GPU
    with torch.autocast("cuda", dtype=torch.bfloat16):
        params = {k: v.detach().clone().requires_grad_(True) for k, v in ref_port.trainable(sdg).items()}
        full = dict(sdg)
        full.update(params)
        ob = ref_port.forward(full, x.cuda(), dense, True, None)
        lb = ref_port.soft_dice_loss(ob, t.cuda()) if dense == "seg" else ref_port.ssi_loss(ob, t.cuda(), 0.1)
    lb.backward()
    # logits / loss: 2e-2 relative (north_star)
    assert _rel(out.detach(), o64) < 2e-2
    assert abs(loss.item() - l64.item()) < 1e-2 * abs(l64.item())
    # gradients: 2e-2 relative L2 / cosine 0.999 against the exact result where bf16 allows it; where the
    # reference's own bf16-autocast run cannot reach that (it does not at B=4 with random init: measured cosine
    # ~0.985 for seg and negative for depth), we must be at least as close to the truth as that run.
    # The plain criterion on a well-conditioned state is tests/test_gpu_wellcond.py; here the count is printed.
    bad, strict, strict_ref, total = [], 0, 0, 0
    for n, p in m.named_parameters():
        if g64.get(n) is None:
            continue
        e, e_ref = _l2(p.grad, g64[n]), _l2(params[n].grad, g64[n])
        c, c_ref = _cos(p.grad, g64[n]), _cos(params[n].grad, g64[n])
        total += 1
        strict += e <= 2e-2 and c >= 0.999
        strict_ref += e_ref <= 2e-2 and c_ref >= 0.999
        if e > max(2e-2, 1.25 * e_ref) or c < min(0.999, c_ref - 0.01):
            bad.append((n, e, e_ref, c, c_ref))
    print(f"[{dense}/{backend} random init B=4] strict_pass ours {strict}/{total}, reference bf16 autocast {strict_ref}/{total}")
    assert not bad, (len(bad), bad[:10])


def test_eval_mode_and_state_dict_roundtrip(tmp_path):
    m, sd = _build("seg", "fp32")
    x, _ = _inputs("seg")
    m.eval()
    with torch.no_grad():
        out = m(x.cuda())
    sdg = {k: v.cuda() for k, v in sd.items()}
    ref = ref_port.forward(sdg, x.cuda(), "seg", training=False)
    assert _rel(out, ref) < 1e-4
    path = tmp_path / "ckpt.pth"
    torch.save({"model_state_dict": m.state_dict()}, path)
    from ssl4gie_b200 import utils
    m2 = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
    m2.load_state_dict(torch.load(path)["model_state_dict"], strict=True)
    m2.precision = "fp32"
    m2 = m2.cuda().eval()
    with torch.no_grad():
        out2 = m2(x.cuda())
    assert torch.equal(out, out2)


def test_frozen_encoder():
    from ssl4gie_b200 import SoftDiceLoss, utils
    sd = ref_port.make_state("seg", seed=42)
    m = utils.get_ImageNet_or_random_ViT(False, 1, True, "seg", False, False)
    m.load_state_dict(sd, strict=True)
    m.precision = "fp32"
    m.decoder.output_conv[3].p = 0.0
    m = m.cuda().train()
    x, t = _inputs("seg")
    loss = SoftDiceLoss()(m(x.cuda()), t.cuda())
    loss.backward()
    for n, p in m.named_parameters():
        if not n.startswith("decoder."):
            assert p.grad is None, n  # encoder under no_grad (Models/models.py:341-343)
    sdg = {k: v.cuda() for k, v in sd.items()}
    _, l2, g2 = ref_port.loss_and_grads(sdg, x.cuda(), t.cuda(), "seg", None)
    assert abs(loss.item() - l2.item()) < 1e-4 * abs(l2.item())
    n = "decoder.act_postprocess12.0.weight"
    assert _l2(m.get_parameter(n).grad, g2[n]) < 1e-2 and _cos(m.get_parameter(n).grad, g2[n]) > 0.9999


def test_wrong_input_size_and_cpu_input_raise():
    m, _ = _build("depth", "fp32")
    with pytest.raises(AssertionError):
        m(torch.randn(1, 3, 192, 192).cuda())
    with pytest.raises(RuntimeError):
        m(torch.randn(1, 3, 224, 224))


def test_device_prefetcher_yields_the_batches_in_order():
    """Ring of depth + 1 reused staging sets: every batch arrives intact and in order even when the consumer is slow
    (the copy of batch i + 3 must wait for the step that read batch i), including a ragged last batch."""
    from ssl4gie_b200.prefetch import DevicePrefetcher
    g = torch.Generator().manual_seed(3)
    host = [(torch.randn(4, 3, 64, 64, generator=g).pin_memory(), torch.randint(0, 5, (4,), generator=g).pin_memory()) for _ in range(9)]
    host.append((torch.randn(3, 3, 64, 64, generator=g).pin_memory(), torch.randint(0, 5, (3,), generator=g).pin_memory()))
    sums, labels, ptrs = [], [], set()
    for x, t in DevicePrefetcher(iter(host)):
        assert x.is_cuda and t.is_cuda
        ptrs.add(x.data_ptr())
        torch.cuda._sleep(20_000_000)  # ~10 ms: the host is several batches ahead of the device
        sums.append(x.double().sum())
        labels.append(t.clone())
    torch.cuda.synchronize()
    assert len(sums) == 10 and len(ptrs) <= 4  # 3 ring slots + the ragged batch
    for s_d, t_d, (xh, th) in zip(sums, labels, host):
        assert abs(s_d.item() - xh.double().sum().item()) < 1e-9 and torch.equal(t_d.cpu(), th)
    single = list(x.clone() for x in DevicePrefetcher(iter([h[0] for h in host[:3]]), depth=1))
    assert all(torch.equal(a.cpu(), h[0]) for a, h in zip(single, host))

