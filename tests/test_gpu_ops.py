"""Operator-level parity (through the C ABI) against the oracle / plain torch fp32 on the same seeded inputs."""
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_port  # noqa: E402  (checker only)


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30)).item()


@pytest.fixture(scope="module")
def C():
    from ssl4gie_b200 import _C
    _C.init(0)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return _C


def _dt(C, dtype):
    return (C.F32, torch.float32) if dtype == "fp32" else (C.BF16, torch.bfloat16)


# ------------------------------------------------------------------------------------------------ losses
def test_dice_kats(C, golden_dir):
    from ssl4gie_b200 import SoftDiceLoss
    kat = json.load(open(os.path.join(golden_dir, "loss_kats.json")))
    logits = torch.linspace(-3, 3, 128, dtype=torch.float64).view(2, 1, 8, 8).float().cuda().requires_grad_(True)
    targets = (torch.arange(128) % 3 == 0).float().view(2, 1, 8, 8).cuda()
    loss = SoftDiceLoss()(logits, targets)
    loss.backward()
    assert abs(loss.item() - kat["dice_loss"]) < 1e-6
    g = torch.tensor(kat["dice_grad"])
    assert _rel(logits.grad.flatten().cpu(), g) < 1e-4
    z = SoftDiceLoss()(torch.zeros(2, 1, 4, 4).cuda(), torch.ones(2, 1, 4, 4).cuda())
    assert abs(z.item() - kat["dice_zeros_ones"]) < 1e-6


@pytest.mark.parametrize("alpha", [0.1, 0.0])
def test_ssi_kats(C, golden_dir, alpha):
    from ssl4gie_b200 import ScaleAndShiftInvariantLoss
    kat = json.load(open(os.path.join(golden_dir, "loss_kats.json")))
    i = torch.arange(512, dtype=torch.float64).view(2, 1, 16, 16)
    pred = (0.5 + 0.4 * torch.sin(0.37 * i)).float()
    targ = (0.5 + 0.45 * torch.cos(0.11 * i)).float()
    targ[(torch.arange(512) % 7 == 0).view(2, 1, 16, 16)] = 0
    p = pred.cuda().requires_grad_(True)
    fn = ScaleAndShiftInvariantLoss(alpha=alpha)
    loss = fn(p, targ.cuda())
    loss.backward()
    assert abs(loss.item() - kat[f"ssi_loss_a{alpha}"]) < 2e-6
    g = torch.tensor(kat[f"ssi_grad_a{alpha}"])
    assert _rel(p.grad.flatten().cpu(), g) < 2e-4
    # prediction_ssi property (losses.py:148-151)
    s = torch.tensor(kat["ssi_scale"]).view(2, 1, 1)
    sh = torch.tensor(kat["ssi_shift"]).view(2, 1, 1)
    assert _rel(fn.prediction_ssi.cpu(), s * pred.squeeze(1) + sh) < 1e-5


def test_ssi_random_and_empty_mask(C, golden_dir):
    from ssl4gie_b200 import ScaleAndShiftInvariantLoss
    kat = json.load(open(os.path.join(golden_dir, "loss_kats.json")))
    g = torch.Generator().manual_seed(3)
    p = torch.rand(3, 1, 32, 32, generator=g)
    t = torch.rand(3, 1, 32, 32, generator=g)
    t[torch.rand(3, 1, 32, 32, generator=g) < 0.2] = 0
    t[2] = 0
    pp = p.cuda().requires_grad_(True)
    loss = ScaleAndShiftInvariantLoss(alpha=0.1)(pp, t.cuda())
    loss.backward()
    assert abs(loss.item() - kat["ssi_rand_loss"]) < 2e-6
    assert abs(pp.grad.double().norm().item() - kat["ssi_rand_grad_norm"]) < 2e-4 * kat["ssi_rand_grad_norm"]
    assert _rel(pp.grad.flatten()[:16].cpu(), torch.tensor(kat["ssi_rand_grad_head"])) < 5e-4
    assert pp.grad[2].abs().max().item() == 0.0  # all-masked image contributes nothing
    # everything masked: loss is exactly 0 (reduction_batch_based returns 0, losses.py:34-35)
    z = ScaleAndShiftInvariantLoss(alpha=0.1)(p.cuda(), torch.zeros_like(p).cuda())
    assert z.item() == 0.0


@pytest.mark.parametrize("tag", ["img", "img_a0", "s6", "s5_img"])
def test_ssi_image_based_and_more_scales(C, golden_dir, tag):
    """reduction="image-based" and scales > 4 against known answers from the reference's own loss code."""
    from ssl4gie_b200 import ScaleAndShiftInvariantLoss
    from tests.test_oracle import _ssi_ext_cases
    kat = json.load(open(os.path.join(golden_dir, "loss_kats.json")))
    p, t, kw = _ssi_ext_cases()[tag]
    pp = p.cuda().requires_grad_(True)
    loss = ScaleAndShiftInvariantLoss(**kw)(pp, t.cuda())
    loss.backward()
    assert abs(loss.item() - kat[f"ssi_{tag}_loss"]) < 2e-6
    assert abs(pp.grad.double().norm().item() - kat[f"ssi_{tag}_grad_norm"]) < 2e-4 * kat[f"ssi_{tag}_grad_norm"]
    assert _rel(pp.grad.flatten()[:16].cpu(), torch.tensor(kat[f"ssi_{tag}_grad_head"])) < 5e-4


def test_losses_are_bit_reproducible(C):
    """Two-stage fixed-order reductions: the same inputs give bit-identical loss and gradient, run after run."""
    from ssl4gie_b200 import ScaleAndShiftInvariantLoss, SoftDiceLoss
    g = torch.Generator().manual_seed(11)
    x = torch.randn(16, 1, 224, 224, generator=g).cuda()
    t = torch.rand(16, 1, 224, 224, generator=g).cuda()
    t[t < 0.1] = 0
    for fn, a, b in ((SoftDiceLoss(), x, (t > 0.5).float()), (ScaleAndShiftInvariantLoss(alpha=0.1), torch.sigmoid(x), t)):
        outs = []
        for _ in range(3):
            p = a.clone().requires_grad_(True)
            loss = fn(p, b)
            loss.backward()
            outs.append((loss.detach().clone(), p.grad.clone()))
        assert all(torch.equal(outs[0][0], o[0]) and torch.equal(outs[0][1], o[1]) for o in outs[1:])


def test_losses_full_size_vs_oracle(C):
    from ssl4gie_b200 import ScaleAndShiftInvariantLoss, SoftDiceLoss
    g = torch.Generator().manual_seed(5)
    B = 8
    logits = (torch.randn(B, 1, 224, 224, generator=g) * 2).cuda()
    tgt = (torch.rand(B, 1, 224, 224, generator=g) > 0.5).float().cuda()
    a = logits.clone().requires_grad_(True)
    b = logits.clone().requires_grad_(True)
    la = SoftDiceLoss()(a, tgt)
    lb = ref_port.soft_dice_loss(b, tgt)
    la.backward(); lb.backward()
    assert abs(la.item() - lb.item()) < 1e-5
    assert _rel(a.grad, b.grad) < 1e-4
    pred = torch.sigmoid(torch.randn(B, 1, 224, 224, generator=g)).cuda()
    dt = torch.rand(B, 1, 224, 224, generator=g)
    dt[torch.rand(B, 1, 224, 224, generator=g) < 0.1] = 0
    dt[:, :, :8] = 0
    dt = dt.cuda()
    a = pred.clone().requires_grad_(True)
    b = pred.clone().requires_grad_(True)
    la = ScaleAndShiftInvariantLoss(alpha=0.1)(a, dt)
    lb = ref_port.ssi_loss(b, dt, 0.1)
    la.backward(); lb.backward()
    assert abs(la.item() - lb.item()) < 1e-4 * abs(lb.item()) + 1e-6
    assert _rel(a.grad, b.grad) < 2e-3  # fp32 reference solve is itself ill-conditioned (a00*a11 - a01^2)


# ------------------------------------------------------------------------------------------------ contractions
@pytest.mark.parametrize("backend", ["simt", "tensor"])
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
@pytest.mark.parametrize("M,N,K", [(394, 2304, 768), (12608, 768, 3072), (1000, 96, 768), (392, 1536, 96)])
def test_linear_fwd(C, backend, dtype, M, N, K):
    if backend == "tensor" and dtype == "fp32":
        pytest.skip("tensor backend is bf16")
    code, tdt = _dt(C, dtype)
    g = torch.Generator().manual_seed(0)
    A = torch.randn(M, K, generator=g).cuda().to(tdt)
    W = (torch.randn(N, K, generator=g) * 0.05).cuda().to(tdt)
    bias = torch.randn(N, generator=g).cuda()
    res = torch.randn(M, N, generator=g).cuda()
    for act, use_res, out_f32 in ((0, False, 0), (1, False, 0), (0, True, 1)):
        out = torch.empty(M, N, device="cuda", dtype=torch.float32 if out_f32 else tdt)
        C.check(C.lib().s4g_linear_fwd(C.ptr(A), C.ptr(W), C.ptr(bias), C.ptr(res) if use_res else None, C.ptr(out), M, N, K,
                                       act, out_f32, code, C.TENSOR if backend == "tensor" else C.SIMT, C.stream_ptr()))
        ref = A.float() @ W.float().t() + bias
        if act == 1:
            ref = torch.nn.functional.gelu(ref)
        if use_res:
            ref = ref + res
        tol = 1e-4 if dtype == "fp32" else 1e-2
        assert _rel(out.float(), ref) < tol, (act, use_res)


@pytest.mark.parametrize("backend", ["simt", "tensor"])
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
@pytest.mark.parametrize("M,N,K", [(394, 2304, 768), (12608, 768, 768), (6272, 1536, 96)])
def test_linear_wgrad(C, backend, dtype, M, N, K):
    if backend == "tensor" and dtype == "fp32":
        pytest.skip("tensor backend is bf16")
    code, tdt = _dt(C, dtype)
    g = torch.Generator().manual_seed(1)
    dY = torch.randn(M, N, generator=g).cuda().to(tdt)
    X = torch.randn(M, K, generator=g).cuda().to(tdt)
    dW = torch.empty(N, K, device="cuda")
    nb = C.lib().s4g_wgrad_scratch_bytes(M, 1, 1, K, N, 1)
    scratch = torch.empty(max(nb, 16), dtype=torch.uint8, device="cuda")
    C.check(C.lib().s4g_linear_wgrad(C.ptr(dY), C.ptr(X), C.ptr(dW), M, N, K, code,
                                     C.TENSOR if backend == "tensor" else C.SIMT, C.ptr(scratch), nb, C.stream_ptr()))
    ref = dY.float().t() @ X.float()
    assert _rel(dW, ref) < (1e-4 if dtype == "fp32" else 5e-3)


@pytest.mark.parametrize("backend", ["simt", "tensor"])
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
@pytest.mark.parametrize("B,H,Cin,Cout", [(2, 7, 768, 256), (2, 14, 256, 256), (3, 28, 192, 256), (2, 56, 96, 256),
                                          (2, 112, 256, 128), (1, 224, 128, 32)])
def test_conv3x3(C, backend, dtype, B, H, Cin, Cout):
    if backend == "tensor" and dtype == "fp32":
        pytest.skip("tensor backend is bf16")
    code, tdt = _dt(C, dtype)
    be = C.TENSOR if backend == "tensor" else C.SIMT
    g = torch.Generator().manual_seed(2)
    x = torch.randn(B, Cin, H, H, generator=g).cuda()
    w = (torch.randn(Cout, Cin, 3, 3, generator=g) * 0.05).cuda()
    bias = torch.randn(Cout, generator=g).cuda()
    xq, wq = x.to(tdt).float(), w.to(tdt).float()
    ref = torch.nn.functional.conv2d(xq, wq, bias, padding=1)
    x_nhwc = xq.permute(0, 2, 3, 1).contiguous().to(tdt)
    wp = wq.permute(0, 2, 3, 1).contiguous().to(tdt)  # [Co][r][s][Ci]
    out = torch.empty(B, H, H, Cout, device="cuda", dtype=tdt)
    C.check(C.lib().s4g_conv3x3_fwd(C.ptr(x_nhwc), C.ptr(wp), C.ptr(bias), C.ptr(out), B, H, H, Cin, Cout, code, be, C.stream_ptr()))
    assert _rel(out.float().permute(0, 3, 1, 2), ref) < (1e-4 if dtype == "fp32" else 1e-2)
    # wgrad
    dy = torch.randn(B, Cout, H, H, generator=g).cuda().to(tdt).float()
    dw_ref = torch.nn.grad.conv2d_weight(xq, w.shape, dy, padding=1)
    dy_nhwc = dy.permute(0, 2, 3, 1).contiguous().to(tdt)
    dwp = torch.empty(Cout, 3, 3, Cin, device="cuda")
    nb = C.lib().s4g_wgrad_scratch_bytes(B, H, H, Cin, Cout, 9)
    scratch = torch.empty(max(nb, 16), dtype=torch.uint8, device="cuda")
    C.check(C.lib().s4g_conv3x3_wgrad(C.ptr(dy_nhwc), C.ptr(x_nhwc), C.ptr(dwp), B, H, H, Cin, Cout, code, be,
                                      C.ptr(scratch), nb, C.stream_ptr()))
    assert _rel(dwp.permute(0, 3, 1, 2), dw_ref) < (1e-4 if dtype == "fp32" else 5e-3)


# ------------------------------------------------------------------------------------------------ norm / attention / resample
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
def test_layernorm(C, dtype):
    code, tdt = _dt(C, dtype)
    g = torch.Generator().manual_seed(3)
    rows = 2 * 197
    x = (torch.randn(rows, 768, generator=g) * 2 + 0.3).cuda()
    gamma = (1 + 0.1 * torch.randn(768, generator=g)).cuda()
    beta = (0.1 * torch.randn(768, generator=g)).cuda()
    y = torch.empty(rows, 768, device="cuda", dtype=tdt)
    mean = torch.empty(rows, device="cuda"); rstd = torch.empty(rows, device="cuda")
    C.check(C.lib().s4g_layernorm_fwd(C.ptr(x), C.ptr(gamma), C.ptr(beta), C.ptr(y), C.ptr(mean), C.ptr(rstd), rows, 768, 1e-6, code, C.stream_ptr()))
    xr = x.clone().requires_grad_(True); gr = gamma.clone().requires_grad_(True); br = beta.clone().requires_grad_(True)
    ref = torch.nn.functional.layer_norm(xr, (768,), gr, br, 1e-6)
    assert _rel(y.float(), ref) < (1e-5 if dtype == "fp32" else 1e-2)
    dy = torch.randn(rows, 768, generator=g).cuda().to(tdt)
    dres = torch.randn(rows, 768, generator=g).cuda()
    ref.backward(dy.float())
    dx = torch.empty_like(x); dg = torch.empty(768, device="cuda"); db = torch.empty(768, device="cuda")
    C.check(C.lib().s4g_layernorm_bwd(C.ptr(dy), C.ptr(x), C.ptr(gamma), C.ptr(mean), C.ptr(rstd), C.ptr(dres), C.ptr(dx), C.ptr(dg),
                                      C.ptr(db), rows, 768, code, C.stream_ptr()))
    assert _rel(dx, xr.grad + dres) < 1e-4
    assert _rel(dg, gr.grad) < 1e-4
    assert _rel(db, br.grad) < 1e-4


@pytest.mark.parametrize("backend", ["simt", "tensor"])
@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
def test_attention(C, backend, dtype):
    if backend == "tensor" and dtype == "fp32":
        pytest.skip("tensor backend is bf16")
    code, tdt = _dt(C, dtype)
    be = C.TENSOR if backend == "tensor" else C.SIMT
    g = torch.Generator().manual_seed(4)
    B, N, H = 3, 197, 12
    qkv = torch.randn(B * N, 3 * H * 64, generator=g).cuda().to(tdt)
    o = torch.empty(B * N, H * 64, device="cuda", dtype=tdt)
    lse = torch.empty(B, H, N, device="cuda")
    C.check(C.lib().s4g_attention_fwd(C.ptr(qkv), C.ptr(o), C.ptr(lse), B, N, H, code, be, C.stream_ptr()))
    qr = qkv.float().clone().requires_grad_(True)
    q, k, v = qr.view(B, N, 3, H, 64).permute(2, 0, 3, 1, 4).unbind(0)
    attn = ((q @ k.transpose(-2, -1)) * 0.125).softmax(-1)
    ref = (attn @ v).transpose(1, 2).reshape(B * N, H * 64)
    tol = 1e-4 if dtype == "fp32" else 2e-2
    assert _rel(o.float(), ref) < tol
    do = torch.randn(B * N, H * 64, generator=g).cuda().to(tdt)
    ref.backward(do.float())
    dqkv = torch.empty_like(qkv)
    C.check(C.lib().s4g_attention_bwd(C.ptr(qkv), C.ptr(o), C.ptr(do), C.ptr(lse), C.ptr(dqkv), B, N, H, code, be, C.stream_ptr()))
    assert _rel(dqkv.float(), qr.grad) < tol


@pytest.mark.parametrize("dtype", ["fp32", "bf16"])
@pytest.mark.parametrize("H", [7, 14, 56])
def test_bilinear(C, dtype, H):
    code, tdt = _dt(C, dtype)
    g = torch.Generator().manual_seed(5)
    B, Cc = 2, 256
    x = torch.randn(B, Cc, H, H, generator=g).cuda().to(tdt).float().requires_grad_(True)
    ref = torch.nn.functional.interpolate(x, scale_factor=2, mode="bilinear", align_corners=True)
    xn = x.detach().permute(0, 2, 3, 1).contiguous().to(tdt)
    y = torch.empty(B, 2 * H, 2 * H, Cc, device="cuda", dtype=tdt)
    C.check(C.lib().s4g_bilinear2x_fwd(C.ptr(xn), C.ptr(y), B, H, H, Cc, code, C.stream_ptr()))
    tol = 1e-5 if dtype == "fp32" else 1e-2
    assert _rel(y.float().permute(0, 3, 1, 2), ref) < tol
    dy = torch.randn(B, Cc, 2 * H, 2 * H, generator=g).cuda().to(tdt).float()
    ref.backward(dy)
    dx = torch.empty(B, H, H, Cc, device="cuda", dtype=tdt)
    C.check(C.lib().s4g_bilinear2x_bwd(C.ptr(dy.permute(0, 2, 3, 1).contiguous().to(tdt)), C.ptr(dx), B, H, H, Cc, code, C.stream_ptr()))
    assert _rel(dx.float().permute(0, 3, 1, 2), x.grad) < tol


def test_adamw(C):
    g = torch.Generator().manual_seed(6)
    n = 4096
    p = torch.randn(n, generator=g).cuda(); p2 = p.clone()
    m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
    m2 = m.clone(); v2 = v.clone()
    for step in range(1, 4):
        gr = torch.randn(n, generator=g).cuda()
        C.check(C.lib().s4g_adamw_step(C.ptr(p), C.ptr(gr), C.ptr(m), C.ptr(v), n, 1e-4, 0.9, 0.999, 1e-8, 0.01, step, 1.0, C.stream_ptr()))
        ref_port.adamw_step(p2, gr, m2, v2, step)
    assert _rel(p, p2) < 1e-6
