"""Parity at BASELINE.json's full size (configs[1]/[2]: 64 images per GPU, bf16 tensor-core path -- the exact handle
bench.py times): one training step against the oracle restatement run on the same GPU in fp32 (TF32 off) and, as the
yardstick for what bf16 can reach, against the same oracle under bf16 autocast; plus the
size-independent property the path offers: without BatchNorm (depth) an image's output does not depend on its batch."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_port  # noqa: E402  (checker only)

from .test_gpu_model import _build, _cos, _inputs, _l2, _loss, _rel  # noqa: E402

B = 64


def _oracle(sdg, x, t, dense, autocast):
    params = {k: v.detach().clone().requires_grad_(True) for k, v in ref_port.trainable(sdg).items()}
    full = dict(sdg)
    full.update(params)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        out = ref_port.forward(full, x, dense, True, None)
        loss = ref_port.soft_dice_loss(out, t) if dense == "seg" else ref_port.ssi_loss(out, t, 0.1)
    loss.backward()
    grads = {k: v.grad for k, v in params.items() if v.grad is not None}
    return out.detach().float(), loss.detach().float(), grads


@pytest.mark.parametrize("dense", ["seg", "depth"])
def test_full_size_bf16_step_tracks_oracle(dense):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    m, sd = _build(dense, "bf16", "tensor")
    x, t = _inputs(dense, B)
    x, t = x.cuda(), t.cuda()
    out = m(x)
    loss = _loss(dense)(out, t)
    loss.backward()
    ours = {n: p.grad.clone() for n, p in m.named_parameters() if p.grad is not None}
    out, loss = out.detach().clone(), loss.detach().clone()
    sdg = {k: v.cuda() for k, v in sd.items()}
    o32, l32, g32 = _oracle(sdg, x, t, dense, False)
    torch.cuda.empty_cache()
    ob, lb, gb = _oracle(sdg, x, t, dense, True)
    torch.cuda.empty_cache()
    assert sorted(ours) == sorted(g32)  # same None-gradient set at full size
    assert _rel(out, o32) < 2e-2, _rel(out, o32)
    assert abs(loss.item() - l32.item()) < 1e-2 * abs(l32.item()), (loss.item(), l32.item())
    bad, strict, strict_ref = [], 0, 0
    for n, g in ours.items():
        e, e_ref = _l2(g, g32[n]), _l2(gb[n], g32[n])
        c, c_ref = _cos(g, g32[n]), _cos(gb[n], g32[n])
        strict += e <= 2e-2 and c >= 0.999
        strict_ref += e_ref <= 2e-2 and c_ref >= 0.999
        # 2e-2 / cosine 0.999 where bf16 allows it, else at least as close to the fp32 result as the reference's own
        # bf16-autocast run on this GPU
        if g.numel() == 1:
            # a scalar (depth head's last bias): its gradient is a sum over all pixels that cancels almost completely
            # (the scale-and-shift-invariant loss ignores a global shift), so bf16 noise decides even its sign -- the
            # reference's own autocast run is off by 61 %; hold it to twice that run's error, cosine is meaningless
            if e > max(2e-2, 2.0 * e_ref):
                bad.append((n, round(e, 4), round(e_ref, 4)))
            continue
        if e > max(2e-2, 1.25 * e_ref) or c < min(0.999, c_ref - 0.01):
            bad.append((n, round(e, 4), round(e_ref, 4), round(c, 5), round(c_ref, 5)))
    print(f"[{dense} random init B={B}] strict_pass ours {strict}/{len(ours)}, reference bf16 autocast {strict_ref}/{len(ours)}")
    assert not bad, (len(bad), bad[:10])


def test_full_size_depth_output_is_batch_independent():
    """No BatchNorm on the depth path: image i of the 64-image step equals the same image run in a batch of 2 (a
    different handle, different tile counts and split-K plans)."""
    m, _ = _build("depth", "bf16", "tensor")
    x, _ = _inputs("depth", B)
    x = x.cuda()
    with torch.no_grad():
        big = m(x).clone()
        small = m(x[10:12]).clone()
    assert _rel(big[10:12], small) < 1e-2, _rel(big[10:12], small)
