"""Developer diagnostic: per-tensor bf16 gradient error on the well-conditioned state of tests/_states.py, next to the
reference's own bf16-autocast error.  Usage: python scripts/dev_wellcond.py [seg|depth] [B] [steps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_port  # noqa: E402
from ssl4gie_b200 import ScaleAndShiftInvariantLoss, SoftDiceLoss, utils  # noqa: E402
from tests import _states  # noqa: E402


def l2(a, b):
    return ((a.double() - b.double()).norm() / (b.double().norm() + 1e-300)).item()


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


def main():
    dense = sys.argv[1] if len(sys.argv) > 1 else "seg"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sd0 = ref_port.make_state(dense, seed=42)
    x = _states.structured_images(B).cuda()
    t = (_states.blob_masks(B) if dense == "seg" else _states.smooth_depth(B)).cuda()
    sd = _states.warm_state(dense, sd0, x, t, steps)
    m = utils.get_ImageNet_or_random_ViT(False, 1 if dense == "seg" else None, False, dense, False, False)
    m.load_state_dict({k: v.cpu() for k, v in sd.items()}, strict=True)
    if dense == "seg":
        m.decoder.output_conv[3].p = 0.0
    m.precision, m.backend = "bf16", "tensor"
    m = m.cuda().train()
    loss_fn = SoftDiceLoss() if dense == "seg" else ScaleAndShiftInvariantLoss(alpha=0.1)
    out = m(x)
    loss = loss_fn(out, t)
    loss.backward()
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
    o64, l64, g64 = ref_port.loss_and_grads(sd64, x.double(), t.double(), dense, None, 0.1)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        params = {k: v.detach().clone().requires_grad_(True) for k, v in ref_port.trainable(sd).items()}
        full = dict(sd)
        full.update(params)
        ob = ref_port.forward(full, x, dense, True, None)
        lb = ref_port.soft_dice_loss(ob, t) if dense == "seg" else ref_port.ssi_loss(ob, t, 0.1)
    lb.backward()
    rel = lambda a, b: ((a.double() - b.double()).abs().max() / b.double().abs().max()).item()  # noqa: E731
    print(f"{dense} B={B} warm={steps}: loss ours {loss.item():.6f} fp64 {l64.item():.6f} autocast {lb.item():.6f}; "
          f"logits rel ours {rel(out.detach(), o64):.3e} autocast {rel(ob.float(), o64):.3e}")
    rows = []
    for n, p in m.named_parameters():
        if g64.get(n) is None:
            continue
        rows.append((l2(p.grad, g64[n]), cos(p.grad, g64[n]), l2(params[n].grad, g64[n]), cos(params[n].grad, g64[n]), g64[n].norm().item(), n))
    strict = sum(1 for r in rows if r[0] <= 2e-2 and r[1] >= 0.999)
    strict_ref = sum(1 for r in rows if r[2] <= 2e-2 and r[3] >= 0.999)
    print(f"strict pass (l2 <= 2e-2 and cos >= 0.999): ours {strict}/{len(rows)}, reference autocast {strict_ref}/{len(rows)}")
    relinf = lambda a, b: ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-300)).item()  # noqa: E731
    inf_pass = sum(1 for n, p in m.named_parameters() if g64.get(n) is not None and relinf(p.grad, g64[n]) <= 2e-2)
    print(f"max-norm criterion (|a-b|inf / |b|inf <= 2e-2): ours {inf_pass}/{len(rows)}")
    rows.sort(reverse=True)
    for r in rows[:25]:
        print("l2 %.3e cos %.6f | autocast l2 %.3e cos %.6f | |g| %.3e %s" % r)


if __name__ == "__main__":
    main()
