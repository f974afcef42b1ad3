"""Developer diagnostic: per-tensor bf16 gradient error on the warmed-up states of tests/_states.py, next to the
reference's own bf16-autocast error.  Usage: python tests/tools/dev_wellcond.py [seg|depth] [B] [warm steps ...]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import ref_port  # noqa: E402
from ssl4gie_b200 import ScaleAndShiftInvariantLoss, SoftDiceLoss, utils  # noqa: E402
from tests import _states  # noqa: E402


def l2(a, b):
    return ((a.double() - b.double()).norm() / (b.double().norm() + 1e-300)).item()


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return (a @ b / (a.norm() * b.norm() + 1e-300)).item()


def q(v, f):
    v = sorted(v)
    return v[min(len(v) - 1, int(f * len(v)))]


def run(dense, B, steps):
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    sd0 = ref_port.make_state(dense, seed=42)
    sd = _states.warm_state(dense, sd0, None, None, steps, fresh=4, device=torch.device("cuda"))
    x, t = (v.cuda() for v in _states.task_batch(dense, B, 7))
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
    if B <= 8:
        sdx = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
        o64, l64, g64 = ref_port.loss_and_grads(sdx, x.double(), t.double(), dense, None, 0.1)
    else:
        o64, l64, g64 = ref_port.loss_and_grads(sd, x, t, dense, None, 0.1)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        params = {k: v.detach().clone().requires_grad_(True) for k, v in ref_port.trainable(sd).items()}
        full = dict(sd)
        full.update(params)
        ob = ref_port.forward(full, x, dense, True, None)
        lb = ref_port.soft_dice_loss(ob, t) if dense == "seg" else ref_port.ssi_loss(ob, t, 0.1)
    lb.backward()
    rel = lambda a, b: ((a.double() - b.double()).abs().max() / b.double().abs().max()).item()  # noqa: E731
    print(f"== {dense} B={B} warm={steps} (fresh batches): loss ours {loss.item():.6f} exact {l64.item():.6f} autocast {lb.item():.6f}; "
          f"logits rel ours {rel(out.detach(), o64):.3e} autocast {rel(ob.float(), o64):.3e}")
    rows = []
    for n, p in m.named_parameters():
        if g64.get(n) is None:
            continue
        rows.append((l2(p.grad, g64[n]), cos(p.grad, g64[n]), l2(params[n].grad, g64[n]), cos(params[n].grad, g64[n]), g64[n].norm().item(), n))
    names = [r[5] for r in rows]
    ga = torch.cat([m.get_parameter(n).grad.flatten().double() for n in names])
    gb = torch.cat([params[n].grad.flatten().double() for n in names])
    gr = torch.cat([g64[n].flatten().double() for n in names])
    print(f"   whole gradient: ours l2 {l2(ga, gr):.3e} cos {cos(ga, gr):.6f} | autocast l2 {l2(gb, gr):.3e} cos {cos(gb, gr):.6f}")
    for thr in (2e-2, 3e-2, 5e-2):
        print(f"   l2 <= {thr:g} and cos >= 0.999: ours {sum(1 for r in rows if r[0] <= thr and r[1] >= 0.999)}/{len(rows)}, "
              f"autocast {sum(1 for r in rows if r[2] <= thr and r[3] >= 0.999)}/{len(rows)}")
    w = [r for r in rows if r[5].endswith("weight") and g64[r[5]].dim() > 1]
    print(f"   matrices only ({len(w)}): ours l2 median {q([r[0] for r in w], .5):.3e} p90 {q([r[0] for r in w], .9):.3e} max {max(r[0] for r in w):.3e}; "
          f"autocast median {q([r[2] for r in w], .5):.3e} p90 {q([r[2] for r in w], .9):.3e} max {max(r[2] for r in w):.3e}")
    print(f"   all tensors: ours l2 median {q([r[0] for r in rows], .5):.3e} p90 {q([r[0] for r in rows], .9):.3e} max {max(r[0] for r in rows):.3e}; "
          f"autocast median {q([r[2] for r in rows], .5):.3e} p90 {q([r[2] for r in rows], .9):.3e} max {max(r[2] for r in rows):.3e}")
    rows.sort(reverse=True)
    for r in rows[:6]:
        print("   l2 %.3e cos %.6f | autocast l2 %.3e cos %.6f | |g| %.3e %s" % r)


if __name__ == "__main__":
    dense = sys.argv[1] if len(sys.argv) > 1 else "seg"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    for st in [int(a) for a in sys.argv[3:]] or [30]:
        run(dense, B, st)
