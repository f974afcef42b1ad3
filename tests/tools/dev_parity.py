"""Developer diagnostic: per-tensor error of our step vs an fp64 oracle run, next to the oracle's own fp32 / bf16-autocast error."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from oracle import ref_port
from tests.test_gpu_model import _inputs, _build, _loss, _rel, _l2

dense, prec, backend, B = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
m, sd = _build(dense, prec, backend)
x, t = _inputs(dense, B)
out = m(x.cuda()); loss = _loss(dense)(out, t.cuda()); loss.backward()
sdg = {k: v.cuda() for k, v in sd.items()}
o32, l32, g32 = ref_port.loss_and_grads(sdg, x.cuda(), t.cuda(), dense, None, 0.1)
sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sdg.items()}
o64, l64, g64 = ref_port.loss_and_grads(sd64, x.cuda().double(), t.cuda().double(), dense, None, 0.1)
with torch.autocast("cuda", dtype=torch.bfloat16):
    params = {k: v.detach().clone().requires_grad_(True) for k, v in ref_port.trainable(sdg).items()}
    full = dict(sdg); full.update(params)
    ob = ref_port.forward(full, x.cuda(), dense, True, None)
    lb = ref_port.soft_dice_loss(ob, t.cuda()) if dense == "seg" else ref_port.ssi_loss(ob, t.cuda(), 0.1)
lb.backward()
gb = {k: v.grad for k, v in params.items()}
print("out: mine %.3e ref32 %.3e autocast %.3e | loss mine %.8f ref64 %.8f ref32 %.8f autocast %.8f" % (_rel(out.detach(), o64), _rel(o32, o64), _rel(ob.float(), o64), loss.item(), l64.item(), l32.item(), lb.item()))
rows = []
for n, p in m.named_parameters():
    if g64.get(n) is None: continue
    a, b = p.grad.double().flatten(), g64[n].flatten()
    c = gb[n].double().flatten()
    cos = (a @ b / (a.norm() * b.norm() + 1e-300)).item()
    cosb = (c @ b / (c.norm() * b.norm() + 1e-300)).item()
    rows.append((_l2(p.grad, g64[n]), _l2(g32[n], g64[n]), _l2(gb[n], g64[n]), cos, cosb, n))
rows.sort(key=lambda r: -r[0] / max(r[1] if sys.argv[2] == 'fp32' else r[2], 1e-6))
for r in rows[:25]: print("mine %.2e ref32 %.2e autocast %.2e cos %.5f cos_autocast %.5f %s" % r)
print("median mine %.2e" % sorted(r[0] for r in rows)[len(rows)//2])
