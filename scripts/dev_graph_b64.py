"""Developer diagnostic: which ingredient of bench.py's process breaks the CUDA-graph capture of the step at B = 64."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import synth  # noqa: E402
from ssl4gie_b200 import FusedAdamW, GraphedStep, SoftDiceLoss, utils  # noqa: E402
from ssl4gie_b200.ddp import GradientAllReduce  # noqa: E402

variant = sys.argv[1]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
torch.manual_seed(42)
model = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
model.precision, model.backend = "bf16", "tensor"
model = model.cuda().train()
loss_fn = SoftDiceLoss()
opt = FusedAdamW(model, lr=1e-4)
x_h, t_h = synth("seg", B, 0)
x_h, t_h = x_h.pin_memory(), t_h.pin_memory()
x_d, t_d = x_h.cuda(non_blocking=True), t_h.cuda(non_blocking=True)
if "ddp" in variant:
    ddp = GradientAllReduce(model)
    ddp.attach(opt)
    ddp.prepare(B)
if "prefetch" in variant:
    from ssl4gie_b200.prefetch import DevicePrefetcher
    for x, t in DevicePrefetcher((x_h, t_h) for _ in range(4)):
        loss_fn(model(x), t).backward()
        opt.step()
    del x, t
for _ in range(3):
    loss_fn(model(x_d), t_d).backward()
    opt.step()
torch.cuda.synchronize()
try:
    gs = GraphedStep(model, loss_fn, opt, eager_steps=1)
    out = [gs(x_d, t_d).item() for _ in range(4)]
    print(variant, B, "OK", out)
except Exception as e:  # noqa: BLE001
    print(variant, B, "FAILED", repr(e)[:300])
