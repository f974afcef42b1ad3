"""First e2e step after a synchronize: where do the extra ~12 ms go?  Host timestamps + GPU events."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4gie_b200 import FusedAdamW, SoftDiceLoss, utils
from ssl4gie_b200.prefetch import DevicePrefetcher

B = 64
torch.manual_seed(42)
model = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
model.precision, model.backend = "bf16", "tensor"
model = model.cuda().train()
loss_fn = SoftDiceLoss()
opt = FusedAdamW(model, lr=1e-4)
x_h = torch.randn(B, 3, 224, 224).pin_memory()
t_h = (torch.rand(B, 1, 224, 224) > 0.5).float().pin_memory()
x_d, t_d = x_h.cuda(), t_h.cuda()


def step(x, t):
    loss = loss_fn(model(x), t)
    loss.backward()
    opt.step()


for _ in range(5):
    step(x_d, t_d)
for x, t in DevicePrefetcher((x_h, t_h) for _ in range(4)):
    step(x, t)

for rep in range(3):
    for mode in ("resident", "prefetch", "copy-only-first"):
        torch.cuda.synchronize()
        ev = lambda: torch.cuda.Event(enable_timing=True)
        e0, e_in, e_fwd, e_end = ev(), ev(), ev(), ev()
        h0 = time.perf_counter()
        e0.record()
        if mode == "resident":
            it = iter([(x_d, t_d)] * 3)
        else:
            it = iter(DevicePrefetcher((x_h, t_h) for _ in range(3)))
        x, t = next(it)
        h1 = time.perf_counter()
        e_in.record()
        out = model(x)
        h2 = time.perf_counter()
        e_fwd.record()
        loss = loss_fn(out, t)
        loss.backward()
        opt.step()
        h3 = time.perf_counter()
        e_end.record()
        for x, t in it:
            step(x, t)
        torch.cuda.synchronize()
        print(f"{mode:16s} host: next {1e3*(h1-h0):6.2f} fwd-enq {1e3*(h2-h1):6.2f} bwd-enq {1e3*(h3-h2):6.2f} | "
              f"gpu: input ready {e0.elapsed_time(e_in):6.2f} fwd {e_in.elapsed_time(e_fwd):6.2f} bwd+opt {e_fwd.elapsed_time(e_end):6.2f}", flush=True)
