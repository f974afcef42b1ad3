"""Where does the end-to-end (host buffers) step lose time against the device-resident step?  1 GPU."""
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
loss_h = torch.empty((), dtype=torch.float32).pin_memory()


def step(x, t, readback):
    loss = loss_fn(model(x), t)
    loss.backward()
    opt.step()
    if readback:
        loss_h.copy_(loss.detach(), non_blocking=True)


def timed(name, n, it, readback):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    cpu, evs = [], [e0]
    for x, t in it:
        c0 = time.perf_counter()
        step(x, t, readback)
        cpu.append(time.perf_counter() - c0)
        ev = torch.cuda.Event(enable_timing=True); ev.record(); evs.append(ev)
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    per = [evs[i].elapsed_time(evs[i + 1]) for i in range(len(evs) - 1)]
    if max(per) > 30:
        print("   per-step ms:", " ".join(f"{p:.1f}" for p in per), flush=True)
    print(f"{name:34s} gpu {e0.elapsed_time(e1)/n:7.3f} ms/step   host enqueue {1e3*(t1-t0)/n:7.3f} ms/step   max host step {1e3*max(cpu):7.3f} ms", flush=True)


def manual_copy(n):
    for _ in range(n):
        yield x_h.cuda(non_blocking=True), t_h.cuda(non_blocking=True)


for _ in range(5):
    step(x_d, t_d, False)
n = 30
for rep in range(4):
    timed("resident", n, ((x_d, t_d) for _ in range(n)), False)
    timed("resident + loss readback", n, ((x_d, t_d) for _ in range(n)), True)
    timed("same-stream H2D copies", n, manual_copy(n), False)
    timed("prefetcher depth 2", n, DevicePrefetcher((x_h, t_h) for _ in range(n)), False)
    timed("prefetcher depth 2 + readback", n, DevicePrefetcher((x_h, t_h) for _ in range(n)), True)
    timed("prefetcher depth 1 + readback", n, DevicePrefetcher(((x_h, t_h) for _ in range(n)), depth=1), True)
print("mem reserved GB", torch.cuda.memory_reserved() / 1e9)
