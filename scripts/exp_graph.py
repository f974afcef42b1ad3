"""Experiment: whole training step captured in one CUDA graph vs eager launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4gie_b200 import FusedAdamW, SoftDiceLoss, utils
from ssl4gie_b200.ddp import GradientAllReduce
from bench import synth
B = 64
torch.manual_seed(42)
model = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
model.precision, model.backend = "bf16", "tensor"
model = model.cuda().train()
loss_fn = SoftDiceLoss()
opt = FusedAdamW(model, lr=1e-4)
ddp = GradientAllReduce(model); ddp.prepare(B)
x, t = synth("seg", B, 0); x, t = x.cuda(), t.cuda()
def step():
    out = model(x); loss = loss_fn(out, t); loss.backward(); opt.step(); return loss
def timeit(f, n=10):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): step()
torch.cuda.current_stream().wait_stream(s)
print("eager ms/step", timeit(step))
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    static_loss = step()
print("graph ms/step", timeit(g.replay), "loss", static_loss.item())
