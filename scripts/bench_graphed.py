"""The training step captured in a CUDA graph (ssl4gie_b200.GraphedStep): GPU time per step and host time per replay next to the
eager step of the same process.  Prints one JSON object.  Usage: python scripts/bench_graphed.py [seg|depth|cls|mae] [B]"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ssl4gie_b200  # noqa: E402
from bench import synth  # noqa: E402
from ssl4gie_b200 import CrossEntropyLoss, FusedAdamW, GraphedStep, ScaleAndShiftInvariantLoss, SoftDiceLoss, utils  # noqa: E402


def main():
    task = sys.argv[1] if len(sys.argv) > 1 else "seg"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else (128 if task == "cls" else 64)
    torch.manual_seed(42)
    if task == "cls":
        model, loss_fn = utils.get_ImageNet_or_random_ViT(True, 23, False, None, False, False), CrossEntropyLoss(torch.ones(23)).cuda()
    elif task == "mae":
        model, loss_fn = ssl4gie_b200.mae_vit_base_patch16(norm_pix_loss=True), None
    else:
        model = utils.get_ImageNet_or_random_ViT(False, 1 if task == "seg" else None, False, task, False, False)
        loss_fn = SoftDiceLoss() if task == "seg" else ScaleAndShiftInvariantLoss(alpha=0.1)
    model.precision, model.backend = "bf16", "tensor"
    model = model.cuda().train()
    if task == "mae":
        opt = FusedAdamW(model, lr=1e-4, betas=(0.9, 0.95), param_groups=ssl4gie_b200.add_weight_decay(model, 0.05))
    else:
        opt = FusedAdamW(model, lr=1e-4)
    x, t = (v.cuda() for v in synth(task, B, 0))

    def eager():
        loss = model(x, noise=t)[0] if task == "mae" else loss_fn(model(x), t)
        loss.backward()
        opt.step()

    def timed(fn, n=20):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        host = []
        e0.record()
        for _ in range(n):
            t0 = time.perf_counter()
            fn()
            host.append(time.perf_counter() - t0)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n, 1e3 * sorted(host)[n // 2]

    for _ in range(4):
        eager()
    # host time of an eager step on an idle device: one step at a time, synchronised (the launch queue never fills)
    idle = []
    for _ in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eager()
        idle.append(time.perf_counter() - t0)
    ms_eager, _ = timed(eager)
    gs = GraphedStep(model, loss_fn, opt, eager_steps=0)
    for _ in range(3):
        gs(x, t)
    ms_graph, host_graph = timed(lambda: gs(x, t))
    print(json.dumps({"task": task, "images_per_gpu": B, "eager_ms_per_step": ms_eager, "eager_host_enqueue_ms": 1e3 * sorted(idle)[2],
                      "graph_ms_per_step": ms_graph, "graph_host_ms_per_replay": host_graph, "steps": 20,
                      "note": "fwd + loss + bwd + AdamW replayed from one CUDA graph; lr / step count / dropout seed live in device memory"}))


if __name__ == "__main__":
    main()
