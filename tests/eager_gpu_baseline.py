"""the comparison number "the reference's modules under torch eager + bf16 autocast on the same B200".
Not a test (not collected) and not on any product path: runs the oracle restatement (oracle/ref_port.py, the reference's
own op sequence -- cuBLASLt / cuDNN / ATen kernels) the way the reference trains (train_segmentation.py:41-52: autocast
forward + loss, backward, torch.optim.AdamW step), batch 64, synthetic inputs, and prints img/s.

    python tests/eager_gpu_baseline.py [seg|depth] [batch] [steps]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_port  # noqa: E402


def main():
    dense = sys.argv[1] if len(sys.argv) > 1 else "seg"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    sd = {k: v.cuda() for k, v in ref_port.make_state(dense, seed=42).items()}
    params = {k: v.clone().requires_grad_(True) for k, v in ref_port.trainable(sd).items()}
    full = dict(sd)
    full.update(params)
    opt = torch.optim.AdamW(list(params.values()), lr=1e-4)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(B, 3, 224, 224, generator=g).cuda()
    t = ((torch.rand(B, 1, 224, 224, generator=g) > 0.5).float() if dense == "seg" else torch.rand(B, 1, 224, 224, generator=g)).cuda()

    def step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = ref_port.forward(full, x, dense, True, None)
            loss = ref_port.soft_dice_loss(out, t) if dense == "seg" else ref_port.ssi_loss(out, t, 0.1)
        loss.backward()
        opt.step()
        return loss

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(f"torch {torch.__version__} eager + bf16 autocast, {dense}, B={B}: {ms:.2f} ms/step = {B / ms * 1e3:.1f} img/s "
          f"(loss {loss.item():.4f}, peak memory {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB)")


if __name__ == "__main__":
    main()
