"""Throughput of the GPU input pipeline (ssl4gie_b200.GpuInputPipeline) next to the reference's own per-sample transforms
(Pillow + torchvision, as Binary_segmentation/Data/dataset.py:34-64 + dataloaders.py:62-76 run them in a DataLoader worker)
on one host core.  Usage: python scripts/bench_augment.py [B] [h] [w]"""
import os
import random
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    h = int(sys.argv[2]) if len(sys.argv) > 2 else 531      # Kvasir-SEG images are ~ 500-600 px
    w = int(sys.argv[3]) if len(sys.argv) > 3 else 622
    rng = np.random.default_rng(0)
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(B)]
    masks = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(B)]
    from ssl4gie_b200 import GpuInputPipeline
    pipe = GpuInputPipeline(train=True)
    for _ in range(3):
        pipe(imgs, masks)
    torch.cuda.synchronize()
    n = 10
    t0 = time.perf_counter()
    for _ in range(n):
        x, y = pipe(imgs, masks)
    torch.cuda.synchronize()
    gpu = n * B / (time.perf_counter() - t0)
    # device time alone: the raw bytes already in HBM, the nine kernels of one batch (2 x 3 resize, jitter, blur, mask, geometric)
    params = [pipe.draw() for _ in range(B)]
    px, pm = pipe._pack(imgs, 3), pipe._pack(masks, 3)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        x8 = pipe._resize_dev(*px)
        y = pipe._mask_to_tensor(pipe._resize_dev(*pm))
        pipe._jitter(x8, params)
        pipe._geometric(pipe._blur_normalize(x8, [p.sigma for p in params]), y, params)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    raw = 2.0 * B * h * w * 3
    print(f"GPU pipeline: {gpu:.0f} img/s end to end from host arrays ({h}x{w} sources, batch {B}, host-side packing into pinned memory and "
          f"H2D of the raw bytes included); kernels alone {ms:.2f} ms per batch = {B / ms * 1e3:.0f} img/s, {raw / ms / 1e6:.0f} GB/s of source bytes")
    try:
        from PIL import Image
        import torchvision.transforms as T
        import torchvision.transforms.functional as TF
    except ImportError:
        print("Pillow / torchvision missing: no host baseline")
        return
    tin = T.Compose([T.ColorJitter(brightness=0.4, contrast=0.5, saturation=0.25, hue=0.01), T.GaussianBlur((25, 25), sigma=(0.001, 2.0)),
                     T.ToTensor(), T.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    ttg = T.Compose([T.ToTensor(), T.Grayscale()])
    torch.set_num_threads(1)
    k = min(B, 16)
    t0 = time.perf_counter()
    for i in range(k):
        x = tin(Image.fromarray(imgs[i]).resize((224, 224)))
        y = ttg(Image.fromarray(masks[i]).resize((224, 224)))
        if random.uniform(0, 1) > 0.5:
            x, y = TF.hflip(x), TF.hflip(y)
        if random.uniform(0, 1) > 0.5:
            x, y = TF.vflip(x), TF.vflip(y)
        a = (random.uniform(-180, 180), (random.uniform(-28, 28), random.uniform(-28, 28)), random.uniform(0.5, 1.5), random.uniform(-22.5, 22.5))
        x, y = TF.affine(x, *a, fill=-1.0), TF.affine(y, *a, fill=0.0)
    cpu = k / (time.perf_counter() - t0)
    print(f"reference transforms (Pillow + torchvision, decode excluded) on one host core: {cpu:.1f} img/s "
          f"=> {8 * cpu:.0f} img/s for the reference's 8 DataLoader workers (dataloaders.py:100)")


if __name__ == "__main__":
    main()
