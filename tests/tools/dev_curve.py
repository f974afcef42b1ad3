"""Developer diagnostic: per-step deviation of the 200-step loss curve (bf16 product vs fp32 oracle).
Usage: python tests/tools/dev_curve.py [seg|depth|cls ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tests import test_gpu_train as T
from oracle import ref_port
for task in (sys.argv[1:] or ["seg"]):
    sd = T._start_state(task)
    data = T._data(task, 4)
    ref = T._curve_oracle(task, sd, data)
    a = T._curve_ours(task, sd, data, "bf16")
    b = T._curve_ours(task, sd, data, "bf16")
    scale = torch.maximum(ref.abs(), 0.05 * ref[0].abs()) if task == "cls" else ref.abs()
    da = (a - ref).abs() / scale
    print("%s: max dev %.4f at %d (first 100: %.4f), run1 == run2: %s" % (task, da.max(), da.argmax(), da[:100].max(), torch.equal(a, b)))
    print("  dev every 10:", [round(v, 4) for v in da[::10].tolist()])
    print("  ref every 20:", [round(v, 5) for v in ref[::20].tolist()])
    print("  ours every 20:", [round(v, 5) for v in a[::20].tolist()])
