"""Developer diagnostic: per-step deviation of the 200-step loss curve (bf16 product vs fp32 oracle), run twice."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests import test_gpu_train as T
from oracle import ref_port
sd = ref_port.make_state("seg", seed=42)
x, t = T._data("seg", 4)
ref = T._curve_oracle("seg", sd, x, t)
a = T._curve_ours("seg", sd, x, t, "bf16")
b = T._curve_ours("seg", sd, x, t, "bf16")
da = ((a - ref).abs() / ref.abs())
db = ((b - ref).abs() / ref.abs())
print("max dev run1 %.4f at %d, run2 %.4f at %d, run1-vs-run2 %.4f" % (da.max(), da.argmax(), db.max(), db.argmax(), ((a - b).abs() / ref.abs()).max()))
print("dev first 30:", [round(v, 4) for v in da[:30].tolist()])
print("ref first 12:", [round(v, 4) for v in ref[:12].tolist()])
print("ours first 12:", [round(v, 4) for v in a[:12].tolist()])
