"""Developer timing: fc1 GEMM with the fused bias+GELU (+pre-activation copy is internal to the model path) epilogue."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4gie_b200 import _C as C
C.init(0); L = C.lib()
M, N, K = 12608, 3072, 768
A = torch.randn(M, K, device="cuda").bfloat16(); W = (torch.randn(N, K, device="cuda") * 0.03).bfloat16()
bias = torch.randn(N, device="cuda")
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for act in (0, 1, 3):
    f = lambda: C.check(L.s4g_linear_fwd(C.ptr(A), C.ptr(W), C.ptr(bias), None, C.ptr(out), M, N, K, act, 0, C.BF16, C.TENSOR, C.stream_ptr()))
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    ref = A.float() @ W.float().t() + bias
    if act == 1: ref = torch.nn.functional.gelu(ref)
    if act == 3: ref = torch.relu(ref)
    err = ((out.float() - ref).abs().max() / ref.abs().max()).item()
    print("act", act, "%.1f us %.0f TF/s rel err %.2e" % (ms * 1e3, 2.0 * M * N * K / ms / 1e9, err))
