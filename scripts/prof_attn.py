"""Developer profiling target: attention fwd/bwd (B=64) through the C ABI (run under ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4gie_b200 import _C as C
C.init(0)
L = C.lib()
B, N, H = 64, 197, 12
qkv = torch.randn(B * N, 3 * H * 64, device="cuda").bfloat16()
o = torch.empty(B * N, H * 64, device="cuda", dtype=torch.bfloat16)
do = torch.randn(B * N, H * 64, device="cuda").bfloat16()
dqkv = torch.empty_like(qkv)
lse = torch.empty(B, H, N, device="cuda")
for _ in range(3):
    C.check(L.s4g_attention_fwd(C.ptr(qkv), C.ptr(o), C.ptr(lse), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
    C.check(L.s4g_attention_bwd(C.ptr(qkv), C.ptr(o), C.ptr(do), C.ptr(lse), C.ptr(dqkv), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
torch.cuda.synchronize()
print("ok")
