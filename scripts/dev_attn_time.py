"""Timing only (S4G_ATTN_DBG experiments): attention fwd / bwd at several batch sizes (rounds per SM = ceil(12 B / 148))."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4gie_b200 import _C as C
C.init(0)
L = C.lib()
N, H = 197, 12
for B in [int(a) for a in sys.argv[1:]] or [64]:
    qkv = torch.randn(B * N, 3 * H * 64, device="cuda").bfloat16()
    o = torch.empty(B * N, H * 64, device="cuda", dtype=torch.bfloat16)
    do = torch.randn(B * N, H * 64, device="cuda").bfloat16()
    dqkv = torch.empty_like(qkv)
    lse = torch.zeros(B, H, N, device="cuda")
    def f(): C.check(L.s4g_attention_fwd(C.ptr(qkv), C.ptr(o), C.ptr(lse), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
    def b(): C.check(L.s4g_attention_bwd(C.ptr(qkv), C.ptr(o), C.ptr(do), C.ptr(lse), C.ptr(dqkv), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
    for fn in (f, b):
        for _ in range(3): fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): fn()
        e1.record(); torch.cuda.synchronize()
        print("dbg=%s B=%d rounds=%d %s %.1f us" % (os.environ.get("S4G_ATTN_DBG", "0"), B, -(-12 * B // 148), fn.__name__, e0.elapsed_time(e1) * 50), flush=True)
