"""Developer diagnostic for the tcgen05 engine: each case prints its error; a hang is cut by the caller's timeout."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4gie_b200 import _C as C
C.init(0)
def rel(a, b): return ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30)).item()
which = sys.argv[1] if len(sys.argv) > 1 else "all"
g = torch.Generator().manual_seed(0)
if which in ("all", "lin"):
    for (M, N, K) in [(128, 256, 64), (128, 256, 768), (394, 2304, 768), (12608, 768, 3072), (1000, 96, 768), (392, 1536, 96), (300, 32, 128)]:
        A = torch.randn(M, K, generator=g).cuda().bfloat16(); W = (torch.randn(N, K, generator=g) * 0.05).cuda().bfloat16()
        bias = torch.randn(N, generator=g).cuda(); out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
        C.check(C.lib().s4g_linear_fwd(C.ptr(A), C.ptr(W), C.ptr(bias), None, C.ptr(out), M, N, K, 0, 0, C.BF16, C.TENSOR, C.stream_ptr()))
        torch.cuda.synchronize()
        ref = A.float() @ W.float().t() + bias
        print("linear_fwd", M, N, K, "rel", rel(out.float(), ref), flush=True)
if which in ("all", "wgrad"):
    for (M, N, K) in [(64, 128, 64), (128, 128, 64), (394, 2304, 768), (12608, 768, 768), (6272, 1536, 96), (500, 32, 128)]:
        dY = torch.randn(M, N, generator=g).cuda().bfloat16(); X = torch.randn(M, K, generator=g).cuda().bfloat16()
        dW = torch.empty(N, K, device="cuda")
        nb = C.lib().s4g_wgrad_scratch_bytes(M, 1, 1, K, N, 1); sc = torch.empty(nb, dtype=torch.uint8, device="cuda")
        C.check(C.lib().s4g_linear_wgrad(C.ptr(dY), C.ptr(X), C.ptr(dW), M, N, K, C.BF16, C.TENSOR, C.ptr(sc), nb, C.stream_ptr()))
        torch.cuda.synchronize()
        ref = dY.float().t() @ X.float()
        print("linear_wgrad", M, N, K, "rel", rel(dW, ref), flush=True)
if which in ("all", "conv"):
    for (B, H, Cin, Cout) in [(2, 8, 64, 64), (2, 7, 768, 256), (2, 14, 256, 256), (3, 28, 192, 256), (2, 56, 96, 256), (2, 112, 256, 128), (1, 224, 128, 32)]:
        x = torch.randn(B, Cin, H, H, generator=g).cuda().bfloat16().float(); w = (torch.randn(Cout, Cin, 3, 3, generator=g) * 0.05).cuda().bfloat16().float()
        bias = torch.randn(Cout, generator=g).cuda()
        ref = torch.nn.functional.conv2d(x, w, bias, padding=1)
        xn = x.permute(0, 2, 3, 1).contiguous().bfloat16(); wp = w.permute(0, 2, 3, 1).contiguous().bfloat16()
        out = torch.zeros(B, H, H, Cout, device="cuda", dtype=torch.bfloat16)
        C.check(C.lib().s4g_conv3x3_fwd(C.ptr(xn), C.ptr(wp), C.ptr(bias), C.ptr(out), B, H, H, Cin, Cout, C.BF16, C.TENSOR, C.stream_ptr()))
        torch.cuda.synchronize()
        print("conv_fwd", B, H, Cin, Cout, "rel", rel(out.float().permute(0, 3, 1, 2), ref), flush=True)
        dy = torch.randn(B, Cout, H, H, generator=g).cuda().bfloat16().float()
        dw_ref = torch.nn.grad.conv2d_weight(x, w.shape, dy, padding=1)
        dyn = dy.permute(0, 2, 3, 1).contiguous().bfloat16()
        dwp = torch.empty(Cout, 3, 3, Cin, device="cuda")
        nb = C.lib().s4g_wgrad_scratch_bytes(B, H, H, Cin, Cout, 9); sc = torch.empty(nb, dtype=torch.uint8, device="cuda")
        C.check(C.lib().s4g_conv3x3_wgrad(C.ptr(dyn), C.ptr(xn), C.ptr(dwp), B, H, H, Cin, Cout, C.BF16, C.TENSOR, C.ptr(sc), nb, C.stream_ptr()))
        torch.cuda.synchronize()
        print("conv_wgrad", B, H, Cin, Cout, "rel", rel(dwp.permute(0, 3, 1, 2), dw_ref), flush=True)
