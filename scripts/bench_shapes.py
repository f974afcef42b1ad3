"""Developer benchmark: every contraction shape of one ViT-B/16 + DPT seg training step (B images), timed alone
through the C ABI with CUDA events; prints ms, TFLOP/s and the step-weighted total.  Usage: bench_shapes.py [B] [filter]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ssl4gie_b200 import _C as C

_script = __name__ == "__main__"
B = int(sys.argv[1]) if _script and len(sys.argv) > 1 else int(os.environ.get("S4G_SHAPES_B", 64))
flt = sys.argv[2] if _script and len(sys.argv) > 2 else ""
C.init(0)
L = C.lib()
T = B * 197
P = B * 196
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n


def lin_fwd(M, N, K):
    A = torch.randn(M, K, device="cuda").bfloat16()
    W = torch.randn(N, K, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    return lambda: C.check(L.s4g_linear_fwd(C.ptr(A), C.ptr(W), None, None, C.ptr(out), M, N, K, 0, 0, C.BF16, C.TENSOR, C.stream_ptr()))


def lin_wgrad(M, N, K):
    dY = torch.randn(M, N, device="cuda").bfloat16()
    X = torch.randn(M, K, device="cuda").bfloat16()
    dW = torch.empty(N, K, device="cuda")
    nb = L.s4g_wgrad_scratch_bytes(M, 1, 1, K, N, 1)
    sc = torch.empty(nb, dtype=torch.uint8, device="cuda")
    return lambda: C.check(L.s4g_linear_wgrad(C.ptr(dY), C.ptr(X), C.ptr(dW), M, N, K, C.BF16, C.TENSOR, C.ptr(sc), nb, C.stream_ptr()))


def conv_fwd(H, Ci, Co):
    x = torch.randn(B, H, H, Ci, device="cuda").bfloat16()
    w = torch.randn(Co, 3, 3, Ci, device="cuda").bfloat16()
    out = torch.empty(B, H, H, Co, device="cuda", dtype=torch.bfloat16)
    return lambda: C.check(L.s4g_conv3x3_fwd(C.ptr(x), C.ptr(w), None, C.ptr(out), B, H, H, Ci, Co, C.BF16, C.TENSOR, C.stream_ptr()))


def conv_wgrad(H, Ci, Co):
    x = torch.randn(B, H, H, Ci, device="cuda").bfloat16()
    dy = torch.randn(B, H, H, Co, device="cuda").bfloat16()
    dw = torch.empty(Co, 3, 3, Ci, device="cuda")
    nb = L.s4g_wgrad_scratch_bytes(B, H, H, Ci, Co, 9)
    sc = torch.empty(nb, dtype=torch.uint8, device="cuda")
    return lambda: C.check(L.s4g_conv3x3_wgrad(C.ptr(dy), C.ptr(x), C.ptr(dw), B, H, H, Ci, Co, C.BF16, C.TENSOR, C.ptr(sc), nb, C.stream_ptr()))


rows = []  # (name, count per step, flops, fn factory)



def lin(name, M, N, K, cnt, dgrad=True):
    f = 2.0 * M * N * K
    rows.append((f"{name} fwd   M{M} N{N} K{K}", cnt, f, lambda: lin_fwd(M, N, K)))
    if dgrad:
        rows.append((f"{name} dgrad M{M} N{K} K{N}", cnt, f, lambda: lin_fwd(M, K, N)))
    rows.append((f"{name} wgrad M{M} N{N} K{K}", cnt, f, lambda: lin_wgrad(M, N, K)))


def conv(name, H, Ci, Co, cnt):
    f = 2.0 * B * H * H * Ci * Co * 9
    rows.append((f"{name} fwd   H{H} {Ci}->{Co}", cnt, f, lambda: conv_fwd(H, Ci, Co)))
    rows.append((f"{name} dgrad H{H} {Co}->{Ci}", cnt, f, lambda: conv_fwd(H, Co, Ci)))
    rows.append((f"{name} wgrad H{H} {Ci}->{Co}", cnt, f, lambda: conv_wgrad(H, Ci, Co)))


lin("patch", P, 768, 768, 1, dgrad=False)
lin("qkv", T, 2304, 768, 12)
lin("proj", T, 768, 768, 12)
lin("fc1", T, 3072, 768, 12)
lin("fc2", T, 768, 3072, 12)
for n in (96, 192, 384, 768):
    lin(f"post1x1_{n}", P, n, 768, 1)
lin("convT4", P, 1536, 96, 1)
lin("convT2", P, 768, 192, 1)
lin("conv_s2", B * 49, 768, 6912, 1)
conv("layer1_rn", 56, 96, 256, 1)
conv("layer2_rn", 28, 192, 256, 1)
conv("layer3_rn", 14, 384, 256, 1)
conv("layer4_rn", 7, 768, 256, 1)
conv("rcu7", 7, 256, 256, 2)
conv("rcu14", 14, 256, 256, 4)
conv("rcu28", 28, 256, 256, 4)
conv("rcu56", 56, 256, 256, 4)
conv("head112", 112, 256, 256, 1)
for H in (14, 28, 56, 112):
    lin(f"out_conv{H}", B * H * H, 256, 256, 1)



def attention(n=5):
    """(fwd ms, bwd ms, algorithmic fwd FLOP) of one attention layer (timm Attention: 4 N^2 d per head forward, 2x backward)."""
    N, H = 197, 12
    qkv = torch.randn(B * N, 3 * H * 64, device="cuda").bfloat16()
    o = torch.empty(B * N, H * 64, device="cuda", dtype=torch.bfloat16)
    do = torch.randn(B * N, H * 64, device="cuda").bfloat16()
    dqkv = torch.empty_like(qkv)
    lse = torch.empty(B, H, N, device="cuda")
    f = timeit(lambda: C.check(L.s4g_attention_fwd(C.ptr(qkv), C.ptr(o), C.ptr(lse), B, N, H, C.BF16, C.TENSOR, C.stream_ptr())), n)
    b = timeit(lambda: C.check(L.s4g_attention_bwd(C.ptr(qkv), C.ptr(o), C.ptr(do), C.ptr(lse), C.ptr(dqkv), B, N, H, C.BF16, C.TENSOR, C.stream_ptr())), n)
    return f, b, 4.0 * N * N * 64 * B * H


def run(verbose=True, n=5):
    """Times every shape; returns {family: (ms per step, FLOP per step)} for the families fprop_dgrad / wgrad / attention."""
    fam = {"fprop_dgrad": [0.0, 0.0], "wgrad": [0.0, 0.0], "attention": [0.0, 0.0]}
    tot_ms = tot_f = 0.0
    for name, cnt, f, mk in rows:
        if flt and flt not in name:
            continue
        fn = mk()
        ms = timeit(fn, n)
        del fn
        torch.cuda.empty_cache()
        tot_ms += ms * cnt
        tot_f += f * cnt
        k = "wgrad" if " wgrad " in name else "fprop_dgrad"
        fam[k][0] += ms * cnt
        fam[k][1] += f * cnt
        if verbose:
            print(f"{name:42s} x{cnt:2d} {ms:8.3f} ms {f / ms / 1e9:8.1f} TF/s  step {ms * cnt:7.3f} ms", flush=True)
    if verbose:
        print(f"TOTAL {tot_ms:.3f} ms/step  {tot_f / 1e12:.2f} TFLOP  {tot_f / tot_ms / 1e9:.1f} TF/s")
    if not flt:
        af, ab, aflop = attention(n)
        fam["attention"] = [12 * (af + ab), 12 * 3 * aflop]
        if verbose:
            print(f"attention fwd {af * 1e3:.1f} us ({aflop / af / 1e9:.0f} TF/s)  bwd {ab * 1e3:.1f} us ({2 * aflop / ab / 1e9:.0f} TF/s)  x12 layers: "
                  f"{12 * (af + ab):.3f} ms/step")
    return fam


if __name__ == "__main__":
    run()
