"""Developer diagnostic for the tcgen05 attention kernels (csrc/attention_umma.cu): forward / backward against torch fp32 on
the same bf16 inputs, error split by query-row region, bit-reproducibility, timing.  Usage: python scripts/dev_attn.py [B ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ssl4gie_b200 import _C as C
C.init(0)
L = C.lib()


def rel(a, b):
    if a.numel() == 0:
        return 0.0
    return ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30)).item()


def run(B, N=197, H=12, scale_in=1.0, time_it=False):
    g = torch.Generator().manual_seed(4 + B)
    qkv = (torch.randn(B * N, 3 * H * 64, generator=g) * scale_in).cuda().bfloat16()
    do = torch.randn(B * N, H * 64, generator=g).cuda().bfloat16()
    o = torch.zeros(B * N, H * 64, device="cuda", dtype=torch.bfloat16)
    lse = torch.zeros(B, H, N, device="cuda")
    dqkv = torch.zeros_like(qkv)
    C.check(L.s4g_attention_fwd(C.ptr(qkv), C.ptr(o), C.ptr(lse), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
    torch.cuda.synchronize()
    qr = qkv.float().clone().requires_grad_(True)
    q, k, v = qr.view(B, N, 3, H, 64).permute(2, 0, 3, 1, 4).unbind(0)
    s = (q @ k.transpose(-2, -1)) * 0.125
    ref = (s.softmax(-1) @ v).transpose(1, 2).reshape(B * N, H * 64)
    lse_ref = torch.logsumexp(s, -1)
    of = o.float().view(B, N, H * 64)
    rf = ref.detach().view(B, N, H * 64)
    print(f"B={B} N={N} x{scale_in}: fwd o rel {rel(of, rf):.3e} (rows<128 {rel(of[:, :128], rf[:, :128]):.3e}, rows>=128 {rel(of[:, 128:], rf[:, 128:]):.3e}) "
          f"lse abs {(lse - lse_ref).abs().max().item():.3e} nan {int(torch.isnan(of).sum())}", flush=True)
    C.check(L.s4g_attention_bwd(C.ptr(qkv), C.ptr(o), C.ptr(do), C.ptr(lse), C.ptr(dqkv), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
    torch.cuda.synchronize()
    ref.backward(do.float())
    gd = dqkv.float().view(B, N, 3, H * 64)
    gr = qr.grad.view(B, N, 3, H * 64)
    for j, name in enumerate(("dq", "dk", "dv")):
        print(f"   {name} rel {rel(gd[:, :, j], gr[:, :, j]):.3e} (rows<128 {rel(gd[:, :128, j], gr[:, :128, j]):.3e}, rows>=128 {rel(gd[:, 128:, j], gr[:, 128:, j]):.3e}) "
              f"nan {int(torch.isnan(gd[:, :, j]).sum())}", flush=True)
    o2, d2 = torch.zeros_like(o), torch.zeros_like(dqkv)
    C.check(L.s4g_attention_fwd(C.ptr(qkv), C.ptr(o2), C.ptr(lse), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
    C.check(L.s4g_attention_bwd(C.ptr(qkv), C.ptr(o2), C.ptr(do), C.ptr(lse), C.ptr(d2), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
    torch.cuda.synchronize()
    print(f"   second run bit-identical: fwd {torch.equal(o, o2)} bwd {torch.equal(dqkv, d2)}", flush=True)
    if time_it:
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        for _ in range(3):
            C.check(L.s4g_attention_fwd(C.ptr(qkv), C.ptr(o), C.ptr(lse), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
            C.check(L.s4g_attention_bwd(C.ptr(qkv), C.ptr(o), C.ptr(do), C.ptr(lse), C.ptr(dqkv), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
        e[0].record()
        for _ in range(20):
            C.check(L.s4g_attention_fwd(C.ptr(qkv), C.ptr(o), C.ptr(lse), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
        e[1].record()
        for _ in range(20):
            C.check(L.s4g_attention_bwd(C.ptr(qkv), C.ptr(o), C.ptr(do), C.ptr(lse), C.ptr(dqkv), B, N, H, C.BF16, C.TENSOR, C.stream_ptr()))
        e[2].record()
        torch.cuda.synchronize()
        tf, tb = e[0].elapsed_time(e[1]) * 50, e[1].elapsed_time(e[2]) * 50
        fl = 4.0 * N * N * 64 * B * H
        print(f"   time fwd {tf:.1f} us ({fl / tf / 1e6:.0f} TFLOP/s)  bwd {tb:.1f} us ({2 * fl / tb / 1e6:.0f} TFLOP/s algorithmic 2x fwd)", flush=True)


if __name__ == "__main__":
    Bs = [int(a) for a in sys.argv[1:]] or [1, 3, 64]
    for B in Bs:
        run(B, time_it=(B == 64))
    run(2, scale_in=4.0)      # large logits: the row maximum matters
    run(5, N=50)              # MAE encoder sequence (49 kept patches + cls)
