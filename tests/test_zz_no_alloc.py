"""s4g_layernorm_bwd_static: the allocation-free variant of s4g_layernorm_bwd (scratch in a static __device__ array,
csrc/no_alloc.cu; no counterpart in the reference).  Collected last on purpose:
a failure here must not hide the parity tests.

Checked through the C ABI against the existing entry point on the same inputs, two calls in a row with different shapes (the
second call finds the first call's partial rows in the shared array), and with the library's cudaMalloc'ed scratch never set
up by this test: the new entry point must work without s4g_init."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30)).item()


def _case(lib, C, rows, Cdim, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    code, tdt = (C.F32, torch.float32) if dtype == "fp32" else (C.BF16, torch.bfloat16)
    x = (torch.randn(rows, Cdim, generator=g) * 2 + 0.3).cuda()
    gamma = (1 + 0.1 * torch.randn(Cdim, generator=g)).cuda()
    beta = (0.1 * torch.randn(Cdim, generator=g)).cuda()
    y = torch.empty(rows, Cdim, device="cuda", dtype=tdt)
    mean = torch.empty(rows, device="cuda")
    rstd = torch.empty(rows, device="cuda")
    C.check(lib.s4g_layernorm_fwd(C.ptr(x), C.ptr(gamma), C.ptr(beta), C.ptr(y), C.ptr(mean), C.ptr(rstd), rows, Cdim, 1e-6, code, C.stream_ptr()))
    dy = torch.randn(rows, Cdim, generator=g).cuda().to(tdt)
    dres = torch.randn(rows, Cdim, generator=g).cuda()
    outs = []
    for fn in (lib.s4g_layernorm_bwd_static, lib.s4g_layernorm_bwd):
        dx = torch.full_like(x, float("nan"))
        dg = torch.full((Cdim,), float("nan"), device="cuda")
        db = torch.full((Cdim,), float("nan"), device="cuda")
        if fn is lib.s4g_layernorm_bwd:
            C.init(0)   # the existing entry point needs the library scratch that s4g_init allocates; the static one ran before it
        C.check(fn(C.ptr(dy), C.ptr(x), C.ptr(gamma), C.ptr(mean), C.ptr(rstd), C.ptr(dres), C.ptr(dx), C.ptr(dg), C.ptr(db), rows, Cdim,
                   code, C.stream_ptr()))
        outs.append((dx, dg, db))
    torch.cuda.synchronize()
    # reference for good measure: torch autograd
    xr, gr, br = x.clone().requires_grad_(True), gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    torch.nn.functional.layer_norm(xr, (Cdim,), gr, br, 1e-6).backward(dy.float())
    return outs, (xr.grad + dres, gr.grad, br.grad)


def test_static_scratch_layernorm_bwd_matches_the_allocating_entry_point():
    from ssl4gie_b200 import _C as C
    lib = C.lib()
    torch.zeros(1, device="cuda")   # a CUDA context; deliberately NOT _C.init before the first static call
    # two calls in a row, different shapes and dtypes: 148 blocks then a short grid that leaves most of the first call's partial
    # rows in the shared array, and the narrower C = 512 layout on top of the C = 768 one
    for rows, Cdim, dtype, seed in ((64 * 197, 768, "bf16", 1), (37, 512, "fp32", 2), (3 * 197, 768, "fp32", 3)):
        (got, want), ref = _case(lib, C, rows, Cdim, dtype, seed)
        for a, b, r, name in zip(got, want, ref, ("dx", "dgamma", "dbeta")):
            assert torch.equal(a, b), (rows, Cdim, dtype, name)          # same kernels, same fixed-order reduction: bit-identical
            assert _rel(a, r) < 1e-4, (rows, Cdim, dtype, name)           # and within the tolerance tests/test_gpu_ops.py uses


def test_static_scratch_rejects_sizes_beyond_the_array():
    from ssl4gie_b200 import _C as C
    lib = C.lib()
    t = torch.zeros(4, 1024, device="cuda")
    v = torch.zeros(1024, device="cuda")
    rc = lib.s4g_layernorm_bwd_static(C.ptr(t), C.ptr(t), C.ptr(v), C.ptr(v), C.ptr(v), None, C.ptr(t), C.ptr(v), C.ptr(v), 4, 1024, C.F32,
                                      C.stream_ptr())
    assert rc < 0
    msg = lib.s4g_last_error().decode()
    assert "static array" in msg and "768" in msg, msg
    with pytest.raises(RuntimeError):
        C.check(rc)
