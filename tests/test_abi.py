"""CPU checks of the drop-in boundary: libssl4gie_b200.so loads, exports every symbol include/ssl4gie_b200.h
declares, the ctypes table covers the header, and the error plumbing works without a GPU (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ssl4gie_b200.h")


@pytest.fixture(scope="module")
def lib_path():
    from ssl4gie_b200 import _C, _build
    if not os.path.exists(_C.LIB_PATH):
        _build.build()
    return _C.LIB_PATH


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(s4g_[a-z0-9_]+)\s*\(", src)
    return sorted(set(n for n in names if n != "s4g_allreduce_fn"))


def test_header_declares_the_documented_surface():
    names = _declared_symbols()
    for must in ("s4g_version", "s4g_last_error", "s4g_init", "s4g_model_create", "s4g_model_bind", "s4g_model_forward",
                 "s4g_model_backward", "s4g_linear_fwd", "s4g_linear_wgrad", "s4g_conv3x3_fwd", "s4g_conv3x3_wgrad",
                 "s4g_layernorm_fwd", "s4g_layernorm_bwd", "s4g_attention_fwd", "s4g_attention_bwd", "s4g_bilinear2x_fwd",
                 "s4g_bilinear2x_bwd", "s4g_dice_fwd", "s4g_dice_bwd", "s4g_ssi_fwd", "s4g_ssi_bwd", "s4g_adamw_step"):
        assert must in names, must


def test_library_exports_every_declared_symbol(lib_path):
    lib = ctypes.CDLL(lib_path)
    for name in _declared_symbols():
        assert hasattr(lib, name), f"{name} is declared in include/ssl4gie_b200.h but not exported"


def test_ctypes_table_matches_header(lib_path):
    from ssl4gie_b200 import _C
    assert sorted(_C.SIGNATURES) == _declared_symbols()
    # argument counts agree with the prototypes
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, (_, args) in _C.SIGNATURES.items():
        m = re.search(r"\b" + name + r"\s*\(([^;]*?)\)\s*;", src, flags=re.S)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ("void", "") else len([p for p in params.split(",") if p.strip()])
        assert n == len(args), (name, n, len(args))


def test_version_and_error_plumbing_without_gpu(lib_path):
    from ssl4gie_b200 import _C
    lib = _C.lib()
    assert lib.s4g_version() >= 100
    # descriptor validation happens on the host: no device needed
    desc = _C.S4GModelDesc(batch=0, dense=1, dtype=1, backend=1, frozen=0, world_size=1, num_classes=1, out_token=0, dropout_p=0.1)
    h = ctypes.c_void_p()
    assert lib.s4g_model_create(ctypes.byref(desc), ctypes.byref(h)) == -1  # S4G_ERR_SHAPE
    assert b"descriptor" in lib.s4g_last_error()
    desc.batch, desc.dtype = 2, 0  # fp32 on the tensor backend is refused
    assert lib.s4g_model_create(ctypes.byref(desc), ctypes.byref(h)) == -6  # S4G_ERR_STATE
    with pytest.raises(RuntimeError, match="libssl4gie_b200 error"):
        _C.check(-6)


def test_no_cpu_fallback_in_the_product():
    """The product package never imports the oracle and refuses CPU tensors."""
    import torch

    import ssl4gie_b200
    from ssl4gie_b200 import SoftDiceLoss
    pkg = os.path.dirname(ssl4gie_b200.__file__)
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert "oracle" not in src.replace("oracle/", ""), f
    with pytest.raises(RuntimeError):
        SoftDiceLoss()(torch.zeros(2, 1, 8, 8), torch.zeros(2, 1, 8, 8))
