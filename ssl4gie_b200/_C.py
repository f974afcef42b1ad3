"""ctypes binding of libssl4gie_b200.so (include/ssl4gie_b200.h).  Loading never falls back to anything."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libssl4gie_b200.so")

F32, BF16 = 0, 1
DENSE = {None: 0, False: 0, "seg": 1, "depth": 2, "mae": 3}
SIMT, TENSOR = 0, 1


class S4GModelDesc(ctypes.Structure):
    _fields_ = [("batch", ctypes.c_int), ("dense", ctypes.c_int), ("dtype", ctypes.c_int), ("backend", ctypes.c_int),
                ("frozen", ctypes.c_int), ("world_size", ctypes.c_int), ("num_classes", ctypes.c_int),
                ("out_token", ctypes.c_int), ("dropout_p", ctypes.c_float)]


ALLREDUCE_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p)

_vp, _i, _ll, _f, _sz, _u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_size_t, ctypes.c_uint64

# name -> (restype, argtypes): one row per symbol declared in include/ssl4gie_b200.h
SIGNATURES = {
    "s4g_version": (_i, []),
    "s4g_last_error": (ctypes.c_char_p, []),
    "s4g_init": (_i, [_i]),
    "s4g_model_create": (_i, [ctypes.POINTER(S4GModelDesc), ctypes.POINTER(_vp)]),
    "s4g_model_destroy": (None, [_vp]),
    "s4g_model_bind": (_i, [_vp, ctypes.c_char_p, _vp, _vp, _ll]),
    "s4g_model_set_grad_arena": (_i, [_vp, _vp, _sz]),
    "s4g_model_workspace_bytes": (_sz, [_vp]),
    "s4g_model_set_workspace": (_i, [_vp, _vp, _sz]),
    "s4g_model_forward": (_i, [_vp, _vp, _vp, _i, _u64, _vp, _vp]),
    "s4g_model_backward": (_i, [_vp, _vp, _vp]),
    "s4g_model_set_mae_io": (_i, [_vp, _vp, _vp, _vp]),
    "s4g_model_set_seed_device": (_i, [_vp, _vp]),
    "s4g_model_set_allreduce": (_i, [_vp, ALLREDUCE_FN, _vp]),
    "s4g_peer_mailbox_bytes": (_sz, []),
    "s4g_peer_alloc": (_i, [_sz, ctypes.POINTER(_vp), ctypes.c_char_p]),
    "s4g_peer_open": (_i, [ctypes.c_char_p, ctypes.POINTER(_vp)]),
    "s4g_peer_close": (_i, [_vp]),
    "s4g_peer_free": (_i, [_vp]),
    "s4g_model_set_peer_exchange": (_i, [_vp, ctypes.POINTER(_vp), _i, _i]),
    "s4g_peer_sum_f64": (_i, [_vp, _i, ctypes.POINTER(_vp), _i, _i, _vp]),
    "s4g_peer_check": (_i, [_vp]),
    "s4g_model_set_bucket_events": (_i, [_vp, ctypes.POINTER(_vp), _i]),
    "s4g_model_bucket_params": (ctypes.c_char_p, [_vp, _i, _i]),
    "s4g_model_last_launches": (_i, [_vp]),
    "s4g_linear_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _ll, _i, _i, _i, _i, _i, _i, _vp]),
    "s4g_linear_wgrad": (_i, [_vp, _vp, _vp, _ll, _i, _i, _i, _i, _vp, _sz, _vp]),
    "s4g_conv3x3_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "s4g_conv3x3_wgrad": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "s4g_wgrad_scratch_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "s4g_layernorm_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _f, _i, _vp]),
    "s4g_layernorm_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "s4g_layernorm_bwd_static": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "s4g_attention_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "s4g_attention_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "s4g_bilinear2x_fwd": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "s4g_bilinear2x_bwd": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "s4g_dice_scratch_bytes": (_sz, [_i]),
    "s4g_dice_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "s4g_dice_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "s4g_ssi_scratch_bytes": (_sz, [_i]),
    "s4g_ssi_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _i, _i, _vp]),
    "s4g_ssi_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _i, _vp]),
    "s4g_ce_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "s4g_ce_bwd": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "s4g_adamw_step": (_i, [_vp, _vp, _vp, _vp, _ll, _f, _f, _f, _f, _f, _i, _f, _vp]),
    "s4g_aug_resize_scratch_bytes": (_sz, [_vp, _i, _i]),
    "s4g_aug_resize": (_i, [_vp, _vp, _vp, _i, _i, _vp, _sz, _vp, _vp]),
    "s4g_aug_color_jitter": (_i, [_vp, _vp, _i, _vp]),
    "s4g_aug_blur_normalize": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "s4g_aug_mask_to_tensor": (_i, [_vp, _vp, _i, _i, _vp]),
    "s4g_aug_depth_target": (_i, [_vp, _vp, _vp, _i, _vp]),
    "s4g_aug_geometric": (_i, [_vp, _vp, _vp, _vp, _vp, _f, _f, _i, _vp]),
    "s4g_adamw_step_device": (_i, [_vp, _vp, _vp, _vp, _ll, _vp, _vp, _vp, _vp]),
    "s4g_device_hyper_set": (_i, [_vp, _f, _f, _f, _f, _f, _f, _vp]),
    "s4g_device_counter_advance": (_i, [_vp, _vp, _vp]),
    "s4g_adamw_step_groups": (_i, [_vp, _vp, _vp, _vp, _ll, _f, _f, _f, _f, _vp, _i, _f, _vp]),
    "s4g_mae_loss_scratch_bytes": (_sz, [_i]),
    "s4g_mae_loss_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "s4g_mae_loss_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
}

_lib = None
_inited = set()


def lib():
    """The loaded library; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(ssl4gie_b200 has no CPU / PyTorch fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype = res
            fn.argtypes = args
    return _lib


def check(code):
    if code != 0:
        raise RuntimeError(f"libssl4gie_b200 error {code}: {lib().s4g_last_error().decode()}")


def init(device_index):
    if device_index not in _inited:
        check(lib().s4g_init(int(device_index)))
        _inited.add(device_index)


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"ssl4gie_b200: {what} must be a CUDA tensor — this package has no CPU path")
    init(t.device.index if t.device.index is not None else 0)
