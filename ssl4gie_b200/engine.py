"""Host side of the static schedule: flat fp32 parameter/gradient arenas, per-shape C model handles, the single
autograd.Function that wraps model forward/backward, SyncBatchNorm exchange callback."""
import ctypes
import itertools
import os
import weakref

import torch
import torch.nn as nn

from . import _C

_ALIGN = 64  # floats (256 B): every parameter starts on a 256 B boundary inside the arena


def _unused_in_dense(name):
    # never receive a gradient in dense mode (reference: final norm skipped, Models/models.py:338,456; refinenet4 is
    # called with one input so resConfUnit1 never runs, Models/DPT_decoder.py:532,288)
    return name.startswith("norm.") or name.startswith("decoder.refinenet4.resConfUnit1.")


def arena_layout(model):
    """Pure host logic (no CUDA): placement of every parameter in the flat fp32 arena.  Parameters that receive a
    gradient come first (their offsets are also the offsets in the flat gradient arena), every slot is padded to
    256 B.  Returns (offsets: name -> (offset, numel) in floats, n_active, total)."""
    named = list(model.named_parameters())
    e = Engine.__new__(Engine)
    active = [(n, p) for n, p in named if Engine.produces_grad(e, n, p, model)]
    inactive = [(n, p) for n, p in named if not Engine.produces_grad(e, n, p, model)]
    offsets, off, n_active = {}, 0, 0
    for i, (n, p) in enumerate(active + inactive):
        offsets[n] = (off, p.numel())
        off += (p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        if i == len(active) - 1:
            n_active = off
    return offsets, n_active, off


class Engine:
    _registry = weakref.WeakKeyDictionary()

    @classmethod
    def of(cls, model):
        e = cls._registry.get(model)
        if e is None:
            e = cls(model)
            cls._registry[model] = e
        return e

    def __init__(self, model):
        self.model_ref = weakref.ref(model)
        self.handles = {}
        self.flat_params = None
        self.flat_grads = None
        self.n_active = 0
        self.direct_grads = False  # True: p.grad are views of flat_grads and autograd returns None for parameters
        self.views = {}
        self.grad_views = {}
        self._seed = itertools.count(1)
        self._cb_keepalive = []
        self.world_size_override = None
        self.bucket_events = None
        self.seed_dev = None   # device-resident dropout seed (enable_device_seed): CUDA-graph replays

    # ------------------------------------------------------------------------------------------ arenas
    def _needs_materialize(self, model):
        if self.flat_params is None:
            return True
        lo = self.flat_params.data_ptr()
        hi = lo + self.flat_params.numel() * 4
        for p in model.parameters():
            if not (lo <= p.data_ptr() < hi):
                return True
        return False

    def produces_grad(self, name, p, model):
        if not p.requires_grad:
            return False
        if getattr(model, "dense", None) == "mae":
            return True  # every trainable tensor of MaskedAutoencoderViT is used (the sin-cos tables are frozen)
        if getattr(model, "dense", None):
            if _unused_in_dense(name):
                return False
            if getattr(model, "frozen", False) and not name.startswith("decoder."):
                return False
        elif getattr(model, "frozen", False) and not name.startswith("lin_head."):
            return False  # frozen classification: the encoder runs under no_grad (Models/models.py:341-343)
        return True

    @torch.no_grad()
    def materialize(self):
        model = self.model_ref()
        if not self._needs_materialize(model):
            return
        named = list(model.named_parameters())
        dev = named[0][1].device
        if dev.type != "cuda":
            raise RuntimeError("ssl4gie_b200: move the model to a CUDA device first — there is no CPU path")
        _C.init(dev.index or 0)
        for _, p in named:
            if p.dtype != torch.float32:
                raise RuntimeError("ssl4gie_b200 keeps fp32 master parameters; do not .half()/.bfloat16() the model")
        offsets, n_active, total = arena_layout(model)
        flat = torch.zeros(total, dtype=torch.float32, device=dev)
        grads = torch.zeros(n_active, dtype=torch.float32, device=dev)
        self.views, self.grad_views = {}, {}
        for n, p in named:
            off = offsets[n][0]
            v = flat[off:off + p.numel()].view_as(p)
            v.copy_(p.data)
            p.data = v
            self.views[n] = v
            if off < n_active:
                self.grad_views[n] = grads[off:off + p.numel()].view_as(p)
        self.flat_params, self.flat_grads, self.n_active = flat, grads, n_active
        self.active_names = [n for n, _ in named if offsets[n][0] < n_active]
        for h in self.handles.values():
            h.destroy()
        self.handles = {}
        if self.direct_grads:
            self.enable_direct_grads()

    def enable_device_seed(self):
        """Dropout seed of the seg head read from device memory (advanced by FusedAdamW's device-state step), so that a captured
        graph of the step draws a new mask at every replay."""
        if self.seed_dev is None:
            self.materialize()
            seed = (torch.initial_seed() * 1000003 + 12345) & 0x7FFFFFFFFFFFFFFF
            self.seed_dev = torch.tensor([seed], dtype=torch.int64, device=self.flat_params.device)
            for h in self.handles.values():
                _C.check(_C.lib().s4g_model_set_seed_device(h.ptr, _C.ptr(self.seed_dev)))
        return self

    def enable_direct_grads(self):
        """Fast path used by FusedAdamW / DataParallelStep: p.grad aliases the flat gradient arena."""
        self.direct_grads = True
        if self.flat_params is None:
            return
        model = self.model_ref()
        for n, p in model.named_parameters():
            if n in self.grad_views:
                p.grad = self.grad_views[n]

    # ------------------------------------------------------------------------------------------ handles
    def config(self, model):
        prec = os.environ.get("SSL4GIE_B200_PRECISION", getattr(model, "precision", "auto"))
        if prec == "auto":
            prec = "bf16" if torch.is_autocast_enabled() else "fp32"
        backend = os.environ.get("SSL4GIE_B200_BACKEND", getattr(model, "backend", "auto"))
        if backend == "auto":
            backend = "tensor" if prec == "bf16" else "simt"
        return prec, backend

    def world_size(self, model):
        if self.world_size_override is not None:
            return self.world_size_override
        import torch.distributed as dist
        if any(isinstance(m, nn.SyncBatchNorm) for m in model.modules()) and dist.is_available() and dist.is_initialized():
            return dist.get_world_size()
        return 1

    def handle(self, batch):
        model = self.model_ref()
        self.materialize()
        prec, backend = self.config(model)
        # Evaluation never exchanges BatchNorm statistics (running stats are used), so an eval-mode handle is a
        # world-size-1 handle: building it involves no collective and the reference's "rank 0 evaluates while the
        # others wait in a barrier" (train_segmentation.py:293-301) cannot deadlock on handle construction.
        ws = self.world_size(model) if model.training else 1
        key = (batch, prec, backend, ws, bool(getattr(model, "frozen", False)))
        h = self.handles.get(key)
        if h is None:
            h = _Handle(self, model, batch, prec, backend, ws)
            self.handles[key] = h
        return h


class _Handle:
    def __init__(self, engine, model, batch, prec, backend, world_size):
        lib = _C.lib()
        self.engine = engine
        dense = getattr(model, "dense", None)
        drop = 0.0
        if dense == "seg":
            drop = float(model.decoder.output_conv[3].p)
        desc = _C.S4GModelDesc(batch=batch, dense=_C.DENSE[dense], dtype=_C.BF16 if prec == "bf16" else _C.F32,
                               backend=_C.TENSOR if backend == "tensor" else _C.SIMT,
                               frozen=int(bool(getattr(model, "frozen", False))), world_size=world_size,
                               num_classes=int(getattr(model, "num_classes_head", 0) or 0),
                               out_token=0 if getattr(model, "out_token", "cls") == "cls" else 1, dropout_p=drop)
        self.ptr = ctypes.c_void_p()
        _C.check(lib.s4g_model_create(ctypes.byref(desc), ctypes.byref(self.ptr)))
        self.batch, self.dense = batch, dense
        self.out_features = int(getattr(model, "num_classes_head", 0) or 0) or 768
        for name, p in model.named_parameters():
            g = engine.grad_views.get(name)
            _C.check(lib.s4g_model_bind(self.ptr, name.encode(), _C.ptr(p.data), _C.ptr(g), p.numel()))
        for name, b in model.named_buffers():
            _C.check(lib.s4g_model_bind(self.ptr, name.encode(), _C.ptr(b), None, b.numel()))
        nbytes = lib.s4g_model_workspace_bytes(self.ptr)
        if nbytes == 0:
            raise RuntimeError("libssl4gie_b200: " + lib.s4g_last_error().decode())
        dev = engine.flat_params.device
        self.workspace = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
        base = self.workspace.data_ptr()
        self.ws_off = (-base) % 256
        _C.check(lib.s4g_model_set_workspace(self.ptr, ctypes.c_void_p(base + self.ws_off), nbytes))
        self.buffers = [b for _, b in model.named_buffers()]
        if engine.seed_dev is not None:
            _C.check(lib.s4g_model_set_seed_device(self.ptr, _C.ptr(engine.seed_dev)))
        self.syncbn_via = None  # "peer" (NVLink mailboxes, csrc/peer.cu) | "nccl" (callback) when world_size > 1
        if world_size > 1:
            self._install_allreduce()

    def _install_allreduce(self):
        import torch.distributed as dist
        from .peer import shared_mailboxes, use_peer_exchange
        if use_peer_exchange() and dist.get_world_size() <= 8 and dist.get_backend() == "nccl":
            # one-CTA NVLink exchange kernel instead of an NCCL collective per statistic vector (csrc/peer.cu).  The
            # mailboxes are created ONCE per process (collective: the first training forward, which every rank
            # makes) and shared by every later handle, whose construction is then collective-free.
            boxes = shared_mailboxes()
            if boxes.ok:
                _C.check(_C.lib().s4g_model_set_peer_exchange(self.ptr, boxes.ptrs, boxes.rank, boxes.world))
                self.syncbn_via = "peer"
                return
            print("ssl4gie_b200: peer mailboxes unavailable (%s); SyncBN statistics go through NCCL" % getattr(boxes, "error", "a peer failed"))
        ws_f64 = self.workspace[self.ws_off:self.ws_off + (self.workspace.numel() - 256) // 8 * 8].view(torch.float64)
        base = self.workspace.data_ptr() + self.ws_off

        def cb(user, dptr, count, stream):
            try:
                off = (dptr - base) // 8
                dist.all_reduce(ws_f64[off:off + count])
                return 0
            except Exception as e:  # noqa: BLE001
                print("ssl4gie_b200 SyncBN all-reduce failed:", e)
                return -6

        self._cb = _C.ALLREDUCE_FN(cb)
        self.syncbn_via = "nccl"
        _C.check(_C.lib().s4g_model_set_allreduce(self.ptr, self._cb, None))

    def destroy(self):
        if self.ptr:
            _C.lib().s4g_model_destroy(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:  # noqa: BLE001
            pass

    def forward(self, x, training, seed, mask):
        B = x.shape[0]
        if self.dense:
            out = torch.empty((B, 1, 224, 224), dtype=torch.float32, device=x.device)
        else:
            out = torch.empty((B, self.out_features), dtype=torch.float32, device=x.device)
        _C.check(_C.lib().s4g_model_forward(self.ptr, _C.ptr(x), _C.ptr(out), int(training), seed, _C.ptr(mask), _C.stream_ptr()))
        return out

    def forward_mae(self, x, noise):
        """MaskedAutoencoderViT up to `pred`: returns (pred [B,196,768], mask [B,196], ids_restore [B,196] int32)."""
        B = x.shape[0]
        pred = torch.empty((B, 196, 768), dtype=torch.float32, device=x.device)
        mask = torch.empty((B, 196), dtype=torch.float32, device=x.device)
        ids = torch.empty((B, 196), dtype=torch.int32, device=x.device)
        lib = _C.lib()
        _C.check(lib.s4g_model_set_mae_io(self.ptr, _C.ptr(noise), _C.ptr(mask), _C.ptr(ids)))
        _C.check(lib.s4g_model_forward(self.ptr, _C.ptr(x), _C.ptr(pred), 1, 0, None, _C.stream_ptr()))
        return pred, mask, ids

    def backward(self, gout):
        ev = self.engine.bucket_events
        if ev is not None and torch.cuda.is_current_stream_capturing():
            # a captured single-GPU step (graph.GraphedStep) has no exchange to signal: record no bucket events into the graph
            _C.check(_C.lib().s4g_model_set_bucket_events(self.ptr, None, 0))
        elif ev is not None:
            arr = (ctypes.c_void_p * len(ev))(*[ctypes.c_void_p(e.cuda_event) for e in ev])
            _C.check(_C.lib().s4g_model_set_bucket_events(self.ptr, arr, len(ev)))
        _C.check(_C.lib().s4g_model_backward(self.ptr, _C.ptr(gout), _C.stream_ptr()))

    def launches(self):
        return _C.lib().s4g_model_last_launches(self.ptr)

    def bucket_params(self, i, n):
        return [s for s in _C.lib().s4g_model_bucket_params(self.ptr, i, n).decode().split("\n") if s]


class _ModelFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, engine, handle, x, seed, mask, *params):
        ctx.engine, ctx.handle = engine, handle
        ctx.n = len(params)
        return handle.forward(x, True, seed, mask)

    @staticmethod
    def backward(ctx, gout):
        e = ctx.engine
        ctx.handle.backward(gout.contiguous().float())
        if e.direct_grads:
            return (None,) * (5 + ctx.n)
        grads = []
        for n, p in e.model_ref().named_parameters():
            g = e.grad_views.get(n)
            grads.append(g.clone() if g is not None else None)
        return (None, None, None, None, None, *grads)


class _MaeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, engine, handle, x, noise, *params):
        ctx.engine, ctx.handle = engine, handle
        ctx.n = len(params)
        pred, mask, ids = handle.forward_mae(x, noise)
        ctx.mark_non_differentiable(mask, ids)
        return pred, mask, ids

    @staticmethod
    def backward(ctx, gpred, _gmask, _gids):
        e = ctx.engine
        ctx.handle.backward(gpred.contiguous().float())
        if e.direct_grads:
            return (None,) * (4 + ctx.n)
        grads = []
        for n, p in e.model_ref().named_parameters():
            g = e.grad_views.get(n)
            grads.append(g.clone() if g is not None else None)
        return (None, None, None, None, *grads)


def run_mae(model, x, noise):
    """MaskedAutoencoderViT.forward_encoder + forward_decoder (Models/mae/models_mae.py:150-196) with the masking noise given:
    (pred [B,196,768] fp32, mask [B,196], ids_restore [B,196] int32)."""
    _C.require_cuda(x, "input")
    if x.dim() != 4 or tuple(x.shape[1:]) != (3, 224, 224):
        raise AssertionError(f"Input image size ({tuple(x.shape[2:])}) doesn't match model (224, 224).")
    if tuple(noise.shape) != (x.shape[0], 196):
        raise ValueError("noise must be [B, 196]")
    engine = Engine.of(model)
    x = x.detach().float().contiguous()
    noise = noise.detach().float().contiguous()
    h = engine.handle(x.shape[0])
    if torch.is_grad_enabled():
        params = [p for _, p in model.named_parameters()]
        return _MaeFn.apply(engine, h, x, noise, *params)
    return h.forward_mae(x, noise)


def run_model(model, x):
    """model(x): [B,3,224,224] float -> [B,1,224,224] float32 (dense) or [B,num_classes] / [B,768] (classification)."""
    _C.require_cuda(x, "input")
    if x.dim() != 4 or x.shape[1] != 3 or x.shape[2] != 224 or x.shape[3] != 224:
        # timm PatchEmbed asserts the input size (reference Models/models.py:326)
        raise AssertionError(f"Input image size ({tuple(x.shape[2:])}) doesn't match model (224, 224).")
    engine = Engine.of(model)
    x = x.detach().float().contiguous()
    h = engine.handle(x.shape[0])
    training = model.training
    mask = getattr(model, "_parity_dropout_mask", None)
    if training and torch.is_grad_enabled():
        seed = (torch.initial_seed() * 1000003 + next(engine._seed)) & 0xFFFFFFFFFFFFFFFF
        params = [p for _, p in model.named_parameters()]
        return _ModelFn.apply(engine, h, x, seed, mask, *params)
    return h.forward(x, training, next(engine._seed), mask if training else None)
