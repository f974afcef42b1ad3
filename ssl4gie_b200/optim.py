"""Fused AdamW over the engine's flat parameter arena (reference: torch.optim.AdamW at train_segmentation.py:204)."""
import torch

from . import _C


class FusedAdamW(torch.optim.Optimizer):
    """AdamW with torch's defaults; one kernel per step over the model's flat fp32 arena.

    ``model`` must be an ssl4gie_b200 backbone; call after ``model.cuda()``.  Parameters that never receive a
    gradient (``norm.*``, ``decoder.refinenet4.resConfUnit1.*`` in dense mode) are left untouched, as torch does
    for ``grad is None``.
    """

    def __init__(self, model, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01, grad_scale=1.0, param_groups=None):
        from .engine import Engine
        core = model.module if hasattr(model, "module") else model
        self.engine = Engine.of(core)
        self.engine.materialize()
        # p.grad aliases the flat gradient arena: backward writes it in place (no per-parameter clone, no accumulation
        # into stale .grad tensors -- zero_grad() below is a no-op because every backward overwrites)
        self.engine.enable_direct_grads()
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        # every parameter, in model.parameters() order, as the reference's torch.optim.AdamW(model.parameters()) holds
        # them: the indices in state_dict() are then the reference checkpoint's indices
        # param_groups: torch-style groups of the model's parameters that differ in weight_decay only (timm add_weight_decay,
        # Models/mae/main_pretrain.py:179-180); lr / betas / eps are shared by all groups
        super().__init__(list(core.parameters()) if param_groups is None else param_groups, defaults)
        n = self.engine.n_active
        self.wd_slots = None
        if param_groups is not None:
            self._build_wd_slots()
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=self.engine.flat_params.device)
        self.exp_avg_sq = torch.zeros_like(self.exp_avg)
        self.step_count = 0
        self.grad_scale = grad_scale
        # device-resident scalars (enable_device_state): what a CUDA graph of the step needs (graph.GraphedStep)
        self._hyper_dev = None
        self._step_dev = None
        self._hyper_written = None

    def _build_wd_slots(self):
        """One weight decay per 64-float arena slot (every parameter starts on a slot boundary)."""
        e = self.engine
        base = e.flat_grads.data_ptr()
        by_id = {id(p): g for n, p in e.model_ref().named_parameters() if (g := e.grad_views.get(n)) is not None}
        wd = torch.zeros(e.n_active // 64, dtype=torch.float32)
        seen = 0
        for group in self.param_groups:
            for p in group["params"]:
                g = by_id.get(id(p))
                if g is None:
                    continue
                off = (g.data_ptr() - base) // 4
                wd[off // 64:(off + p.numel() + 63) // 64] = group["weight_decay"]
                seen += 1
        if seen != len(by_id):
            raise ValueError("param_groups must cover every trainable parameter of the model (got %d of %d)" % (seen, len(by_id)))
        self.wd_slots = wd.to(e.flat_params.device)
        self._wd_key = tuple(g["weight_decay"] for g in self.param_groups)

    # ---- bucket-wise stepping (GradientAllReduce.attach(optimizer, fused_step=True)) ----
    def begin_external_step(self):
        """The next step is applied range by range through step_range(); step() then only acknowledges it."""
        self.step_count += 1
        self._external = True

    @torch.no_grad()
    def step_range(self, lo, hi):
        """AdamW on floats [lo, hi) of the arena (lo on a 64-float slot boundary; hi is rounded up to one: padding is zero)."""
        g = self.param_groups[0]
        e = self.engine
        hi = min((hi + 63) // 64 * 64, e.n_active)
        if lo % 64:
            raise ValueError("step_range: lo must be a multiple of 64")
        import ctypes

        def ptr(t):
            return ctypes.c_void_p(t.data_ptr() + 4 * lo)

        if self.wd_slots is not None:
            wd = ctypes.c_void_p(self.wd_slots.data_ptr() + 4 * (lo // 64))
            _C.check(_C.lib().s4g_adamw_step_groups(ptr(e.flat_params), ptr(e.flat_grads), ptr(self.exp_avg), ptr(self.exp_avg_sq),
                                                    hi - lo, g["lr"], g["betas"][0], g["betas"][1], g["eps"], wd, self.step_count,
                                                    self.grad_scale, _C.stream_ptr()))
        else:
            _C.check(_C.lib().s4g_adamw_step(ptr(e.flat_params), ptr(e.flat_grads), ptr(self.exp_avg), ptr(self.exp_avg_sq), hi - lo,
                                             g["lr"], g["betas"][0], g["betas"][1], g["eps"], g["weight_decay"], self.step_count,
                                             self.grad_scale, _C.stream_ptr()))

    # ---- device-resident lr / betas / step count (CUDA-graph capture of the step) ----
    def enable_device_state(self):
        """Keep lr, betas, eps, weight decay, grad_scale and the step count in device memory: step() then launches kernels whose
        arguments never change, so the step can be captured once and replayed (ssl4gie_b200.graph.GraphedStep).  Per-group weight
        decay still comes from the slot table.  Same arithmetic except that the bias corrections are evaluated on the device."""
        if self._hyper_dev is None:
            dev = self.engine.flat_params.device
            self._hyper_dev = torch.zeros(8, dtype=torch.float32, device=dev)
            self._step_dev = torch.full((1,), self.step_count, dtype=torch.int32, device=dev)
        return self

    def sync_device_hyper(self):
        """Push lr / betas / eps / weight_decay / grad_scale to the device if they changed (a 1-thread kernel with by-value
        arguments; call outside graph capture: GraphedStep does before every replay)."""
        g = self.param_groups[0]
        cur = (float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]), float(self.grad_scale))
        if cur != self._hyper_written:
            _C.check(_C.lib().s4g_device_hyper_set(_C.ptr(self._hyper_dev), *cur, _C.stream_ptr()))
            self._hyper_written = cur

    def _step_device(self):
        e = self.engine
        if not torch.cuda.is_current_stream_capturing():
            self.sync_device_hyper()
        wd = _C.ptr(self.wd_slots) if self.wd_slots is not None else None
        _C.check(_C.lib().s4g_adamw_step_device(_C.ptr(e.flat_params), _C.ptr(e.flat_grads), _C.ptr(self.exp_avg), _C.ptr(self.exp_avg_sq),
                                                e.n_active, _C.ptr(self._hyper_dev), _C.ptr(self._step_dev), wd, _C.stream_ptr()))
        seed = getattr(e, "seed_dev", None)
        _C.check(_C.lib().s4g_device_counter_advance(_C.ptr(self._step_dev), _C.ptr(seed), _C.stream_ptr()))

    @torch.no_grad()
    def step(self, closure=None):
        if getattr(self, "_external", False):   # already applied bucket by bucket during the gradient exchange
            self._external = False
            return
        if self._hyper_dev is not None:
            self.step_count += 1
            return self._step_device()
        g = self.param_groups[0]
        self.step_count += 1
        e = self.engine
        if self.wd_slots is not None:
            for o in self.param_groups[1:]:
                if (o["lr"], tuple(o["betas"]), o["eps"]) != (g["lr"], tuple(g["betas"]), g["eps"]):
                    raise ValueError("FusedAdamW: parameter groups may differ in weight_decay only")
            if tuple(o["weight_decay"] for o in self.param_groups) != self._wd_key:
                self._build_wd_slots()
            _C.check(_C.lib().s4g_adamw_step_groups(_C.ptr(e.flat_params), _C.ptr(e.flat_grads), _C.ptr(self.exp_avg),
                                                    _C.ptr(self.exp_avg_sq), e.n_active, g["lr"], g["betas"][0], g["betas"][1],
                                                    g["eps"], _C.ptr(self.wd_slots), self.step_count, self.grad_scale, _C.stream_ptr()))
            return
        _C.check(_C.lib().s4g_adamw_step(_C.ptr(e.flat_params), _C.ptr(e.flat_grads), _C.ptr(self.exp_avg), _C.ptr(self.exp_avg_sq),
                                         e.n_active, g["lr"], g["betas"][0], g["betas"][1], g["eps"], g["weight_decay"],
                                         self.step_count, self.grad_scale, _C.stream_ptr()))

    def zero_grad(self, set_to_none=False):
        # gradients are overwritten by every backward; nothing to clear (keeps the flat views alive)
        return None

    def _slots(self):
        """(index in model.parameters(), offset, parameter) of every parameter that has optimizer state."""
        e = self.engine
        base = e.flat_grads.data_ptr()
        out = []
        for i, (n, p) in enumerate(e.model_ref().named_parameters()):
            g = e.grad_views.get(n)
            if g is not None:
                out.append((i, (g.data_ptr() - base) // 4, p))
        return out

    def state_dict(self):
        """torch.optim.AdamW's format (what the reference stores under "optimizer_state_dict",
        train_segmentation.py:333): per-parameter step / exp_avg / exp_avg_sq keyed by the index in
        model.parameters(); parameters that never received a gradient have no entry, as in torch."""
        n_params = len(self.param_groups[0]["params"])
        proto = torch.optim.AdamW([torch.nn.Parameter(torch.zeros(1))]).state_dict()["param_groups"][0]
        group = dict(proto)
        group.update({k: v for k, v in self.param_groups[0].items() if k != "params"})
        group["params"] = list(range(n_params))
        state = {}
        if self.step_count > 0:
            for i, off, p in self._slots():
                state[i] = {"step": torch.tensor(float(self.step_count)),
                            "exp_avg": self.exp_avg[off:off + p.numel()].view_as(p).clone(),
                            "exp_avg_sq": self.exp_avg_sq[off:off + p.numel()].view_as(p).clone()}
        return {"state": state, "param_groups": [group]}

    @torch.no_grad()
    def load_state_dict(self, sd):
        """Accepts a torch.optim.AdamW state_dict of the same model (a reference checkpoint) or this class's."""
        if "state" not in sd:  # flat format written by earlier versions of this class
            self.step_count = sd["step"]
            self.exp_avg.copy_(sd["exp_avg"])
            self.exp_avg_sq.copy_(sd["exp_avg_sq"])
            return
        g = sd["param_groups"][0]
        if len(g["params"]) != len(self.param_groups[0]["params"]):
            raise ValueError("loaded state dict contains a parameter group that doesn't match the size of optimizer's group")
        for k in ("lr", "betas", "eps", "weight_decay"):
            if k in g:
                self.param_groups[0][k] = tuple(g[k]) if k == "betas" else g[k]
        order = {pid: i for i, pid in enumerate(g["params"])}
        self.exp_avg.zero_()
        self.exp_avg_sq.zero_()
        steps = set()
        slots = {i: (off, p) for i, off, p in self._slots()}
        for pid, st in sd["state"].items():
            i = order[pid]
            if i not in slots:
                continue  # state of a parameter this schedule never updates
            off, p = slots[i]
            self.exp_avg[off:off + p.numel()].copy_(st["exp_avg"].reshape(-1))
            self.exp_avg_sq[off:off + p.numel()].copy_(st["exp_avg_sq"].reshape(-1))
            steps.add(int(st["step"]))
        if len(steps) > 1:
            raise ValueError("FusedAdamW keeps one step count; the loaded state has %s" % sorted(steps))
        self.step_count = steps.pop() if steps else 0
