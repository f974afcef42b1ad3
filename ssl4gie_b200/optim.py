"""Fused AdamW over the engine's flat parameter arena (reference: torch.optim.AdamW at train_segmentation.py:204)."""
import torch

from . import _C


class FusedAdamW(torch.optim.Optimizer):
    """AdamW with torch's defaults; one kernel per step over the model's flat fp32 arena.

    ``model`` must be an ssl4gie_b200 backbone; call after ``model.cuda()``.  Parameters that never receive a
    gradient (``norm.*``, ``decoder.refinenet4.resConfUnit1.*`` in dense mode) are left untouched, as torch does
    for ``grad is None``.
    """

    def __init__(self, model, lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01, grad_scale=1.0):
        from .engine import Engine
        core = model.module if hasattr(model, "module") else model
        self.engine = Engine.of(core)
        self.engine.materialize()
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__([p for p in core.parameters() if p.requires_grad], defaults)
        n = self.engine.n_active
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=self.engine.flat_params.device)
        self.exp_avg_sq = torch.zeros_like(self.exp_avg)
        self.step_count = 0
        self.grad_scale = grad_scale

    @torch.no_grad()
    def step(self, closure=None):
        g = self.param_groups[0]
        self.step_count += 1
        e = self.engine
        _C.check(_C.lib().s4g_adamw_step(_C.ptr(e.flat_params), _C.ptr(e.flat_grads), _C.ptr(self.exp_avg), _C.ptr(self.exp_avg_sq),
                                         e.n_active, g["lr"], g["betas"][0], g["betas"][1], g["eps"], g["weight_decay"],
                                         self.step_count, self.grad_scale, _C.stream_ptr()))

    def zero_grad(self, set_to_none=False):
        # gradients are overwritten by every backward; nothing to clear (keeps the flat views alive)
        return None

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq,
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd):
        self.step_count = sd["step"]
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
