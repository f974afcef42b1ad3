"""The whole training step -- forward, loss, backward, AdamW -- captured once in a CUDA graph and replayed.

The C++ schedule is allocation- and sync-free, so the only things a captured graph would freeze are the scalars that the eager
path passes by value: the learning rate and Adam's step count (bias corrections) and the seg head's dropout seed.  GraphedStep
moves them to device memory (FusedAdamW.enable_device_state, Engine.enable_device_seed), runs the first calls eagerly (one-time
set-up: kernel attributes, tensor maps, handle construction), captures the next call with torch.cuda.graph and replays from then
on.  Learning-rate changes made through optimizer.param_groups (a scheduler, ssl4gie_b200.adjust_learning_rate) are pushed to the
device before every replay.  One GPU; the data-parallel exchange is not captured."""
import torch

from .engine import Engine


class GraphedStep:
    """step = GraphedStep(model, loss_fn, optimizer); loss = step(x, t)

    model: an ssl4gie_b200 backbone (or MaskedAutoencoderViT with loss_fn=None: the model returns its loss);
    optimizer: ssl4gie_b200.FusedAdamW.  x, t must keep their shapes and dtypes; they are copied into static buffers.  The returned
    loss is a static device tensor, overwritten by the next call.  Call under the same torch.autocast context every time."""

    def __init__(self, model, loss_fn, optimizer, eager_steps=2):
        core = model.module if hasattr(model, "module") else model
        if torch.distributed.is_available() and torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1:
            raise NotImplementedError("GraphedStep captures the single-GPU step; use the eager step with GradientAllReduce for N > 1")
        if not hasattr(optimizer, "enable_device_state"):
            raise TypeError("GraphedStep needs ssl4gie_b200.FusedAdamW")
        self.model, self.core, self.loss_fn, self.optimizer = model, core, loss_fn, optimizer
        self.engine = Engine.of(core)
        optimizer.enable_device_state()
        if getattr(core, "dense", None) == "seg":
            self.engine.enable_device_seed()
        self.eager_steps, self.calls = eager_steps, 0
        self.graph = self.x = self.t = self.loss = None

    def _step(self, x, t):
        if self.loss_fn is None:
            loss = self.core(x, noise=t)[0] if getattr(self.core, "dense", None) == "mae" and t is not None else self.core(x)[0]
        else:
            loss = self.loss_fn(self.model(x), t)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    def __call__(self, x, t=None):
        self.calls += 1
        if self.calls <= self.eager_steps:
            return self._step(x, t)
        if self.graph is None:
            self.x = x.detach().clone()
            self.t = None if t is None else t.detach().clone()
            self.autocast = torch.is_autocast_enabled()
            self.optimizer.sync_device_hyper()
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):     # nothing executes here; the first replay below performs this call's step
                self.loss = self._step(self.x, self.t)
            self.optimizer.step_count -= 1          # counted again by the replay below
        else:
            if torch.is_autocast_enabled() != self.autocast:
                raise RuntimeError("GraphedStep: the autocast context changed since capture")
            self.x.copy_(x, non_blocking=True)
            if t is not None:
                self.t.copy_(t, non_blocking=True)
        self.optimizer.sync_device_hyper()
        self.graph.replay()
        self.optimizer.step_count += 1
        return self.loss
