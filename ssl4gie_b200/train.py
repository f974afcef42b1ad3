"""Sync-free drop-ins for the reference's epoch loops and checkpoint dict.

Reference: ``train_epoch`` / ``test`` / the ``torch.save`` dict in Binary_segmentation/train_segmentation.py:25-81,
84-126, 185-197, 328-342 (Depth_estimation/train_depth.py and Classification/train_classification.py have the same
shape).  Same arguments, same log lines, same return values, same checkpoint keys -- but the step never waits for the
GPU: batches arrive through ``DevicePrefetcher`` (copy of batch i+1 under the compute of batch i), the per-step loss
stays on the device (one all-reduce + one read-back per epoch instead of an all_reduce + ``.item()`` + ``barrier`` per
step), progress lines print a loss that was read back asynchronously a few steps earlier, and autocast / GradScaler are
not needed (bf16 operands with fp32 accumulation and fp32 master weights have fp32's exponent range).

Precision: the reference runs these loops under ``torch.cuda.amp.autocast`` (train_segmentation.py:45).  Here the loops
select the bf16 tcgen05 schedule themselves: a model whose ``precision`` attribute is still ``"auto"`` runs as
``"bf16"`` inside ``train_epoch`` / ``test`` (outside them ``"auto"`` means bf16 under ``torch.autocast`` and the fp32
FFMA validation kernels otherwise).  Set ``model.precision = "fp32"`` (or SSL4GIE_B200_PRECISION=fp32) to run the loops
in the validation mode.

Multi-GPU: pass ``grad_sync=ssl4gie_b200.ddp.GradientAllReduce(model)``; it prepares itself on the first ``reduce()``
and folds the 1/world_size mean into ``FusedAdamW`` (``grad_sync.attach(optimizer)`` is called here).
"""
import contextlib
import random
import time

import numpy as np
import torch

from .prefetch import DevicePrefetcher


def _is_dist():
    import torch.distributed as dist
    return dist.is_available() and dist.is_initialized()


@contextlib.contextmanager
def _bf16_by_default(model):
    """Resolve precision "auto" to the bf16 tensor-core schedule for the duration of an epoch loop."""
    core = model.module if hasattr(model, "module") else model
    was = getattr(core, "precision", None)
    if was == "auto":
        core.precision = "bf16"
    try:
        yield
    finally:
        if was == "auto":
            core.precision = was


class _LaggedScalar:
    """Asynchronous device -> pinned-host read-back of a scalar; ``value`` is the latest copy that has completed."""

    def __init__(self, slots=4):
        self.host = torch.zeros(slots, dtype=torch.float32).pin_memory()
        self.events = [None] * slots
        self.i = 0
        self.value = float("nan")

    def push(self, t):
        k = self.i % len(self.events)
        self.host[k:k + 1].copy_(t.detach().reshape(1), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self.events[k] = ev
        self.i += 1
        for j in range(1, len(self.events) + 1):  # newest completed copy
            q = (self.i - j) % len(self.events)
            if self.events[q] is not None and self.events[q].query():
                self.value = float(self.host[q])
                break


def train_epoch(model, rank, world_size, train_loader, train_sampler, optimizer, epoch, loss_fn, log_path, scaler=None,
                grad_sync=None, log_every=10, graph_step=None):
    """One training epoch; returns the epoch-mean of the rank-averaged loss (reference train_segmentation.py:25-81).

    ``scaler`` is accepted for signature compatibility and ignored.  ``grad_sync`` is the
    ``ssl4gie_b200.ddp.GradientAllReduce`` of the model when world_size > 1 (the reference wraps the model in DDP).
    ``graph_step`` (one GPU, constant batch shape -- the reference's loaders drop the last batch): a
    ``ssl4gie_b200.GraphedStep(model, loss_fn, optimizer)`` kept by the caller across epochs; every step is then one CUDA-graph
    replay instead of ~600 kernel launches from Python."""
    with _bf16_by_default(model):
        return _train_epoch(model, rank, world_size, train_loader, train_sampler, optimizer, epoch, loss_fn, log_path,
                            grad_sync, log_every, graph_step)


def _train_epoch(model, rank, world_size, train_loader, train_sampler, optimizer, epoch, loss_fn, log_path, grad_sync,
                 log_every, graph_step=None):
    t = time.time()
    model.train()
    if train_sampler is not None:
        train_sampler.set_epoch(epoch - 1)
    if grad_sync is not None and hasattr(grad_sync, "attach"):
        grad_sync.attach(optimizer)
    n_batches = len(train_loader)
    n_data = len(train_loader.dataset) if hasattr(train_loader, "dataset") else None
    losses = None
    shown = _LaggedScalar()
    seen = 0
    for batch_idx, (data, target) in enumerate(DevicePrefetcher(train_loader, device=torch.device("cuda", rank))):
        if losses is None:
            losses = torch.zeros(n_batches, dtype=torch.float32, device=data.device)
        if graph_step is not None:
            loss = graph_step(data, target)
        else:
            optimizer.zero_grad()
            loss = loss_fn(model(data), target)
            loss.backward()
            if grad_sync is not None:
                grad_sync.reduce()
            optimizer.step()
        losses[batch_idx] = loss.detach()
        seen += len(data) * world_size
        if rank == 0 and batch_idx + 1 < n_batches and (batch_idx + 1) % log_every == 0:
            shown.push(loss)
            print("\rTrain Epoch: {} [{}/{} ({:.1f}%)]\tLoss: {:.6f}\tTime: {:.6f}".format(
                epoch, seen, n_data if n_data is not None else n_batches * len(data) * world_size,
                100.0 * (batch_idx + 1) / n_batches, shown.value, time.time() - t), end="")
    if losses is None:
        return float("nan")
    if world_size > 1 and _is_dist():
        import torch.distributed as dist
        dist.all_reduce(losses)
        losses /= world_size
    mean = float(losses.mean().item())  # the epoch's only host synchronisation
    if rank == 0:
        printout = "Train Epoch: {} [{}/{} ({:.1f}%)]\tAverage loss: {:.6f}\tTime: {:.6f}".format(
            epoch, seen, n_data if n_data is not None else seen, 100.0, mean, time.time() - t)
        print("\r" + printout)
        if log_path:
            with open(log_path, "a") as f:
                f.write(printout)
                f.write("\n")
    return mean


@torch.no_grad()
def test(model, rank, test_loader, epoch, perf_fn, log_path):
    """Sample-weighted mean of ``perf_fn(output, target)`` over the loader (reference train_segmentation.py:84-126)."""
    with _bf16_by_default(model):
        return _test(model, rank, test_loader, epoch, perf_fn, log_path)


def _test(model, rank, test_loader, epoch, perf_fn, log_path):
    t = time.time()
    model.eval()
    acc = None
    N = 0
    for data, target in DevicePrefetcher(test_loader, device=torch.device("cuda", rank)):
        n = len(data)
        N += n
        perf = perf_fn(model(data), target)
        perf = perf if torch.is_tensor(perf) else torch.tensor(float(perf), device=data.device)
        acc = perf.float() * n if acc is None else acc + perf.float() * n
    result = float(acc.item()) / N if N else float("nan")
    n_data = len(test_loader.dataset) if hasattr(test_loader, "dataset") else N
    printout = "Test  Epoch: {} [{}/{} ({:.1f}%)]\tAverage performance: {:.6f}\tTime: {:.6f}".format(
        epoch, N, n_data, 100.0, result, time.time() - t)
    print("\r" + printout)
    if log_path:
        with open(log_path, "a") as f:
            f.write(printout)
            f.write("\n")
    return result


def _default_scaler_state():
    """State of a fresh, enabled GradScaler: the reference's build() loads "scaler_state_dict" into one unconditionally
    (train_segmentation.py:205-207) and torch refuses an empty dict there."""
    return {"scale": 65536.0, "growth_factor": 2.0, "backoff_factor": 0.5, "growth_interval": 2000, "_growth_tracker": 0}


def save_checkpoint(ckpt_path, epoch, model, optimizer, scaler, loss, val_perf, test_perf):
    """The reference's checkpoint dict, key for key (train_segmentation.py:328-342): loadable by the reference's
    build() (:185-206) and by eval_segmentation.py.  ``model`` may be wrapped (``.module``)."""
    core = model.module if hasattr(model, "module") else model
    torch.save({
        "epoch": epoch,
        "model_state_dict": core.state_dict(),
        "optimizer_state_dict": optimizer.state_dict(),
        "scaler_state_dict": scaler.state_dict() if scaler is not None else _default_scaler_state(),
        "loss": loss,
        "val_perf": val_perf,
        "test_perf": test_perf,
        "py_state": random.getstate(),
        "np_state": np.random.get_state(),
        "torch_state": torch.get_rng_state(),
    }, ckpt_path)


def load_checkpoint(ckpt_path, model, optimizer=None, scaler=None):
    """Resume as the reference's build() does (train_segmentation.py:185-206): returns (start_epoch, prev_best_test)."""
    main_dict = torch.load(ckpt_path, map_location="cpu", weights_only=False)
    core = model.module if hasattr(model, "module") else model
    core.load_state_dict(main_dict["model_state_dict"])
    random.setstate(main_dict["py_state"])
    np.random.set_state(main_dict["np_state"])
    torch.set_rng_state(main_dict["torch_state"])
    if optimizer is not None:
        optimizer.load_state_dict(main_dict["optimizer_state_dict"])
    if scaler is not None and main_dict.get("scaler_state_dict"):
        scaler.load_state_dict(main_dict["scaler_state_dict"])
    return main_dict["epoch"] + 1, main_dict["val_perf"]
