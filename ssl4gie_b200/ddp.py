"""Data-parallel gradient exchange for the fused schedule (reference: DDP(model, find_unused_parameters=True),
Binary_segmentation/train_segmentation.py:203, collective C1 in SURVEY.md §2.3).

The C++ backward records one CUDA event per gradient bucket as soon as that bucket's gradients are final
(bucket 0 = DPT decoder, then encoder blocks 11..0 in equal groups, the last bucket also holding patch_embed /
cls_token / pos_embed).  Each bucket is a contiguous slice of the flat fp32 gradient arena, all-reduced with NCCL on
a side stream that waits on the bucket's event, so the exchange overlaps the rest of backward.  The unused
parameters (norm.*, decoder.refinenet4.resConfUnit1.*) are known statically and are simply not in the arena —
no find_unused_parameters graph walk."""
import os

import torch
import torch.distributed as dist

from .engine import Engine


def bucket_slices(names_by_bucket, offsets):
    """[lo, hi) float ranges of the flat gradient arena that make up each bucket: one list of ranges per bucket, a range
    being a run of parameters of that bucket that are adjacent in the arena (alignment padding between them included).
    A bucket is usually ONE range; it is several when the arena order (nn.Module.named_parameters: a module's own
    parameters before its children's, e.g. cls_token / mask_token of MaskedAutoencoderViT) interleaves buckets.  Every
    parameter lies in exactly one range.  names_by_bucket: list of parameter-name lists; offsets: name -> (offset, numel)."""
    owner = {n: b for b, names in enumerate(names_by_bucket) for n in names}
    runs = [[] for _ in names_by_bucket]
    prev = None
    for n in sorted(owner, key=lambda k: offsets[k][0]):
        off, numel = offsets[n]
        b = owner[n]
        if prev == b:
            runs[b][-1][1] = off + numel
        else:
            runs[b].append([off, off + numel])
        prev = b
    return [[(lo, hi) for lo, hi in r] for r in runs]


def allreduce_buckets(flat, slices, group=None, events=None, comm_stream=None, after_bucket=None):
    """Sum-all-reduce every range of every bucket of `flat` (in place).  On CUDA each bucket waits for its event on the
    communication stream; on CPU (gloo, used by the tests) the calls are issued in order.  after_bucket(ranges) is called
    once a bucket's exchanges are enqueued (the fused optimizer step of that bucket)."""
    for i, ranges in enumerate(slices):
        if not ranges:
            continue
        if events is not None and comm_stream is not None:
            comm_stream.wait_event(events[i])
        for lo, hi in ranges:
            dist.all_reduce(flat[lo:hi], group=group)
        if after_bucket is not None:
            after_bucket(ranges)


class GradientAllReduce:
    def __init__(self, model, n_buckets=5, process_group=None):
        core = model.module if hasattr(model, "module") else model
        self.engine = Engine.of(core)
        self.engine.materialize()
        self.engine.enable_direct_grads()
        self.group = process_group
        self.world_size = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.n_buckets = int(os.environ.get("S4G_DDP_BUCKETS", n_buckets))
        self.overlap = os.environ.get("S4G_DDP_OVERLAP", "1") != "0"
        self.comm_stream = None
        self.slices = None
        self.events = None
        self.handles = []
        # fused_step (opt-in, attach(optimizer, fused_step=True)): FusedAdamW updates each bucket on the communication
        # stream right after that bucket's all-reduce, i.e. underneath the rest of backward; the optimizer.step() that
        # follows reduce() is then only the bookkeeping.  Same arithmetic, same result; only for loops that do nothing to
        # the gradients between reduce() and step() (no clipping / unscaling).
        self.fused_optimizer = None
        if self.n_buckets < 1:
            raise ValueError("GradientAllReduce: n_buckets must be >= 1")

    def attach(self, optimizer, fused_step=False):
        """Fold the 1/world_size of DDP's gradient MEAN (train_segmentation.py:203) into the optimizer kernel: the
        all-reduce here is a sum.  Called by train_epoch; call it yourself in a hand-written loop."""
        if fused_step:
            if not hasattr(optimizer, "step_range"):
                raise TypeError("fused_step needs ssl4gie_b200.FusedAdamW")
            self.fused_optimizer = optimizer
        if hasattr(optimizer, "grad_scale"):
            optimizer.grad_scale = 1.0 / self.world_size
        elif self.world_size > 1:
            raise TypeError("GradientAllReduce sums gradients; use ssl4gie_b200.FusedAdamW (grad_scale) or divide "
                            "engine.flat_grads by world_size before optimizer.step()")

    def _setup(self, handle):
        e = self.engine
        offs = {}
        base = e.flat_grads.data_ptr()
        for n, g in e.grad_views.items():
            offs[n] = ((g.data_ptr() - base) // 4, g.numel())
        self.slices = bucket_slices([handle.bucket_params(i, self.n_buckets) for i in range(self.n_buckets)], offs)
        if e.flat_grads.is_cuda:
            self.comm_stream = torch.cuda.Stream()
            self.events = [torch.cuda.Event() for _ in range(self.n_buckets)]
            for ev in self.events:
                ev.record()  # materialise the cudaEvent_t handles
            e.bucket_events = self.events

    def prepare(self, batch):
        """Before the first backward (needs the C model handle for the bucket plan): installs the per-bucket events so
        that the exchange overlaps backward from the first step.  Optional: reduce() prepares itself otherwise, and
        that first step's exchange simply runs after backward."""
        if self.slices is None:
            self._setup(self.engine.handle(batch))

    def reduce(self):
        """After loss.backward(): enqueue the bucket all-reduces (overlapping the tail of backward) and make the
        current stream wait for them."""
        if self.world_size == 1:
            return
        flat = self.engine.flat_grads
        cur = torch.cuda.current_stream()
        if self.slices is None:
            # not prepared: backward recorded no bucket events this step -> one exchange after backward, then install
            # the plan (from any training handle of the model) for the following steps
            dist.all_reduce(flat, group=self.group)
            if self.engine.handles:  # the bucket plan depends on the parameter set only: any handle will do
                self._setup(next(iter(self.engine.handles.values())))
            return
        if not self.overlap:  # one all-reduce of the whole arena after backward (S4G_DDP_OVERLAP=0)
            dist.all_reduce(flat, group=self.group)
            return
        opt = self.fused_optimizer
        with torch.cuda.stream(self.comm_stream):
            if opt is not None:
                opt.begin_external_step()
            allreduce_buckets(flat, self.slices, self.group, self.events, self.comm_stream,
                              None if opt is None else lambda ranges: [opt.step_range(lo, hi) for lo, hi in ranges])
        cur.wait_stream(self.comm_stream)
