"""Pinned-host -> device input prefetch on a side stream (the reference's train_epoch copies each batch
with a blocking ``.cuda(rank)`` at the top of the step, Binary_segmentation/train_segmentation.py:41-43; here the copy of
batch i+1 overlaps the compute of batch i).

    pf = DevicePrefetcher(loader)            # loader yields tuples of CPU tensors (pinned for true overlap)
    for x, t in pf:                          # device tensors, ready on the current stream
        loss = loss_fn(model(x), t); loss.backward(); opt.step()

The device tensors live in a ring of ``depth + 1`` staging sets owned by the prefetcher and are REUSED: a yielded batch
is valid until the next batch is requested (consume it on the current stream before calling ``next``; ``.clone()`` what
must outlive the step).  No allocation happens per batch: a cudaMalloc on a caching-allocator miss is a kernel-mode call
that was measured stalling the host for ~40 ms (and the GPU with it) when it collides with an NVML poll.
"""
import torch

_COPY_STREAMS = {}


def _copy_stream(device):
    """One copy stream per device for every prefetcher: the caching allocator pools blocks per stream, so a fresh stream
    per epoch would cudaMalloc its staging buffers again (3 ms of host time at the top of every epoch) and strand the
    old ones."""
    s = _COPY_STREAMS.get(device.index)
    if s is None:
        s = _COPY_STREAMS[device.index] = torch.cuda.Stream(device=device)
    return s


_SLOT_POOL = {}  # (device index, ((shape, dtype), ...)) -> idle slots of finished prefetchers, staging buffers included


def _signature(device, host):
    return (device.index, tuple((tuple(h.shape), h.dtype) for h in host))


class _Slot:
    __slots__ = ("dev", "host", "ready", "free")

    def __init__(self):
        self.dev, self.host, self.ready, self.free = None, None, torch.cuda.Event(), None


class DevicePrefetcher:
    def __init__(self, batches, device=None, depth=2):
        self.batches = iter(batches)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.stream = _copy_stream(self.device)
        self.depth = max(1, depth)
        self.slots = [_Slot() for _ in range(self.depth + 1)]
        self.issued = 0
        self.queue = []
        self.current = None  # the slot handed out by the previous __next__

    def _issue(self):
        try:
            host = next(self.batches)
        except StopIteration:
            return False
        single = not isinstance(host, (tuple, list))
        if single:
            host = (host,)
        slot = self.slots[self.issued % len(self.slots)]
        self.issued += 1
        with torch.cuda.stream(self.stream):
            fits = slot.dev is not None and len(slot.dev) == len(host) and all(
                d.shape == h.shape and d.dtype == h.dtype for d, h in zip(slot.dev, host))
            if not fits:  # first use of the slot, or a ragged last batch
                idle = _SLOT_POOL.get(_signature(self.device, host))
                if idle:  # a finished prefetcher's slot: same buffers, its `free` event orders the overwrite
                    old = idle.pop()
                    slot.dev, slot.free, slot.ready = old.dev, old.free, old.ready
                else:
                    slot.dev = tuple(torch.empty(h.shape, dtype=h.dtype, device=self.device) for h in host)
                    slot.free = None
            if slot.free is not None:
                self.stream.wait_event(slot.free)  # the step that read this slot's previous batch has been enqueued before it
            for d, h in zip(slot.dev, host):
                d.copy_(h, non_blocking=True)
            slot.ready.record(self.stream)
        slot.host = host  # stays referenced until the slot is overwritten
        self.queue.append((slot, single))
        return True

    def __iter__(self):
        return self

    def __next__(self):
        cur = torch.cuda.current_stream(self.device)
        if self.current is not None:
            # everything that reads the previous batch is on `cur` by now: its slot may be overwritten after this point
            if self.current.free is None:
                self.current.free = torch.cuda.Event()
            self.current.free.record(cur)
            self.current = None
        while len(self.queue) < self.depth and self._issue():
            pass
        if not self.queue:
            self._retire()
            raise StopIteration
        slot, single = self.queue.pop(0)
        cur.wait_event(slot.ready)
        for d in slot.dev:
            d.record_stream(cur)  # when the ring is dropped, the allocator must not recycle a buffer the last steps still read
        self.current = slot
        self._issue()
        return slot.dev[0] if single else slot.dev

    def _retire(self):
        """Loader exhausted: every slot's last reader is ordered by its `free` event, so the staging buffers can serve the
        next prefetcher (next epoch) without touching the allocator."""
        for slot in self.slots:
            if slot.dev is not None and slot.free is not None:
                slot.host = None
                idle = _SLOT_POOL.setdefault(_signature(self.device, slot.dev), [])
                if len(idle) < 8:
                    idle.append(slot)
        self.slots = [_Slot() for _ in range(self.depth + 1)]
