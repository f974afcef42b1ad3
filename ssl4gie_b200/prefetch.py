"""Pinned-host -> device input prefetch on a side stream (the reference's train_epoch copies each batch
with a blocking ``.cuda(rank)`` at the top of the step, Binary_segmentation/train_segmentation.py:41-43; here the copy of
batch i+1 overlaps the compute of batch i).

    pf = DevicePrefetcher(loader)            # loader yields tuples of CPU tensors (pinned for true overlap)
    for x, t in pf:                          # device tensors, ready on the current stream
        loss = loss_fn(model(x), t); loss.backward(); opt.step()
"""
import torch


class DevicePrefetcher:
    def __init__(self, batches, device=None, depth=2):
        self.batches = iter(batches)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.depth = max(1, depth)
        self.queue = []

    def _issue(self):
        try:
            host = next(self.batches)
        except StopIteration:
            return False
        if not isinstance(host, (tuple, list)):
            host = (host,)
        with torch.cuda.stream(self.stream):
            dev = tuple(h.to(self.device, non_blocking=True) for h in host)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.queue.append((dev, ev, host))  # `host` stays referenced until its copy has been consumed
        return True

    def __iter__(self):
        return self

    def __next__(self):
        while len(self.queue) < self.depth and self._issue():
            pass
        if not self.queue:
            raise StopIteration
        dev, ev, _ = self.queue.pop(0)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        for d in dev:
            d.record_stream(cur)  # the caching allocator must not recycle the buffer while the step still reads it
        self._issue()
        return dev if len(dev) > 1 else dev[0]
