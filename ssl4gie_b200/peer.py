"""Peer-memory mailboxes for the SyncBatchNorm statistic exchange (csrc/peer.cu; replaces the NCCL collectives inside
torch.nn.SyncBatchNorm, reference Binary_segmentation/train_segmentation.py:199): every rank cudaMallocs one mailbox,
the 64-byte cudaIpc handles travel through the process group, peers are opened with lazy peer access.  One mailbox set
per process serves every model handle (`shared_mailboxes`)."""
import ctypes
import os

import torch

from . import _C


class PeerMailboxes:
    """`ptrs` = one device pointer per rank (own mailbox at [rank]); collective over `group`: construct on every rank."""

    def __init__(self, group=None):
        import torch.distributed as dist
        lib = _C.lib()
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.own = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        _C.check(lib.s4g_peer_alloc(lib.s4g_peer_mailbox_bytes(), ctypes.byref(self.own), handle))
        dev = torch.device("cuda", torch.cuda.current_device())
        mine = torch.frombuffer(bytearray(handle.raw), dtype=torch.uint8).to(dev)
        every = torch.empty(self.world * 64, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(every, mine, group=group)
        every = every.cpu()
        self.opened = []
        self.ptrs = (ctypes.c_void_p * self.world)()
        ok = 1
        for r in range(self.world):
            if r == self.rank:
                self.ptrs[r] = self.own.value
                continue
            p = ctypes.c_void_p()
            if lib.s4g_peer_open(bytes(every[r * 64:(r + 1) * 64].tolist()), ctypes.byref(p)) != 0:
                ok = 0
                self.error = lib.s4g_last_error().decode()
                break
            self.opened.append(p)
            self.ptrs[r] = p.value
        flag = torch.tensor([ok], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)  # also the barrier: every mailbox is zeroed and mapped
        self.ok = bool(flag.item())
        if not self.ok:
            self.close()

    def close(self):
        lib = _C.lib()
        for p in self.opened:
            lib.s4g_peer_close(p)
        self.opened = []
        if self.own:
            lib.s4g_peer_free(self.own)
            self.own = ctypes.c_void_p()


    def check(self):
        """Synchronises the device; raises if an exchange timed out waiting for a peer (S4G_PEER_TIMEOUT_S)."""
        if self.own:
            _C.check(_C.lib().s4g_peer_check(self.own))


_shared = None


def shared_mailboxes():
    """The process-wide mailbox set of the default group (every model handle shares it: one exchange counter, kept in
    the mailbox itself).  The FIRST call is collective; later calls are not."""
    global _shared
    if _shared is None:
        _shared = PeerMailboxes()
    return _shared


def use_peer_exchange():
    """S4G_SYNCBN=nccl keeps the host-callback all-reduce (A/B measurements, or boxes without P2P)."""
    return os.environ.get("S4G_SYNCBN", "peer") != "nccl"
