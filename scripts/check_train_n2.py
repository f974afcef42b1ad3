"""torchrun --nproc-per-node 2 scripts/check_train_n2.py
ssl4gie_b200.train.train_epoch under data parallelism (SyncBatchNorm peer exchange + GradientAllReduce + FusedAdamW with
grad_scale 1/world): every rank returns the same rank-averaged epoch loss, replicas stay bit-identical, the loss falls."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    from ssl4gie_b200 import FusedAdamW, SoftDiceLoss, train, utils
    from ssl4gie_b200.ddp import GradientAllReduce
    torch.manual_seed(42)
    m = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
    m.decoder.output_conv[3].p = 0.0
    # precision stays "auto": train_epoch / test select the bf16 tensor-core schedule themselves; the optimizer's
    # 1/world and the bucket plan are installed by train_epoch / the first reduce() (no prepare(), no grad_scale here)
    m = torch.nn.SyncBatchNorm.convert_sync_batchnorm(m.cuda(local)).train()
    opt = FusedAdamW(m, lr=1e-4)
    sync = GradientAllReduce(m)
    g = torch.Generator().manual_seed(100 + rank)
    loader = [(torch.randn(4, 3, 224, 224, generator=g).pin_memory(), (torch.rand(4, 1, 224, 224, generator=g) > 0.5).float().pin_memory())
              for _ in range(3)]
    losses = [train.train_epoch(m, local, world, loader, None, opt, ep, SoftDiceLoss(), None, grad_sync=sync) for ep in (1, 2)]
    assert opt.grad_scale == 1.0 / world and sync.slices is not None
    # The reference evaluates on rank 0 only while the other ranks wait in a barrier (train_segmentation.py:293-301), with
    # a ragged last batch (no drop_last): the evaluation handles (new batch sizes, eval mode) must be built without any
    # collective, or rank 0's handle construction would pair with the other ranks' barrier.
    if rank == 0:
        gv = torch.Generator().manual_seed(5)
        val = [(torch.randn(b, 3, 224, 224, generator=gv).pin_memory(), (torch.rand(b, 1, 224, 224, generator=gv) > 0.5).float().pin_memory())
               for b in (4, 3)]
        perf = train.test(m, local, val, 1, lambda o, t: ((o > 0).float() == t).float().mean(), None)
        assert 0.0 <= perf <= 1.0
    dist.barrier()
    losses += [train.train_epoch(m, local, world, loader, None, opt, ep, SoftDiceLoss(), None, grad_sync=sync) for ep in (3, 4)]
    from ssl4gie_b200.peer import shared_mailboxes
    shared_mailboxes().check()   # no exchange timed out
    t = torch.tensor(losses, device="cuda", dtype=torch.float64)
    both = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(both, t)
    assert all(torch.equal(b, both[0]) for b in both), both            # the epoch loss is the rank average on every rank
    flat = torch.cat([p.detach().flatten() for p in m.parameters()])
    digest = torch.stack([flat.double().sum(), flat.double().abs().sum()])
    every = [torch.empty_like(digest) for _ in range(world)]
    dist.all_gather(every, digest)
    assert all(torch.equal(e, every[0]) for e in every), every          # replicas did not drift apart
    assert all(l == l for l in losses) and losses[-1] < losses[0], losses
    if rank == 0:
        print("epoch losses", ["%.5f" % l for l in losses], flush=True)
        print("CHECK_TRAIN_N2 OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
