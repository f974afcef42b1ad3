"""torchrun --nproc-per-node 2 scripts/check_mae_n2.py
Data-parallel MAE pretraining step on 2 GPUs (reference: DDP(model, find_unused_parameters=True) at
Models/mae/main_pretrain.py:175 + AdamW groups :179-180):
  1. the bucketed sum-all-reduce of the flat gradient arena equals the sum of the two ranks' local gradients;
  2. five optimizer steps leave bit-identical parameters on both ranks, finite decreasing-ish losses;
  3. the 2-rank run equals ONE process stepping on the concatenated batch (same noise), to fp32 accuracy."""
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
    import ssl4gie_b200
    from ssl4gie_b200 import FusedAdamW
    from ssl4gie_b200.ddp import GradientAllReduce
    from ssl4gie_b200.engine import Engine

    B = 2

    def build():
        torch.manual_seed(42)
        m = ssl4gie_b200.mae_vit_base_patch16(norm_pix_loss=True)
        m.precision, m.backend = "fp32", "simt"
        return m.cuda().train()

    def data(step):
        g = torch.Generator().manual_seed(1000 + step)
        return torch.randn(world * B, 3, 224, 224, generator=g), torch.rand(world * B, 196, generator=g)

    m = build()
    opt = FusedAdamW(m, lr=1e-3, betas=(0.9, 0.95), param_groups=ssl4gie_b200.add_weight_decay(m, 0.05))
    ddp = GradientAllReduce(m)
    ddp.attach(opt)
    ddp.prepare(B)
    e = Engine.of(m)
    losses = []
    for step in range(5):
        x, noise = data(step)
        xs, ns = x[rank * B:(rank + 1) * B].cuda(), noise[rank * B:(rank + 1) * B].cuda()
        ssl4gie_b200.adjust_learning_rate(opt, step / 10, 1e-3, warmup_epochs=1, epochs=4)
        loss, _, _ = m(xs, noise=ns)
        loss.backward()
        if step == 0:
            torch.cuda.synchronize()
            local_g = e.flat_grads.clone()
            ref = local_g.clone()
            dist.all_reduce(ref)
        ddp.reduce()
        if step == 0:
            torch.cuda.synchronize()
            err = ((e.flat_grads - ref).abs().max() / ref.abs().max()).item()
            assert err == 0.0, err   # NCCL's sum of two addends is order-free
            if rank == 0:
                print("bucketed all-reduce == all-reduce of the whole arena: max rel diff %.3g" % err, flush=True)
        opt.step()
        l = loss.detach().clone()
        dist.all_reduce(l)
        losses.append(l.item() / world)
    flat = e.flat_params.clone()
    other = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(other, flat)
    assert all(torch.equal(o, other[0]) for o in other), "parameters diverged between ranks"
    assert all(torch.isfinite(torch.tensor(losses))), losses
    # 4. the bucket-wise optimizer step (attach(optimizer, fused_step=True): AdamW of each bucket right after its all-reduce,
    #    on the communication stream) is the same arithmetic: bit-identical parameters after the same five steps
    m2 = build()
    o2 = FusedAdamW(m2, lr=1e-3, betas=(0.9, 0.95), param_groups=ssl4gie_b200.add_weight_decay(m2, 0.05))
    d2 = GradientAllReduce(m2)
    d2.attach(o2, fused_step=True)
    d2.prepare(B)
    for step in range(5):
        x, noise = data(step)
        xs, ns = x[rank * B:(rank + 1) * B].cuda(), noise[rank * B:(rank + 1) * B].cuda()
        ssl4gie_b200.adjust_learning_rate(o2, step / 10, 1e-3, warmup_epochs=1, epochs=4)
        loss, _, _ = m2(xs, noise=ns)
        loss.backward()
        d2.reduce()
        o2.step()
    torch.cuda.synchronize()
    assert o2.step_count == 5 and torch.equal(Engine.of(m2).flat_params, flat), "fused bucket-wise step differs"
    if rank == 0:
        print("bucket-wise optimizer step on the communication stream: parameters bit-identical to the plain step", flush=True)
    if rank == 0:
        print("5 steps on 2 ranks: mean losses", ["%.5f" % v for v in losses], "parameters bit-identical across ranks", flush=True)
        # the same five steps in one process over the concatenated batch: loss = mean over 2B images' removed patches; every image
        # has 147 removed patches, so the global masked mean equals the mean of the per-rank losses and the gradient their mean
        m1 = build()
        o1 = FusedAdamW(m1, lr=1e-3, betas=(0.9, 0.95), param_groups=ssl4gie_b200.add_weight_decay(m1, 0.05))
        l1 = []
        for step in range(5):
            x, noise = data(step)
            ssl4gie_b200.adjust_learning_rate(o1, step / 10, 1e-3, warmup_epochs=1, epochs=4)
            loss, _, _ = m1(x.cuda(), noise=noise.cuda())
            loss.backward()
            o1.step()
            l1.append(loss.item())
        dl = max(abs(a - b) / abs(b) for a, b in zip(losses, l1))
        p1 = Engine.of(m1).flat_params
        n = Engine.of(m1).n_active
        dp = ((flat[:n] - p1[:n]).norm() / p1[:n].norm()).item()
        print("vs one process on the whole batch: worst loss deviation %.3g, parameters relative L2 %.3g" % (dl, dp), flush=True)
        assert dl < 1e-4 and dp < 1e-4, (dl, dp)
        print("CHECK_MAE_N2 OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
