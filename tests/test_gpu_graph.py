"""CUDA-graph capture of the whole step (ssl4gie_b200.GraphedStep): the replayed step must be the eager
step -- same losses, same parameters -- with the learning rate still following the scheduler and dropout still drawing a new
mask per step, although nothing of that is passed by value any more."""
import time

import pytest
import torch

from oracle import ref_port  # checker only: initial states

pytestmark = pytest.mark.gpu


def _seg_model(p_drop):
    from ssl4gie_b200 import utils
    m = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
    m.load_state_dict(ref_port.make_state("seg", seed=42), strict=True)
    m.decoder.output_conv[3].p = p_drop
    m.precision, m.backend = "bf16", "tensor"
    return m.cuda().train()


def _batch(B, k):
    g = torch.Generator().manual_seed(50 + k)
    return torch.randn(B, 3, 224, 224, generator=g).cuda(), (torch.rand(B, 1, 224, 224, generator=g) > 0.5).float().cuda()


def _lr(k):
    return 1e-4 * (1.0 + 0.3 * k)   # a schedule that changes every step


def test_graphed_step_equals_eager_step():
    from ssl4gie_b200 import FusedAdamW, GraphedStep, SoftDiceLoss
    from ssl4gie_b200.engine import Engine
    steps, B = 7, 4
    loss_fn = SoftDiceLoss()
    m_e = _seg_model(0.0)
    o_e = FusedAdamW(m_e, lr=1e-4)
    eager = []
    for k in range(steps):
        o_e.param_groups[0]["lr"] = _lr(k)
        x, t = _batch(B, k)
        loss = loss_fn(m_e(x), t)
        loss.backward()
        o_e.step()
        eager.append(loss.item())
    m_g = _seg_model(0.0)
    o_g = FusedAdamW(m_g, lr=1e-4)
    step = GraphedStep(m_g, loss_fn, o_g, eager_steps=2)
    graphed, enqueue = [], []
    for k in range(steps):
        o_g.param_groups[0]["lr"] = _lr(k)
        x, t = _batch(B, k)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        loss = step(x, t)
        enqueue.append(time.perf_counter() - t0)
        graphed.append(loss.item())
    assert step.graph is not None and o_g.step_count == steps
    pe, pg = Engine.of(m_e).flat_params, Engine.of(m_g).flat_params
    rel = ((pe - pg).double().norm() / pe.double().norm()).item()
    worst = max(abs(a - b) / abs(a) for a, b in zip(eager, graphed))
    print(f"[graph] eager vs graphed over {steps} steps: worst loss deviation {worst:.2e}, parameters relative L2 {rel:.2e}; "
          f"host time per replay {1e3 * min(enqueue[3:]):.3f} ms (eager step {1e3 * min(enqueue[:2]):.3f} ms)")
    # identical kernels on identical inputs; the only difference is where Adam's bias corrections are evaluated (device powf vs host)
    assert worst < 1e-5 and rel < 1e-6
    sd = o_g.state_dict()
    assert int(sd["state"][0]["step"]) == steps


def test_graphed_step_draws_a_new_dropout_mask_every_replay():
    from ssl4gie_b200 import FusedAdamW, GraphedStep, SoftDiceLoss

    def run():
        torch.manual_seed(123)
        m = _seg_model(0.5)
        opt = FusedAdamW(m, lr=0.0, weight_decay=0.0)    # parameters frozen: only the mask changes between steps
        step = GraphedStep(m, SoftDiceLoss(), opt, eager_steps=1)
        x, t = _batch(2, 0)
        return [step(x, t).item() for _ in range(6)]

    a, b = run(), run()
    assert len(set(a[1:])) == len(a[1:]), a      # replays 2..6: same input, same parameters, different masks
    assert a == b                                 # and the sequence is a function of torch's seed


def test_graphed_mae_step():
    import ssl4gie_b200
    from ssl4gie_b200 import FusedAdamW, GraphedStep
    sd = ref_port.mae_make_state(seed=42)

    def model():
        m = ssl4gie_b200.mae_vit_base_patch16(norm_pix_loss=True)
        m.load_state_dict(sd, strict=True)
        m.precision, m.backend = "bf16", "tensor"
        return m.cuda().train()

    g = torch.Generator().manual_seed(9)
    data = [(torch.randn(4, 3, 224, 224, generator=g).cuda(), torch.rand(4, 196, generator=g).cuda()) for _ in range(5)]
    m_e = model()
    o_e = FusedAdamW(m_e, lr=1e-3, betas=(0.9, 0.95), param_groups=ssl4gie_b200.add_weight_decay(m_e, 0.05))
    eager = []
    for x, n in data:
        loss = m_e(x, noise=n)[0]
        loss.backward()
        o_e.step()
        eager.append(loss.item())
    m_g = model()
    o_g = FusedAdamW(m_g, lr=1e-3, betas=(0.9, 0.95), param_groups=ssl4gie_b200.add_weight_decay(m_g, 0.05))
    step = GraphedStep(m_g, None, o_g, eager_steps=2)
    graphed = [step(x, n).item() for x, n in data]
    worst = max(abs(a - b) / abs(a) for a, b in zip(eager, graphed))
    print(f"[graph] MAE eager vs graphed: worst loss deviation {worst:.2e}")
    assert worst < 1e-5


def test_train_epoch_with_a_graphed_step():
    """train.train_epoch(..., graph_step=GraphedStep(...)) (reference loop train_segmentation.py:25-81): same epoch means as the
    launch-by-launch loop over two epochs."""
    from ssl4gie_b200 import FusedAdamW, GraphedStep, SoftDiceLoss, train
    loader = [tuple(v.cpu() for v in _batch(2, k)) for k in range(4)]
    loss_fn = SoftDiceLoss()
    m_e = _seg_model(0.0)
    o_e = FusedAdamW(m_e, lr=1e-4)
    eager = [train.train_epoch(m_e, 0, 1, loader, None, o_e, ep, loss_fn, None) for ep in (1, 2)]
    m_g = _seg_model(0.0)
    o_g = FusedAdamW(m_g, lr=1e-4)
    gs = GraphedStep(m_g, loss_fn, o_g)
    graphed = [train.train_epoch(m_g, 0, 1, loader, None, o_g, ep, loss_fn, None, graph_step=gs) for ep in (1, 2)]
    assert gs.graph is not None and o_g.step_count == 8
    assert all(abs(a - b) < 1e-6 * abs(a) for a, b in zip(eager, graphed)), (eager, graphed)


def test_graphed_step_with_bucket_events_installed():
    """bench.py builds a GradientAllReduce even on one GPU (its bucket events are then installed in the C schedule); capturing the
    step must not record them into the graph."""
    from ssl4gie_b200 import FusedAdamW, GraphedStep, SoftDiceLoss
    from ssl4gie_b200.ddp import GradientAllReduce
    m = _seg_model(0.1)
    opt = FusedAdamW(m, lr=1e-4)
    sync = GradientAllReduce(m)
    sync.attach(opt)
    sync.prepare(2)
    x, t = _batch(2, 0)
    loss_fn = SoftDiceLoss()
    for _ in range(2):     # eager steps with the events in place, as the bench runs before it builds the graph
        loss_fn(m(x), t).backward()
        sync.reduce()
        opt.step()
    gs = GraphedStep(m, loss_fn, opt, eager_steps=1)
    losses = [gs(x, t).item() for _ in range(4)]
    assert gs.graph is not None and all(l == l for l in losses)
    torch.rand(4, device="cuda")   # torch's generator is usable after the capture
