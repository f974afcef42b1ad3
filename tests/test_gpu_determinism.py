"""The product path is bit-reproducible: every cross-block reduction is two-stage with a fixed order (csrc/reduce.cu),
the split-K weight gradients are reduced in a fixed order, and nothing uses floating-point atomics.  torch's LayerNorm /
BatchNorm backward and the reference's losses are deterministic on a fixed device too, so a fixed seed reproduces a
reference run (Binary_segmentation/train_segmentation.py:386-388 seeds everything)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_port  # noqa: E402  (checker only: supplies reference-named initial tensors)


def _run(task, B, steps):
    from ssl4gie_b200 import CrossEntropyLoss, FusedAdamW, ScaleAndShiftInvariantLoss, SoftDiceLoss, utils
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, 3, 224, 224, generator=g).cuda()
    if task == "cls":
        sd = ref_port.make_state(None, seed=42, head_classes=23)
        m = utils.get_ImageNet_or_random_ViT(True, 23, False, None, False, False)
        t = torch.randint(0, 23, (B,), generator=g).cuda()
        loss_fn = CrossEntropyLoss(torch.ones(23)).cuda()
    else:
        sd = ref_port.make_state(task, seed=42)
        m = utils.get_ImageNet_or_random_ViT(False, 1 if task == "seg" else None, False, task, False, False)
        t = torch.rand(B, 1, 224, 224, generator=g)
        t = ((t > 0.5).float() if task == "seg" else t * (t > 0.1)).cuda()
        loss_fn = SoftDiceLoss() if task == "seg" else ScaleAndShiftInvariantLoss(alpha=0.1)
    m.load_state_dict(sd, strict=True)
    m.precision, m.backend = "bf16", "tensor"
    m = m.cuda().train()
    torch.manual_seed(7)  # the seg head's Dropout(0.1) draws its mask from torch's seed + a per-model counter
    opt = FusedAdamW(m, lr=1e-4)
    losses, first_grads = [], None
    for i in range(steps):
        out = m(x)
        loss = loss_fn(out, t)
        loss.backward()
        if i == 0:
            first_grads = {n: p.grad.clone() for n, p in m.named_parameters() if p.grad is not None}
            first_out = out.detach().clone()
        opt.step()
        losses.append(loss.detach().clone())
    return first_out, first_grads, torch.stack(losses), {k: v.clone() for k, v in m.state_dict().items()}


@pytest.mark.parametrize("task,B", [("seg", 3), ("depth", 2), ("cls", 4)])
def test_two_runs_of_the_bf16_step_are_bit_identical(task, B):
    a = _run(task, B, 3)
    b = _run(task, B, 3)
    assert torch.equal(a[0], b[0]), "forward output differs between runs"
    diff = [n for n in a[1] if not torch.equal(a[1][n], b[1][n])]
    assert not diff, ("gradients differ between two runs", diff[:8], len(diff))
    assert torch.equal(a[2], b[2]), (a[2].tolist(), b[2].tolist())
    diff = [k for k in a[3] if not torch.equal(a[3][k], b[3][k])]
    assert not diff, ("state after 3 AdamW steps differs between two runs", diff[:8], len(diff))
