"""sync-free epoch loops and the reference's checkpoint dict (Binary_segmentation/
train_segmentation.py:25-81, 84-126, 185-206, 328-342) through ssl4gie_b200.train, and FusedAdamW's state_dict in
torch.optim.AdamW's format (what the reference stores and reloads under "optimizer_state_dict")."""
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _model(seed=42, precision="fp32"):
    from ssl4gie_b200 import utils
    torch.manual_seed(seed)
    m = utils.get_ImageNet_or_random_ViT(False, 1, False, "seg", False, False)
    m.decoder.output_conv[3].p = 0.0
    m.precision = precision
    return m.cuda().train()


def _batches(n, B=2, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(B, 3, 224, 224, generator=g), (torch.rand(B, 1, 224, 224, generator=g) > 0.5).float()) for _ in range(n)]


def _max_param_diff(a, b):
    return max((p.detach() - q.detach()).abs().max().item() for p, q in zip(a.parameters(), b.parameters()))


def test_fused_adamw_state_dict_interchanges_with_torch_adamw():
    from ssl4gie_b200 import FusedAdamW, SoftDiceLoss
    loss_fn = SoftDiceLoss()
    data = [(x.cuda(), t.cuda()) for x, t in _batches(4)]

    def step(m, opt, x, t):
        opt.zero_grad()
        loss_fn(m(x), t).backward()
        opt.step()

    # torch.optim.AdamW for two steps (the reference's optimizer, train_segmentation.py:204) ...
    m_t = _model()
    opt_t = torch.optim.AdamW(m_t.parameters(), lr=1e-4)
    for x, t in data[:2]:
        step(m_t, opt_t, x, t)
    sd_t = opt_t.state_dict()
    # ... resumed by FusedAdamW on a copy of the model
    m_f = _model(seed=1)
    m_f.load_state_dict(m_t.state_dict())
    opt_f = FusedAdamW(m_f, lr=3e-4)
    opt_f.load_state_dict(sd_t)
    assert opt_f.step_count == 2 and opt_f.param_groups[0]["lr"] == 1e-4
    step(m_t, opt_t, *data[2])
    step(m_f, opt_f, *data[2])
    assert _max_param_diff(m_t, m_f) < 2e-6
    # and back: FusedAdamW's state_dict loads into a fresh torch.optim.AdamW over model.parameters()
    sd_f = opt_f.state_dict()
    assert sorted(sd_f) == ["param_groups", "state"]
    assert sd_f["param_groups"][0]["params"] == list(range(len(list(m_f.parameters()))))
    assert sorted(sd_f["state"]) == sorted(opt_t.state_dict()["state"])  # same parameters carry state (grad-less ones do not)
    m_t2 = _model(seed=2)
    m_t2.load_state_dict(m_f.state_dict())
    opt_t2 = torch.optim.AdamW(m_t2.parameters(), lr=1e-4)
    opt_t2.load_state_dict(sd_f)
    step(m_t2, opt_t2, *data[3])
    step(m_f, opt_f, *data[3])
    assert _max_param_diff(m_t2, m_f) < 2e-6


def test_train_epoch_checkpoint_and_resume(tmp_path, golden_dir, capsys):
    from ssl4gie_b200 import FusedAdamW, SoftDiceLoss, train
    loader = _batches(4)
    log = str(tmp_path / "log.txt")
    ckpt = str(tmp_path / "ckpt.pth")
    loss_fn = SoftDiceLoss()

    m_a = _model()
    opt_a = FusedAdamW(m_a, lr=1e-4)
    l1 = train.train_epoch(m_a, 0, 1, loader, None, opt_a, 1, loss_fn, log, scaler=None, log_every=2)
    # the same four steps written out by hand
    m_ref = _model()
    opt_ref = FusedAdamW(m_ref, lr=1e-4)
    by_hand = []
    for x, t in loader:
        loss = loss_fn(m_ref(x.cuda()), t.cuda())
        loss.backward()
        opt_ref.step()
        by_hand.append(loss.item())
    assert abs(l1 - sum(by_hand) / 4) < 1e-5 * abs(l1)
    out = capsys.readouterr().out
    assert "Train Epoch: 1 [8/8 (100.0%)]\tAverage loss: " in out and "Train Epoch: 1 [4/8 (50.0%)]\tLoss: " in out
    assert open(log).read().startswith("Train Epoch: 1 [8/8 (100.0%)]\tAverage loss: ")

    perf = train.test(m_a, 0, loader[:2], 1, lambda o, t: ((o > 0).float() == t).float().mean(), log)
    assert 0.0 <= perf <= 1.0
    train.save_checkpoint(ckpt, 1, m_a, opt_a, None, l1, perf, perf)
    l2 = train.train_epoch(m_a, 0, 1, loader, None, opt_a, 2, loss_fn, log)

    # the dict is the reference's, key for key, and the reference's own resume code accepts it
    d = torch.load(ckpt, map_location="cpu", weights_only=False)
    assert sorted(d) == sorted(["epoch", "model_state_dict", "optimizer_state_dict", "scaler_state_dict", "loss", "val_perf",
                                "test_perf", "py_state", "np_state", "torch_state"])
    manifest = json.load(open(os.path.join(golden_dir, "manifest_seg.json")))
    assert sorted(d["model_state_dict"]) == sorted(manifest)
    m_b = _model(seed=3)
    torch.optim.AdamW(m_b.parameters(), lr=1e-4).load_state_dict(d["optimizer_state_dict"])  # train_segmentation.py:206
    # the reference's build() loads "scaler_state_dict" into an ENABLED GradScaler unconditionally (:205-207)
    sc = torch.amp.GradScaler("cuda")
    sc.load_state_dict(d["scaler_state_dict"])
    assert sc.get_scale() == 65536.0

    # resume in a fresh process state: same second epoch
    opt_b = FusedAdamW(m_b, lr=1e-4)
    start, best = train.load_checkpoint(ckpt, m_b, opt_b)
    assert start == 2 and best == perf and opt_b.step_count == 4
    sd_b = opt_b.state_dict()["state"]  # the optimizer state survives the round trip exactly
    for i, st in d["optimizer_state_dict"]["state"].items():
        assert torch.equal(st["exp_avg"], sd_b[i]["exp_avg"].cpu()) and torch.equal(st["exp_avg_sq"], sd_b[i]["exp_avg_sq"].cpu())
    assert all(torch.equal(v.cpu(), d["model_state_dict"][k]) for k, v in m_b.state_dict().items())
    l2b = train.train_epoch(m_b, 0, 1, loader, None, opt_b, 2, loss_fn, log)
    assert abs(l2 - l2b) < 1e-4 * abs(l2), (l2, l2b)
    # parameters: Adam's g / sqrt(v) turns the run-to-run fp32 atomic-order noise of near-zero gradient entries into
    # full-size (lr = 1e-4) steps of either sign, so single entries differ by up to ~lr after an epoch; the bulk agrees
    mean_diff = sum((p.detach() - q.detach()).abs().sum().item() for p, q in zip(m_a.parameters(), m_b.parameters())) / \
        sum(p.numel() for p in m_a.parameters())
    assert mean_diff < 2e-6 and _max_param_diff(m_a, m_b) < 4e-4, (mean_diff, _max_param_diff(m_a, m_b))


def test_epoch_loops_select_the_bf16_schedule_by_default():
    """precision "auto" outside autocast means the fp32 validation kernels for a bare model(x) call, but the drop-in
    epoch loops (the reference runs them under autocast, train_segmentation.py:45) pick the bf16 tcgen05 schedule."""
    from ssl4gie_b200 import FusedAdamW, SoftDiceLoss, train
    from ssl4gie_b200.engine import Engine
    m = _model(precision="auto")
    opt = FusedAdamW(m, lr=1e-4)
    train.train_epoch(m, 0, 1, _batches(2), None, opt, 1, SoftDiceLoss(), None)
    train.test(m, 0, _batches(1), 1, lambda o, t: ((o > 0).float() == t).float().mean(), None)
    keys = list(Engine.of(m).handles)
    assert keys and all(k[1] == "bf16" and k[2] == "tensor" for k in keys), keys
    assert m.precision == "auto"
    m.precision = "fp32"   # explicit choice is respected
    train.test(m, 0, _batches(1), 1, lambda o, t: ((o > 0).float() == t).float().mean(), None)
    assert any(k[1] == "fp32" for k in Engine.of(m).handles)
