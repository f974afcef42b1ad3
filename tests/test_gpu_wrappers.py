"""The MAE- and MoCo-v3-based wrappers on the GPU (reference Models/models.py:360-475, :478-578): a synthetic encoder
checkpoint is loaded the way the reference does (load_my_state_dict / strict load before the decoder exists), then one
fwd + loss + bwd is compared with the oracle on the same tensors.  The wrappers differ from the timm-based one in what is
trainable: pos_embed (both) and decoder_pos_embed (MAE) are fixed sin-cos tables, patch_embed.proj is forced trainable
in the MoCo wrapper (Models/models.py:392-399, 506-511)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import ref_port  # noqa: E402  (checker only)


def _l2(a, b):
    return ((a.double() - b.double()).norm() / (b.double().norm() + 1e-300)).item()


def _rel(a, b):
    return ((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-30)).item()


def _encoder_ckpt(seed):
    sd = ref_port.make_state(None, seed=seed)
    return {k: v for k, v in sd.items()}  # cls_token, pos_embed, patch_embed.proj.*, blocks.*, norm.*


def _check_step(m, dense, B, frozen_names):
    from ssl4gie_b200 import ScaleAndShiftInvariantLoss, SoftDiceLoss
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    g = torch.Generator().manual_seed(0)
    x = torch.randn(B, 3, 224, 224, generator=g).cuda()
    if dense == "seg":
        t = (torch.rand(B, 1, 224, 224, generator=g) > 0.5).float().cuda()
        loss_fn = SoftDiceLoss()
        m.decoder.output_conv[3].p = 0.0
    else:
        t = torch.rand(B, 1, 224, 224, generator=g)
        t[:, :, :8, :] = 0.0
        t = t.cuda()
        loss_fn = ScaleAndShiftInvariantLoss(alpha=0.1)
    m.precision = "fp32"
    m = m.cuda().train()
    before = {n: p.detach().clone() for n, p in m.named_parameters() if n in frozen_names}
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    out = m(x)
    loss = loss_fn(out, t)
    loss.backward()
    o32, l32, g32 = ref_port.loss_and_grads(sd, x, t, dense, None, 0.1)
    sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
    o64, l64, g64 = ref_port.loss_and_grads(sd64, x.double(), t.double(), dense, None, 0.1)
    assert _rel(out.detach(), o64) < max(1e-4, 3 * _rel(o32, o64))
    assert abs(loss.item() - l64.item()) < 1e-4 * abs(l64.item())
    bad = []
    for n, p in m.named_parameters():
        if n in frozen_names:
            assert not p.requires_grad and p.grad is None, n            # fixed tables: never a gradient ...
            assert torch.equal(p.detach(), before[n]), n                # ... and never written
            continue
        if n.startswith("norm.") or n.startswith("decoder.refinenet4.resConfUnit1.") or n == "decoder_pos_embed":
            assert p.grad is None, n
            continue
        assert p.grad is not None, n
        e, e_ref = _l2(p.grad, g64[n]), _l2(g32[n], g64[n])
        if e > max(1e-4, 3 * e_ref):
            bad.append((n, e, e_ref))
    assert not bad, bad[:10]
    return m


def test_mae_wrapper_step(tmp_path, capsys):
    from ssl4gie_b200 import utils
    enc = _encoder_ckpt(11)
    ckpt = dict(enc)
    ckpt["pos_embed"] = torch.randn(1, 197, 768) * 0.02           # a pretraining run saves its own table
    ckpt["decoder_pos_embed"] = torch.randn(1, 197, 512) * 0.02
    ckpt["mask_token"] = torch.zeros(1, 1, 512)                    # MAE-decoder entries: skipped silently (:417-425)
    ckpt["decoder_blocks.0.attn.qkv.weight"] = torch.zeros(1536, 512)
    path = str(tmp_path / "mae.pth")
    torch.save({"model": ckpt}, path)
    m = utils.get_MAE_backbone(path, False, 1, False, "seg", False)
    assert "Successfully loaded params for" in capsys.readouterr().out
    sd = m.state_dict()
    for k in enc:
        assert torch.equal(sd[k], ckpt[k]), k
    assert torch.equal(sd["decoder_pos_embed"], ckpt["decoder_pos_embed"])
    _check_step(m, "seg", 2, {"pos_embed", "decoder_pos_embed"})
    assert m.patch_embed.proj.weight.grad is not None and m.cls_token.grad is not None


def test_moco_wrapper_step(tmp_path):
    from ssl4gie_b200 import utils
    enc = _encoder_ckpt(12)
    path = str(tmp_path / "moco.pth")
    torch.save(enc, path)
    m = utils.get_MoCoV3_backbone(path, "vit_b", False, None, False, "depth", False)
    for k, v in enc.items():
        assert torch.equal(m.state_dict()[k], v), k
    # a checkpoint that also holds head weights must be refused: the reference strict-loads before the decoder exists
    bad = dict(enc)
    bad["head.weight"] = torch.zeros(4096, 768)
    torch.save(bad, str(tmp_path / "bad.pth"))
    with pytest.raises(RuntimeError):
        utils.get_MoCoV3_backbone(str(tmp_path / "bad.pth"), "vit_b", False, None, False, "depth", False)
    m = _check_step(m, "depth", 2, {"pos_embed"})
    # forced trainable (Models/models.py:510-511) and trained
    assert m.patch_embed.proj.weight.requires_grad and m.patch_embed.proj.weight.grad is not None
    assert m.patch_embed.proj.bias.grad is not None
