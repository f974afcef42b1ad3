"""DPT decoder parameter container — same constructor, sub-module names and state_dict keys as
reference Models/DPT_decoder.py:315-499.  The arithmetic lives in csrc/engine.cu (build_decoder)."""
import torch.nn as nn


class ResidualConvUnit_custom(nn.Module):
    """reference Models/DPT_decoder.py:170-210"""

    def __init__(self, features, activation, bn):
        super().__init__()
        self.bn = bn
        self.groups = 1
        self.conv1 = nn.Conv2d(features, features, kernel_size=3, stride=1, padding=1, bias=not self.bn, groups=1)
        self.conv2 = nn.Conv2d(features, features, kernel_size=3, stride=1, padding=1, bias=not self.bn, groups=1)
        if self.bn:
            self.bn1 = nn.BatchNorm2d(features)
            self.bn2 = nn.BatchNorm2d(features)
        self.activation = activation


class FeatureFusionBlock_custom(nn.Module):
    """reference Models/DPT_decoder.py:238-279"""

    def __init__(self, features, activation, deconv=False, bn=False, expand=False, align_corners=True):
        super().__init__()
        self.deconv = deconv
        self.align_corners = align_corners
        self.groups = 1
        self.expand = expand
        out_features = features // 2 if expand else features
        self.out_conv = nn.Conv2d(features, out_features, kernel_size=1, stride=1, padding=0, bias=True, groups=1)
        self.resConfUnit1 = ResidualConvUnit_custom(features, activation, bn)
        self.resConfUnit2 = ResidualConvUnit_custom(features, activation, bn)


def _make_fusion_block(features, use_bn):
    return FeatureFusionBlock_custom(features, nn.ReLU(False), deconv=False, bn=use_bn, expand=False, align_corners=True)


class _Marker(nn.Module):
    """Parameter-free stage kept so Sequential indices (output_conv.0/.1/.4 ...) match the reference."""

    def __init__(self, what, **kw):
        super().__init__()
        self.what = what
        for k, v in kw.items():
            setattr(self, k, v)

    def extra_repr(self):
        return self.what


class DPT_decoder(nn.Module):
    def __init__(self, num_classes=1, dense="seg", vit_features=768, features=[96, 192, 384, 768], fusion_features=256,
                 use_readout="ignore", size=[224, 224], patch_size=[16, 16]):
        super().__init__()
        assert use_readout in ("ignore", "add", "project"), \
            "wrong operation for readout token, use_readout can be 'ignore', 'add', or 'project'"
        if (use_readout != "ignore" or vit_features != 768 or list(features) != [96, 192, 384, 768] or fusion_features != 256
                or list(size) != [224, 224] or list(patch_size) != [16, 16]):
            raise NotImplementedError("the B200 schedule is specialised for the reference's default DPT configuration")
        if dense == "seg" and num_classes != 1:
            raise NotImplementedError("binary segmentation head (num_classes=1) only, as in train_segmentation.py:157-176")
        f = features
        self.act_postprocess12 = nn.Sequential(nn.Conv2d(vit_features, f[0], 1), nn.ConvTranspose2d(f[0], f[0], 4, stride=4))
        self.act_postprocess22 = nn.Sequential(nn.Conv2d(vit_features, f[1], 1), nn.ConvTranspose2d(f[1], f[1], 2, stride=2))
        self.act_postprocess32 = nn.Sequential(nn.Conv2d(vit_features, f[2], 1))
        self.act_postprocess42 = nn.Sequential(nn.Conv2d(vit_features, f[3], 1), nn.Conv2d(f[3], f[3], 3, stride=2, padding=1))
        self.layer1_rn = nn.Conv2d(f[0], fusion_features, 3, padding=1, bias=False)
        self.layer2_rn = nn.Conv2d(f[1], fusion_features, 3, padding=1, bias=False)
        self.layer3_rn = nn.Conv2d(f[2], fusion_features, 3, padding=1, bias=False)
        self.layer4_rn = nn.Conv2d(f[3], fusion_features, 3, padding=1, bias=False)
        use_bn = dense == "seg"
        self.refinenet1 = _make_fusion_block(fusion_features, use_bn)
        self.refinenet2 = _make_fusion_block(fusion_features, use_bn)
        self.refinenet3 = _make_fusion_block(fusion_features, use_bn)
        self.refinenet4 = _make_fusion_block(fusion_features, use_bn)
        if dense == "depth":
            self.output_conv = nn.Sequential(
                nn.Conv2d(fusion_features, fusion_features // 2, 3, padding=1),
                _Marker("Interpolate(scale_factor=2, bilinear, align_corners=True)"),
                nn.Conv2d(fusion_features // 2, 32, 3, padding=1),
                _Marker("ReLU"),
                nn.Conv2d(32, 1, 1),
                _Marker("Sigmoid"),
            )
        else:
            self.output_conv = nn.Sequential(
                nn.Conv2d(fusion_features, fusion_features, 3, padding=1, bias=False),
                nn.BatchNorm2d(fusion_features),
                _Marker("ReLU"),
                nn.Dropout(0.1, False),
                nn.Conv2d(fusion_features, num_classes, 1),
                _Marker("Interpolate(scale_factor=2, bilinear, align_corners=True)"),
            )
        self.dense = dense

    def forward(self, activations):
        raise RuntimeError("DPT_decoder runs fused inside the backbone schedule; call the backbone module instead")
