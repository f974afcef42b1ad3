"""Model factories with the reference's exact signatures (reference utils.py:4-82)."""
from . import models


def get_BarlowTwins_backbone(weight_path, head, num_classes, frozen, dense):
    raise NotImplementedError("ResNet50 backbones (utils.py:4-5, :53-56) are outside the ViT-B/16 + DPT hot path")


def get_MAE_backbone(weight_path, head, num_classes, frozen, dense, det, fixed_size=None, out_token="cls"):
    return models.ViT_from_MAE(weight_path, head, num_classes, frozen, dense, det, fixed_size, 768, 12, 12, out_token)


def get_MoCoV3_backbone(weight_path, arch, head, num_classes, frozen, dense, det, fixed_size=None, out_token="cls"):
    if arch == "vit_b":
        return models.ViT_from_MoCoV3(weight_path, head, num_classes, frozen, dense, det, fixed_size, embed_dim=768,
                                      out_token=out_token)
    raise NotImplementedError("arch='resnet50' is outside the ViT-B/16 + DPT hot path")


def get_ImageNet_or_random_ResNet(head, num_classes, frozen, dense, ImageNet_weights):
    raise NotImplementedError("ResNet50 backbones are outside the ViT-B/16 + DPT hot path")


def get_ImageNet_or_random_ViT(head, num_classes, frozen, dense, det, ImageNet_weights, fixed_size=None, out_token="cls"):
    return models.VisionTransformer_from_Any(head, num_classes, frozen, dense, det, fixed_size, 768, 12, 12, out_token,
                                             ImageNet_weights)
