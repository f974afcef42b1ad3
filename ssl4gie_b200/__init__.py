"""ssl4gie_b200 — B200-native (sm_100a) ViT-B/16 + DPT fine-tuning hot path of ESandML/SSL4GIE.

Same constructors, nn.Module call signatures and state_dict layout as the reference
(Models/models.py, Models/DPT_decoder.py, utils.py, */Metrics/losses.py); forward/backward run through
hand-written CUDA kernels behind the C ABI in include/ssl4gie_b200.h.  There is no CPU or PyTorch-op fallback:
every compute call raises if libssl4gie_b200.so is missing or the device is not sm_100.
"""
from . import utils  # noqa: F401
from .augment import GpuInputPipeline  # noqa: F401
from .dpt_decoder import DPT_decoder  # noqa: F401
from .graph import GraphedStep  # noqa: F401
from .losses import CrossEntropyLoss, ScaleAndShiftInvariantLoss, SoftDiceLoss  # noqa: F401
from .mae import MaskedAutoencoderViT, add_weight_decay, adjust_learning_rate, mae_vit_base_patch16  # noqa: F401
from .models import ViT_from_MAE, ViT_from_MoCoV3, VisionTransformer_from_Any  # noqa: F401
from .optim import FusedAdamW  # noqa: F401
from .prefetch import DevicePrefetcher  # noqa: F401

__version__ = "0.1.0"
