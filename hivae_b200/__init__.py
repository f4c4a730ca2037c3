"""hivae_b200 — B200-native (sm_100a) implementation of the Hi-VAE hot path.

Python host code mirrors the reference's nn.Module API (model/model_AMD.py, model/transformer.py,
model/modules.py, model/wavelet.py, model/frequency_utils.py, model/rectified_flow.py,
model/attention.py); all device work is done by the hand-written CUDA kernels in csrc/ behind the
C ABI of include/hivae_b200.h (bound with ctypes in _lib.py).  No CPU fallback.
"""
from . import configs, synth  # noqa: F401
from ._lib import LIB_PATH, load as load_library  # noqa: F401
from .wavelet import DWT, IWT, dwt_init, iwt_init, dwt_pyramid, iwt_pyramid  # noqa: F401
from .frequency_utils import gaussian_low_pass_filter, freq_3d_filter  # noqa: F401
from .rectified_flow import RectifiedFlow  # noqa: F401
from .modules import (BasicTransformerBlock, BasicCrossTransformerBlock, AMDTransformerBlock, BasicDiTBlock,  # noqa: F401
                      PatchEmbed, AMDLayerNormZero, AMDLayerNormZero_OneVariable, AMDLayerNormZero2Condition, AdaLayerNorm,
                      AMDTransformerMotionBlock)
from .transformer import (MotionEncoderLearnTokenTransformer, MotionEncoderLearnTokenOnlyTemporalTransformer,  # noqa: F401
                          MotionEncoderLearnTokenTemporalTransformer,
                          AMDDiffusionTransformerModelImgSpatialTempMotion, AMDDiffusionTransformerModelTempMotion,
                          AMDDiffusionTransformerModelImgSpatial, AMDDiffusionTransformerModel, AMDReconstructTransformerModel)
from .model_AMD import AMDModel, AMDModel_Camera, AMDModel_New, AMDModel_Rec, AMD_N, AMD_models  # noqa: F401
from . import sharding  # noqa: F401
from . import optim  # noqa: F401
from . import utils  # noqa: F401
from .utils import find_latest_checkpoint, model_load_pretrain, frozen_model  # noqa: F401
from .pipeline import AMDPipeLine_single  # noqa: F401

__version__ = "0.1.0"
