"""stofnet_b200 - B200-native (sm_100a) StofNet echo-waveform inference path.

Drop-in for the hot path of hahnec/stofnet: ``models.stofnet.StofNet`` forward, the sub-pixel
shuffle, and ``utils.mask2samples`` time-of-flight extraction, behind the reference's own Python
surface (same module / class / function names and signatures).  All compute runs in hand-written
CUDA kernels reached through the C ABI of ``libstofnet_b200.so`` (include/stofnet_b200.h).
"""
from .models.stofnet import StofNet, SemiGlobalBlock          # noqa: F401
from .models.espcn_1d import ESPCN_1D                          # noqa: F401
from .models.edsr_1d import EDSR_1D                            # noqa: F401
from .utils.sample_shuffle import SampleShuffle1D              # noqa: F401
from .utils.mask2samples import (mask2coords, get_maxima_positions, nms_1d, thresholding,  # noqa: F401
                                 coords2mask)
from .utils.metrics import toa_rmse                             # noqa: F401
from .utils.threshold import find_threshold                     # noqa: F401
from .utils.attention_block import AttentionBlock               # noqa: F401
from .utils.gaussian import gaussian_kernel                     # noqa: F401
from .datasets.chirp_dataset import upscale_1d, iq2rf           # noqa: F401
from .loss import mask_loss                                     # noqa: F401
from .train import Trainer, cosine_lr, average_gradients          # noqa: F401
from .checkpoint import load_checkpoint, load_config, build_model  # noqa: F401

__version__ = "0.1"
