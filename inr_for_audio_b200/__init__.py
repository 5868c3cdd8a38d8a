"""inr_for_audio_b200 -- B200-native (sm_100a) INR-fitting hot path behind the reference's models.py / run.py surface.

Host code is Python/PyTorch (device memory, streams, torch.distributed); all arithmetic on the path runs in
hand-written CUDA kernels reached through the C ABI of libinr_b200.so (include/inr_b200.h).  There is no CPU
fallback: importing the model classes works anywhere, calling them without the built library or a GPU raises.
"""
from .models import SineLayer, Siren, SirenWithSnakeTanh, Snake  # noqa: F401
from .encoding import GaussianEncoding, PosEncodingNeRF  # noqa: F401
from .utils import get_coord, calculate_snr, WaveformFitting, MDCTFitting  # noqa: F401
from .engine import ClipBatchFitter, SirenFitter, SirenSpec  # noqa: F401
from .kan import KAN, KANLinear, KanFitter, KanSpec  # noqa: F401
from .run import train  # noqa: F401

__all__ = ["SineLayer", "Siren", "SirenWithSnakeTanh", "Snake", "GaussianEncoding", "PosEncodingNeRF", "get_coord", "calculate_snr", "WaveformFitting",
           "MDCTFitting", "SirenFitter", "SirenSpec", "ClipBatchFitter", "KAN", "KANLinear", "KanFitter", "KanSpec", "train"]
