"""whisper_burn_b200 -- B200 (sm_100a) implementation of whisper-burn's Q4_0 hot path.

The package holds only what the path needs: csrc/ (CUDA kernels + the C ABI of include/wq4.h,
built into libwq4.so) and the host-side mirror of the reference's operator interface
(gguf.py: Q4Tensor / q4_matmul / Q4Linear, layers.py: Q4FFN / Conv1D).
"""
from ._lib import Wq4Error, build, lib, launch_count, reset_launch_count  # noqa: F401
from ._lib import WQ4_EPI_GELU, WQ4_FORCE_GEMM, WQ4_FORCE_GEMV, WQ4_FORCE_SKINNY, WQ4_OPERANDS_BF16  # noqa: F401
from .gguf import Q4Linear, Q4Tensor, q4_matmul  # noqa: F401
from .layers import Conv1D, Q4FFN  # noqa: F401
from .model import InferenceState, WhisperConfig, WhisperModel, lang_token, load_whisper_from_gguf, transcribe, transcribe_mels  # noqa: F401
from .audio import MelSpectrogram  # noqa: F401
