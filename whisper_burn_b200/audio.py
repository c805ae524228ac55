"""Mirror of src/audio/mel.rs + the framing of src/transcribe.rs:46-82 over the C ABI (wq4_logmel_*): batched log-mel
spectrograms computed on the GPU.  torch tensors are device buffers only."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .gguf import _as_device, _check_f32_cuda, _stream

WHISPER_SAMPLE_RATE = 16000        # src/transcribe.rs:16-20
WHISPER_CHUNK_SAMPLES = 480000
WHISPER_MEL_FRAMES = 3000


class MelSpectrogram:
    """MelSpectrogram::new(MelConfig::whisper_with_mels(n_mels)) (mel.rs:45-90)."""

    def __init__(self, n_mels: int = 128, device="cuda"):
        self.device = _as_device(device)
        self.n_mels = int(n_mels)
        self._h = C.c_void_p()
        _lib.check(_lib.lib().wq4_logmel_create(self.n_mels, self.device.index or 0, C.c_void_p(_stream(self.device)),
                                                C.byref(self._h)))

    def compute_log(self, samples: torch.Tensor) -> torch.Tensor:
        """samples [B, n] (or [n]) f32 16 kHz mono on the device -> model input [B, n_mels, 3000] f32: compute_log
        (mel.rs:118-157) of the clip padded / truncated to 30 s, transposed (transcribe.rs:46-82)."""
        squeeze = samples.dim() == 1
        if squeeze:
            samples = samples.unsqueeze(0)
        samples = _check_f32_cuda(samples, "samples")
        if samples.dim() != 2 or samples.shape[1] < 1:
            raise _lib.Wq4Error(2, "samples must be [B, n] with n >= 1")
        b, n = samples.shape
        out = torch.empty((b, self.n_mels, WHISPER_MEL_FRAMES), dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().wq4_logmel_forward(self._h, C.c_void_p(samples.data_ptr()), b, n, C.c_void_p(out.data_ptr()),
                                                 C.c_void_p(_stream(self.device))))
        return out[0] if squeeze else out

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None and h.value:
            try:
                _lib.lib().wq4_logmel_destroy(h)
            except Exception:
                pass
