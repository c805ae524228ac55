"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): NumPy restatement of the reference's log-mel front end.

Follows src/audio/mel.rs (zerr0o/whisper-burn) step by step, in float32 like the Rust code:
  hann_window            mel.rs:315-319   periodic Hann, 0.5*(1 - cos(2*pi*i/N))
  create_mel_filterbank  mel.rs:270-312   Slaney mel scale (hz_to_mel / mel_to_hz, mel.rs:242-267), triangular,
                                          no area normalisation, closed intervals as written there
  stft                   mel.rs:174-227   reflect padding of n_fft/2 as written (index clamps included), frames
                                          i*hop .. i*hop+n_fft, n_frames = (padded - n_fft) / hop, bins 0..n_fft/2
  compute / compute_log  mel.rs:105-157   |X|^2 -> filterbank -> log10(max(.,1e-10)) -> max(., global_max - 8) -> (.+4)/4
  frame_clip             src/transcribe.rs:46-82   pad / truncate to 480 000 samples, transpose to [n_mels, frames],
                                          pad / truncate frames to 3000
Parity status: UNPINNED by the reference (no test or golden vector covers mel.rs; the FFT is rustfft 6.4.1, a
third-party crate, Cargo.lock).  The restatement uses numpy's pocketfft in single precision; both are exact DFTs up to
f32 rounding, and the log / clamp stage is insensitive to that (the clamp floor is 8 decades below the maximum).
"""
from __future__ import annotations

import numpy as np

SAMPLE_RATE = 16000
N_FFT = 400
HOP = 160
CHUNK_SAMPLES = 480000      # 30 s
MEL_FRAMES = 3000

_F32 = np.float32


def hann_window(length: int = N_FFT) -> np.ndarray:
    i = np.arange(length, dtype=_F32)
    return (_F32(0.5) * (_F32(1.0) - np.cos(_F32(2.0) * _F32(np.pi) * i / _F32(length)))).astype(_F32)


def _hz_to_mel(f):
    f_sp = _F32(200.0 / 3.0)
    min_log_hz = _F32(1000.0)
    min_log_mel = min_log_hz / f_sp
    logstep = _F32(0.06875174)
    f = _F32(f)
    return f / f_sp if f < min_log_hz else min_log_mel + np.log(f / min_log_hz, dtype=_F32) / logstep


def _mel_to_hz(m):
    f_sp = _F32(200.0 / 3.0)
    min_log_hz = _F32(1000.0)
    min_log_mel = min_log_hz / f_sp
    logstep = _F32(0.06875174)
    m = _F32(m)
    return m * f_sp if m < min_log_mel else min_log_hz * np.exp((m - min_log_mel) * logstep, dtype=_F32)


def mel_filterbank(n_mels: int, sample_rate: int = SAMPLE_RATE, n_fft: int = N_FFT, fmin: float = 0.0,
                   fmax: float | None = None) -> np.ndarray:
    """[n_mels, n_fft/2+1] float32."""
    fmax = sample_rate / 2.0 if fmax is None else fmax
    n_freqs = n_fft // 2 + 1
    mel_min, mel_max = _hz_to_mel(fmin), _hz_to_mel(fmax)
    mel_points = [mel_min + (mel_max - mel_min) * _F32(i) / _F32(n_mels + 1) for i in range(n_mels + 2)]
    hz = [_mel_to_hz(m) for m in mel_points]
    freqs = [_F32(i) * _F32(sample_rate) / _F32(n_fft) for i in range(n_freqs)]
    fb = np.zeros((n_mels, n_freqs), dtype=_F32)
    for i in range(n_mels):
        lo, ce, up = hz[i], hz[i + 1], hz[i + 2]
        for j, fr in enumerate(freqs):
            if lo <= fr <= ce and ce > lo:
                fb[i, j] = (fr - lo) / (ce - lo)
            elif ce < fr <= up and up > ce:
                fb[i, j] = (up - fr) / (up - ce)
    return fb


def reflect_pad(samples: np.ndarray, pad: int = N_FFT // 2) -> np.ndarray:
    """mel.rs:180-193, including its index clamps for very short inputs."""
    n = len(samples)
    if n == 0:
        return np.zeros(2 * pad, dtype=_F32)
    left = [samples[min(i, max(n - 1, 0))] for i in range(pad, 0, -1)]
    right = [samples[max(max(n - 2, 0) - i, 0)] for i in range(pad)]
    return np.concatenate([np.asarray(left, dtype=_F32), samples.astype(_F32), np.asarray(right, dtype=_F32)])


def power_spectrogram(samples: np.ndarray) -> np.ndarray:
    """[n_frames, 201] float32."""
    padded = reflect_pad(np.asarray(samples, dtype=_F32))
    n_frames = (len(padded) - N_FFT) // HOP
    idx = (np.arange(n_frames) * HOP)[:, None] + np.arange(N_FFT)[None, :]
    frames = padded[idx] * hann_window()[None, :]
    spec = np.fft.rfft(frames.astype(_F32), axis=1)
    return (spec.real.astype(_F32) ** 2 + spec.imag.astype(_F32) ** 2).astype(_F32)


def log_mel(samples: np.ndarray, n_mels: int) -> np.ndarray:
    """compute_log: [n_frames, n_mels] float32."""
    mel = power_spectrogram(samples) @ mel_filterbank(n_mels).T
    lg = np.log10(np.maximum(mel.astype(_F32), _F32(1e-10))).astype(_F32)
    lg = np.maximum(lg, lg.max() - _F32(8.0))
    return ((lg + _F32(4.0)) / _F32(4.0)).astype(_F32)


def frame_clip(samples: np.ndarray, n_mels: int) -> np.ndarray:
    """transcribe.rs:46-82: one clip of 16 kHz mono samples -> model input [n_mels, 3000] float32."""
    s = np.asarray(samples, dtype=_F32)
    if len(s) < CHUNK_SAMPLES:
        s = np.concatenate([s, np.zeros(CHUNK_SAMPLES - len(s), dtype=_F32)])
    else:
        s = s[:CHUNK_SAMPLES]
    m = log_mel(s, n_mels).T                      # [n_mels, frames]
    out = np.zeros((n_mels, MEL_FRAMES), dtype=_F32)
    f = min(m.shape[1], MEL_FRAMES)
    out[:, :f] = m[:, :f]
    return out
