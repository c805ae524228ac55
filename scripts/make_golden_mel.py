#!/usr/bin/env python3
"""Generate tests/golden/test_wav_mel.npz in the build container (BASELINE.json config 4 input).

Reads the reference's audio asset /root/reference/test.wav (16 kHz mono s16, 30 720 samples = 1.92 s; no reference
test reads it and no expected output exists for it) and records
  pcm_s16    the samples (so the GPU box, which has no /root/reference, can run the same clip)
  mel128     oracle/mel_np.frame_clip(pcm, 128)  [128, 3000] f32 -- the Large-V3 encoder input for this clip
  mel80      same with 80 mel bins (Medium)
  stft_check max |oracle - independent torch.stft formulation| (must stay ~1e-5)
Run:  python scripts/make_golden_mel.py
"""
import sys
import wave
from pathlib import Path

import numpy as np
import torch

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
from oracle import mel_np as mn  # noqa: E402

w = wave.open("/root/reference/test.wav")
assert (w.getnchannels(), w.getframerate(), w.getsampwidth()) == (1, 16000, 2)
pcm_s16 = np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16).copy()
pcm = pcm_s16.astype(np.float32) / 32768.0
mel128, mel80 = mn.frame_clip(pcm, 128), mn.frame_clip(pcm, 80)

s = torch.zeros(mn.CHUNK_SAMPLES)
s[: len(pcm)] = torch.from_numpy(pcm)
st = torch.stft(s, 400, 160, window=torch.hann_window(400), return_complex=True)
ms = torch.from_numpy(mn.mel_filterbank(128)) @ (st[..., :-1].abs() ** 2)
lg = torch.clamp(ms, min=1e-10).log10()
lg = (torch.maximum(lg, lg.max() - 8.0) + 4.0) / 4.0
check = float((lg - torch.from_numpy(mel128)).abs().max())
assert check < 1e-4, check
out = REPO / "tests" / "golden" / "test_wav_mel.npz"
np.savez_compressed(out, pcm_s16=pcm_s16, mel128=mel128, mel80=mel80, stft_check=np.float32(check))
print("wrote", out, out.stat().st_size, "bytes; stft cross-check", check)
