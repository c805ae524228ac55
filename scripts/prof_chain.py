#!/usr/bin/env python3
"""Runs a few encoder passes of a 2-layer model (target of `ncu -k regex:q4_chain`).  Usage: prof_chain.py [medium|large_v3] [batch]"""
import dataclasses, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from whisper_burn_b200.gguf_file import random_whisper_gguf
name = sys.argv[1] if len(sys.argv) > 1 else "medium"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda:0")
cfg = dataclasses.replace(getattr(wb.WhisperConfig, name)(), n_audio_layer=2, n_text_layer=1)
model = wb.load_whisper_from_gguf(random_whisper_gguf(cfg, seed=0), cfg, dev)
model.reserve(B)
rng = np.random.default_rng(0)
mel = torch.from_numpy(np.clip(rng.standard_normal((B, cfg.n_mels, 3000)).astype(np.float32) * 0.5, -1, 1.5)).to(dev)
for _ in range(3): model.encode(mel, want_output=False)
torch.cuda.synchronize()
print("ok errflag", wb.lib().wq4_device_error_flag(0))
