#!/usr/bin/env python3
"""Per-phase times of the decode-step kernel: CTA 0 stamps %globaltimer after each grid barrier
(wq4_debug_decode_timeline hook).  Usage: decode_timeline.py [batch] [kv_len]"""
import ctypes as C, dataclasses, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from whisper_burn_b200.gguf_file import random_whisper_gguf
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
kv = int(sys.argv[2]) if len(sys.argv) > 2 else 224
dev = torch.device("cuda:0")
cfg = dataclasses.replace(wb.WhisperConfig.large_v3(), n_audio_layer=1)
model = wb.load_whisper_from_gguf(random_whisper_gguf(cfg, seed=0), cfg, dev)
model.reserve(B)
rng = np.random.default_rng(0)
mel = torch.from_numpy(np.clip(rng.standard_normal((B, cfg.n_mels, 3000)).astype(np.float32) * 0.5, -1, 1.5)).to(dev)
model.encode(mel, want_output=False)
model.decoder_pass([[50258, 50259, cfg.transcribe_token(), cfg.no_timestamps_token()]] * B, 0, 0, recompute_cross=True)
for _ in range(3): model.decode_step_device(B, kv, kv)
torch.cuda.synchronize()
tl = torch.zeros(512, dtype=torch.int64, device=dev)
L = wb.lib(); L.wq4_debug_decode_timeline.argtypes = [C.c_void_p]; L.wq4_debug_decode_timeline.restype = None
L.wq4_debug_decode_timeline(C.c_void_p(tl.data_ptr()))
model.decode_step_device(B, kv, kv); torch.cuda.synchronize()
L.wq4_debug_decode_timeline(None)
t = tl.cpu().numpy(); n = int((t != 0).sum()); t = t[:n]
t = t[t > 0]; n = len(t)
d = np.diff(t) / 1e3
names = ["P1 ln+qkv", "P2 self-attn", "P3 out", "P4 ln+cq", "P5 cross-attn", "P6 cout", "P7 ln+fc1", "P8 fc2"]
print("batch", B, "kv", kv, "barriers", n, "total us", (t[-1] - t[0]) / 1e3)
per = d[: 8 * 32].reshape(32, 8)
print("layer 1 phases (us):", dict(zip(names, np.round(per[1], 2))))
print("mean over layers  :", dict(zip(names, np.round(per.mean(0), 2))))
print("tail (final LN+logits, argmax):", np.round(d[8 * 32:], 2))
