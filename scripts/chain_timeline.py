#!/usr/bin/env python3
"""Tile timeline of one chained encoder-block launch (wq4_chain.cu): %globaltimer stamps of the leader CTA of every pair.
Usage: chain_timeline.py [medium|large_v3] [batch]"""
import ctypes as C, dataclasses, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from whisper_burn_b200.gguf_file import random_whisper_gguf
name = sys.argv[1] if len(sys.argv) > 1 else "medium"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda:0")
cfg = getattr(wb.WhisperConfig, name)()
cfg = dataclasses.replace(cfg, n_audio_layer=3, n_text_layer=1)
model = wb.load_whisper_from_gguf(random_whisper_gguf(cfg, seed=0), cfg, dev)
model.reserve(B)
rng = np.random.default_rng(0)
mel = torch.from_numpy(np.clip(rng.standard_normal((B, cfg.n_mels, 3000)).astype(np.float32) * 0.5, -1, 1.5)).to(dev)
for _ in range(3): model.encode(mel, want_output=False)
torch.cuda.synchronize()
L = wb.lib(); L.wq4_debug_chain_timeline.argtypes = [C.c_void_p]; L.wq4_debug_chain_timeline.restype = None
tl = torch.zeros(74 * 16 * 12, dtype=torch.int64, device=dev)
L.wq4_debug_chain_timeline(C.c_void_p(tl.data_ptr()))
model.encode(mel, want_output=False); torch.cuda.synchronize()
t = tl.cpu().numpy().reshape(74, 16, 12).astype(np.float64)
t0 = t[t > 0].min()
names = ["x:start", "x:dep_ok", "x:issued", "mma:acc_free", "mma:issued", "epi:ready", "epi:acc_full", "epi:done", "epi:chunks", "st:done", "st:start", "-"]
print("end of launch (us):", (t.max() - t0) / 1e3)
for pair in (0, 1, 30, 50, 73):
    print(f"pair {pair}")
    for it in range(16):
        if t[pair, it].max() == 0: break
        print("  tile", it, " ".join(f"{n}={(v - t0) / 1e3:7.2f}" if v > 0 else f"{n}=   -   " for n, v in zip(names, t[pair, it])))
# per-stamp summary over all pairs for each tile iteration
for it in range(8):
    m = t[:, it, :]
    if m.max() == 0: break
    ok = m[:, 7] > 0
    print(f"it {it}: pairs {ok.sum():2d}  dep wait {np.mean((m[ok,1]-m[ok,0]))/1e3:6.2f}  x issue {np.mean(m[ok,2]-m[ok,1])/1e3:6.2f}  mma {np.mean(m[ok,4]-m[ok,3])/1e3:6.2f}  "
          f"epi wait acc {np.mean(m[ok,6]-m[ok,5])/1e3:6.2f}  epi chunks {np.mean(m[ok,8]-m[ok,6])/1e3:6.2f} +signal {np.mean(m[ok,7]-m[ok,8])/1e3:5.2f}  stats {np.mean((m[ok,9]-m[ok,10])[m[ok,9]>0])/1e3 if (m[ok,9]>0).any() else 0:5.2f}  done at {np.mean(m[ok,7]-t0)/1e3:7.2f}")
print("error flag", L.wq4_device_error_flag(0))
