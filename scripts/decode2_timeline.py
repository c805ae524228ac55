#!/usr/bin/env python3
"""Sub-phase times of the second-generation decode-step kernel (wq4_decode2.cu): CTA 0 stamps %globaltimer at every
STAMP() of the kernel (wq4_debug_decode2_timeline hook).  Usage: decode2_timeline.py [kv_len]"""
import ctypes as C, dataclasses, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from whisper_burn_b200.gguf_file import random_whisper_gguf
kv = int(sys.argv[1]) if len(sys.argv) > 1 else 224
dev = torch.device("cuda:0")
cfg = dataclasses.replace(wb.WhisperConfig.large_v3(), n_audio_layer=1)
model = wb.load_whisper_from_gguf(random_whisper_gguf(cfg, seed=0), cfg, dev)
model.reserve(1)
rng = np.random.default_rng(0)
mel = torch.from_numpy(np.clip(rng.standard_normal((1, cfg.n_mels, 3000)).astype(np.float32) * 0.5, -1, 1.5)).to(dev)
model.encode(mel, want_output=False)
model.decoder_pass([[50258, 50259, cfg.transcribe_token(), cfg.no_timestamps_token()]], 0, 0, recompute_cross=True)
for p in range(4, kv):
    model.decode_step_device(1, p, p)
for _ in range(3): model.decode_step_device(1, kv, kv)
torch.cuda.synchronize()
tl = torch.zeros(512, dtype=torch.int64, device=dev)
L = wb.lib(); L.wq4_debug_decode2_timeline.argtypes = [C.c_void_p]; L.wq4_debug_decode2_timeline.restype = None
acc = []
for rep in range(5):
    tl.zero_()
    L.wq4_debug_decode2_timeline(C.c_void_p(tl.data_ptr()))
    model.decode_step_device(1, kv, kv); torch.cuda.synchronize()
    L.wq4_debug_decode2_timeline(None)
    full = tl.cpu().numpy()
    sub = full[480:509].astype(np.float64)
    subs = globals().setdefault("subs", []); subs.append(np.diff(sub) / 1e3)
    t = full[:480]; t = t[t > 0]
    acc.append(np.diff(t) / 1e3)
d = np.median(np.stack(acc), axis=0)
n = len(d) + 1
# stamps: 1 (start) + per layer: [wait (l>0)] qkv-sync, attn, arrive | wait, cq-sync, attn, arrive | wait, fc1-sync, arrive  + tail: wait, logits, argmax
names0 = ["A:ln+qkv+xchg", "A:attn+xchg", "A:out+arrive", "B:wait", "B:red+ln+cq+xchg", "B:attn+xchg", "B:cout+arrive", "C:wait", "C:red+ln+fc1+xchg", "C:fc2+arrive"]
names = ["A:wait", "A:red+ln+qkv+xchg"] + names0[1:]
i = len(names0)
print("kv", kv, "stamps", n, "total us", float(d.sum()))
print("layer 0:", dict(zip(names0, np.round(d[:i], 2))))
per = d[i:i + 31 * len(names)].reshape(31, len(names))
print("mean layers 1..31 (us):", dict(zip(names, np.round(per.mean(0), 2))))
print("layer sum (us):", round(float(per.mean(0).sum()), 2))
print("tail (wait, red+ln+logits, argmax):", np.round(d[i + 31 * len(names):], 2))
sn = ["A:grid_wait", "A:reduce_x", "A:ln", "A:wait_w", "A:gemv_qkv", "A:cl_sync", "A:attn", "A:attn_xchg", "A:gemv_out", "A:arrive",
      "B:grid_wait", "B:reduce_x", "B:ln", "B:wait_w", "B:gemv_cq", "B:cl_sync", "B:kv+attn", "B:attn_xchg", "B:gemv_cout", "B:arrive",
      "C:grid_wait", "C:reduce_x", "C:ln", "C:wait_w", "C:gemv_fc1", "C:cl_sync", "C:gemv_fc2", "C:arrive"]
sd = np.median(np.stack(subs), axis=0)
print("layer 2 sub-phases (us, CTA 0 thread 0):")
for nm, v in zip(sn, sd): print(f"  {nm:14s} {v:6.2f}")
print("error flag", L.wq4_device_error_flag(0))
