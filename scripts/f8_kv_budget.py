#!/usr/bin/env python3
"""What an e4m3 cross K/V cache would cost in logit error: the NumPy oracle with the cross-attention
K/V of every decoder layer rounded to float8_e4m3fn (per-(layer, head) scale = amax / 448, K and V separately) or to f16
(what the CUDA path stores) against the unrounded f32 oracle, teacher-forced over a prompt and a few steps.
CPU only.  Usage: f8_kv_budget.py [dec_layers] [d_model]"""
import sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from oracle import whisper_np as wn

L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
D = int(sys.argv[2]) if len(sys.argv) > 2 else 512
cfg = wn.tiny_config(d=D, heads=D // 64, enc_layers=2, dec_layers=L, n_audio_ctx=1500)
blob = wn.state_to_gguf(wn.random_state(cfg, seed=9, n_vocab=2000))
mel = wn.synthetic_mel(1, cfg.n_mels, 3000, seed=10)

def round_kv(a, kind, heads):
    if kind == "f32": return a
    if kind == "f16": return a.astype(np.float16).astype(np.float32)
    b, t, d = a.shape
    x = torch.from_numpy(a.reshape(b, t, heads, 64).copy())
    s = x.abs().amax(dim=(1, 3), keepdim=True).clamp_min(1e-30) / 448.0
    return ((x / s).to(torch.float8_e4m3fn).to(torch.float32) * s).reshape(b, t, d).numpy()

def run(kind):
    o = wn.WhisperOracle(blob, cfg)
    enc = o.encode(mel)
    cache = o.new_cache()
    prompt = [1, 5, 9, 13]
    out = []
    # init pass computes the cross K/V; round them in place afterwards and run the steps on the rounded cache
    lg = o.forward_prompt(prompt, enc, cache)
    for cb in cache:
        cb["cross_k"] = round_kv(cb["cross_k"], kind, cfg.n_text_head)
        cb["cross_v"] = round_kv(cb["cross_v"], kind, cfg.n_text_head)
    tok, pos = 7, 4
    for step in range(6):
        lg = o.decode_step(tok, pos, enc, cache)
        out.append(lg.copy())
        tok, pos = (tok * 31 + 11) % 1900, pos + 1
    return np.stack(out)

ref = run("f32")
for kind in ("f16", "e4m3"):
    got = run(kind)
    err = np.abs(got - ref).max(axis=1) / np.abs(ref).max(axis=1)
    flips = int((got.argmax(1) != ref.argmax(1)).sum())
    print(f"{kind:5s} cross K/V, {L} decoder layers, D = {D}: worst logit rel err {err.max():.2e} (mean {err.mean():.2e}), arg-max flips {flips}/{len(err)}")
