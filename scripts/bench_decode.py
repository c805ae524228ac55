#!/usr/bin/env python3
"""Config 3 (BASELINE.json): Whisper Large-V3 greedy decoder step, M = batch, KV cache in place.

Times CUDA-graph replays of one decode step (wq4_whisper_decode_step_device: embed -> 32 decoder layers
-> final LN -> logits -> argmax on the device) at a fixed self-attention cache length, for a few batch
sizes.  Reports tokens/s and the HBM rate of the step's algorithmic bytes:
  Q4 weights of the step (412.9 MB at L-V3) + f16 cross K/V read + f16 self K/V read + logits matrix.
"""
import argparse
import dataclasses
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from whisper_burn_b200.gguf_file import random_whisper_gguf


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="large_v3")
    ap.add_argument("--batches", default="1,4,8")
    ap.add_argument("--kv", type=int, default=224)
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--enc-layers", type=int, default=1, help="encoder layers to instantiate (the encoder is not timed)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    cfg = wb.WhisperConfig.medium() if a.model == "medium" else wb.WhisperConfig.large_v3()
    cfg = dataclasses.replace(cfg, n_audio_layer=a.enc_layers)
    blob = random_whisper_gguf(cfg, seed=0)
    model = wb.load_whisper_from_gguf(blob, cfg, dev)
    del blob
    batches = [int(b) for b in a.batches.split(",")]
    model.reserve(max(batches))
    D, F, L, V, T = model.d_model, model.ffn, cfg.n_text_layer, model.vocab, cfg.n_audio_ctx
    q4_bytes = L * (6 * D * D + 2 * D * F) // 32 * 18          # self q,k,v,out + cross q,out + fc1 + fc2
    rng = np.random.default_rng(0)
    for B in batches:
        mel = torch.from_numpy(np.clip(rng.standard_normal((B, cfg.n_mels, 2 * T)).astype(np.float32) * 0.5, -1, 1.5)).to(dev)
        model.encode(mel, want_output=False)
        prompt = [[50258, 50259, cfg.transcribe_token(), cfg.no_timestamps_token()]] * B
        model.decoder_pass(prompt, 0, 0, recompute_cross=True)
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for _ in range(3):
                model.decode_step_device(B, a.kv, a.kv)
            torch.cuda.synchronize()
            wb.reset_launch_count()
            model.decode_step_device(B, a.kv, a.kv)
            launches = wb.launch_count()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                model.decode_step_device(B, a.kv, a.kv)
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            for _ in range(a.reps):
                g.replay()
            e1.record(s)
            torch.cuda.synchronize()
            dt = e0.elapsed_time(e1) * 1e-3 / a.reps
        kv_bytes = L * B * (T + a.kv + 1) * D * 2 * 2               # f16 K and V, cross + self
        logit_bytes = V * D * model_logit_elem_bytes()
        total = q4_bytes + kv_bytes + logit_bytes
        print(json.dumps({"model": a.model, "batch": B, "kv_len": a.kv, "us_per_step": dt * 1e6, "tokens_per_s": B / dt,
                          "launches_per_step": launches, "q4_weight_MB": q4_bytes / 1e6, "kv_MB": kv_bytes / 1e6,
                          "logits_MB": logit_bytes / 1e6, "GBps_q4_only": q4_bytes / dt / 1e9, "GBps_all": total / dt / 1e9}), flush=True)
    print("error flag", wb.lib().wq4_device_error_flag(0))


def model_logit_elem_bytes():
    import ctypes as C
    L = wb.lib()
    try:
        L.wq4_logits_elem_bytes.restype = C.c_int32
        return int(L.wq4_logits_elem_bytes())
    except AttributeError:
        return 2          # the step reads the f16 copy of the embedding matrix


if __name__ == "__main__":
    main()
