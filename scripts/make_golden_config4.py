#!/usr/bin/env python3
"""Generate tests/golden/config4_large_v3.npz and tests/golden/config_medium.npz by running the CPU oracle
(oracle/whisper_np.py) at FULL SIZE in the build container, so that the `-m gpu` tests can hold the CUDA path to
BASELINE.json config 4 (and a Medium-depth composed model) without a multi-minute CPU oracle run on the GPU box.
The GPU tests regenerate the same checkpoints from the seeds (oracle.whisper_np.plain_and_walk_gguf) and compare.

Large-V3 (seed 2):
  walk model  : greedy transcription of the reference's test.wav clip in both prompt modes of whisper.rs:66-88:
                224 token ids, per-step top-1/top-2 margin and max |logit| of the ORACLE.
  plain model : encoder output of the clip (selected rows + sums), and teacher-forced decoder logits of a batch of 18
                clips (6 distinct audio clips, per-clip token sequences whisper_np.teacher_token): prompt + 223 steps
                for clips 0..3, prompt + 39 steps for clips 4..17; stored at the steps in STEPS_* as 512 sampled
                vocabulary columns + top-8 + max |logit|.
Medium (seed 3): encoder rows + prompt and 8 steps of teacher-forced logits for 2 clips.

Run:  python scripts/make_golden_config4.py [--only large|medium]      (~40 min on 8 cores)
"""
import argparse
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import q4                      # noqa: E402
from oracle import whisper_np as wn       # noqa: E402

OUT = ROOT / "tests" / "golden"
STEPS_ALL = [0, 1, 2, 3, 4, 8, 16, 32, 39]          # step 0 = the prompt pass
STEPS_LONG = [64, 128, 200, 223]                    # clips 0..3 only
N_CLIPS, N_LONG, N_AUDIO = 18, 4, 6


def sample_columns(vocab: int, seed: int = 1234, n: int = 512):
    rng = np.random.default_rng(seed)
    fixed = np.array([0, 1, 50256, 50257, 50258, 50259, 50358, 50360, 50364, vocab - 1])
    rest = rng.choice(vocab, n - len(fixed), replace=False)
    return np.unique(np.concatenate([fixed, rest]))[:n].astype(np.int64)


def pack(logits, cols):
    top = np.argsort(logits)[-8:][::-1]
    return logits[cols].astype(np.float32), top.astype(np.int32), logits[top].astype(np.float32), np.float32(np.abs(logits).max())


def enc_summary(enc):
    rows = [0, 1, 500, enc.shape[1] - 1]
    return enc[0, rows].astype(np.float32), np.array([np.abs(enc).max(), enc.astype(np.float64).sum(), (enc.astype(np.float64) ** 2).sum()])


def clips_mels(cfg, mel0, n_audio):
    return np.concatenate([mel0] + [wn.synthetic_mel(1, cfg.n_mels, 3000, seed=300 + i) for i in range(1, n_audio)], axis=0)


def teacher_forced(oracle, cfg, encs, n_clips, n_long, steps_all, steps_long, n_audio, log):
    cols = sample_columns(oracle.tok_emb.shape[0])
    out = {}
    for b in range(n_clips):
        t0 = time.time()
        enc = encs[b % n_audio]
        cache = oracle.new_cache()
        prompt = [wn.SOT, 50259 + b, cfg.transcribe_token(), cfg.no_timestamps_token()]
        n_steps = max(steps_all + (steps_long if b < n_long else []))
        want = set(steps_all + (steps_long if b < n_long else []))
        lg = oracle.forward_prompt(prompt, enc, cache)
        if 0 in want:
            out[(b, 0)] = pack(lg, cols)
        for s in range(1, n_steps + 1):
            lg = oracle.decode_step(wn.teacher_token(b, s), 3 + s, enc, cache)
            if s in want:
                out[(b, s)] = pack(lg, cols)
        log(f"  clip {b}: {n_steps} steps in {time.time() - t0:.0f} s")
    return cols, out


def store_tf(d, prefix, cols, out):
    keys = sorted(out)
    d[prefix + "cols"] = cols
    d[prefix + "keys"] = np.array(keys, dtype=np.int32)
    d[prefix + "vals"] = np.stack([out[k][0] for k in keys])
    d[prefix + "top_ids"] = np.stack([out[k][1] for k in keys])
    d[prefix + "top_vals"] = np.stack([out[k][2] for k in keys])
    d[prefix + "maxabs"] = np.array([out[k][3] for k in keys], dtype=np.float32)


def large(log):
    cfg = wn.WhisperConfig.large_v3()
    t0 = time.time()
    plain, walk, psi = wn.plain_and_walk_gguf(cfg, seed=2)
    log(f"large_v3 checkpoints: {len(plain) / 1e6:.0f} MB each in {time.time() - t0:.0f} s")
    g = np.load(OUT / "test_wav_mel.npz")
    mel0 = np.ascontiguousarray(g["mel128"][None]).astype(np.float32)
    d = {"seed": np.int64(2)}
    # ---- walk model: token parity of the greedy loop
    oracle = wn.WhisperOracle(walk, cfg)
    t0 = time.time()
    enc0 = oracle.encode(mel0)              # the encoder does not depend on the decoder embeddings
    log(f"oracle encoder (test.wav): {time.time() - t0:.0f} s")
    d["enc_rows"], d["enc_stats"] = enc_summary(enc0)
    for mode, lang in (("en", 50259), ("auto", None)):
        t0 = time.time()
        toks, lt, margins = oracle.transcribe(mel0, lang, enc=enc0, return_margins=True)
        d[f"walk_{mode}_tokens"] = np.array(toks, dtype=np.int32)
        d[f"walk_{mode}_lang"] = np.int32(lt)
        d[f"walk_{mode}_margins"] = np.array(margins, dtype=np.float32)
        d[f"walk_{mode}_maxabs"] = np.array(oracle.last_maxabs, dtype=np.float32)
        rel = np.array(margins) / np.array(oracle.last_maxabs)
        log(f"walk {mode}: {len(toks)} tokens, {len(set(toks))} distinct, lang {lt}, min margin/max|logit| {rel.min():.4f} in {time.time() - t0:.0f} s")
    del oracle
    # ---- plain model: teacher-forced logits
    oracle = wn.WhisperOracle(plain, cfg)
    mels = clips_mels(cfg, mel0, N_AUDIO)
    encs = [enc0]
    for i in range(1, N_AUDIO):
        t0 = time.time()
        encs.append(oracle.encode(mels[i:i + 1]))
        log(f"oracle encoder (clip {i}): {time.time() - t0:.0f} s")
    d["enc_rows_all"] = np.stack([enc_summary(e)[0] for e in encs])
    d["enc_stats_all"] = np.stack([enc_summary(e)[1] for e in encs])
    cols, out = teacher_forced(oracle, cfg, encs, N_CLIPS, N_LONG, STEPS_ALL, STEPS_LONG, N_AUDIO, log)
    store_tf(d, "tf_", cols, out)
    np.savez_compressed(OUT / "config4_large_v3.npz", **d)
    log(f"wrote config4_large_v3.npz ({(OUT / 'config4_large_v3.npz').stat().st_size / 1e3:.0f} KB)")


def medium(log):
    cfg = wn.WhisperConfig.medium()
    t0 = time.time()
    plain, _, _ = wn.plain_and_walk_gguf(cfg, seed=3)
    log(f"medium checkpoint: {len(plain) / 1e6:.0f} MB in {time.time() - t0:.0f} s")
    g = np.load(OUT / "test_wav_mel.npz")
    mel0 = np.ascontiguousarray(g["mel80"][None]).astype(np.float32)
    oracle = wn.WhisperOracle(plain, cfg)
    mels = clips_mels(cfg, mel0, 2)
    encs = []
    for i in range(2):
        t0 = time.time()
        encs.append(oracle.encode(mels[i:i + 1]))
        log(f"oracle encoder (medium clip {i}): {time.time() - t0:.0f} s")
    d = {"seed": np.int64(3), "enc_rows_all": np.stack([enc_summary(e)[0] for e in encs]),
         "enc_stats_all": np.stack([enc_summary(e)[1] for e in encs])}
    cols, out = teacher_forced(oracle, cfg, encs, 2, 2, list(range(9)), [], 2, log)
    store_tf(d, "tf_", cols, out)
    np.savez_compressed(OUT / "config_medium.npz", **d)
    log("wrote config_medium.npz")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="", choices=["", "large", "medium"])
    a = ap.parse_args()
    q4.set_threads(q4.host_cores())
    t_start = time.time()

    def log(msg):
        print(f"[{time.time() - t_start:6.0f} s] {msg}", flush=True)
    if a.only in ("", "medium"):
        medium(log)
    if a.only in ("", "large"):
        large(log)


if __name__ == "__main__":
    main()
