"""GPU log-mel front end (csrc/wq4_mel.cu) against the NumPy restatement of src/audio/mel.rs (oracle/mel_np.py).
The reference has no test or golden vector for mel.rs (parity unpinned); the oracle itself is cross-checked
against an independent torch.stft formulation when the fixture is generated (scripts/make_golden_mel.py)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import whisper_burn_b200 as wb
from oracle import mel_np as mn
from oracle import whisper_np as wn

DEV = "cuda:0"
TOL = 2e-4        # absolute, on values in [-1, 1.5]: f32 DFT rounding + log10 ulps


def synth(n, seed):
    rng = np.random.default_rng(seed)
    t = np.arange(n, dtype=np.float64) / 16000.0
    x = 0.3 * np.sin(2 * np.pi * (200 + 300 * t) * t) + 0.1 * np.sin(2 * np.pi * 3100 * t) * (t > 0.5) + 0.01 * rng.standard_normal(n)
    return x.astype(np.float32)


def test_test_wav_clip_matches_oracle_and_golden(golden_dir):
    g = np.load(golden_dir / "test_wav_mel.npz")
    pcm = g["pcm_s16"].astype(np.float32) / 32768.0
    for n_mels, key in ((128, "mel128"), (80, "mel80")):
        want = mn.frame_clip(pcm, n_mels)
        assert np.array_equal(want, g[key]), "oracle drifted from the committed fixture"
        got = wb.MelSpectrogram(n_mels, DEV).compute_log(torch.from_numpy(pcm).to(DEV)).cpu().numpy()
        assert got.shape == (n_mels, 3000)
        err = float(np.abs(got - want).max())
        print(f"test.wav n_mels {n_mels}: max abs err {err:.2e}")
        assert err < TOL


@pytest.mark.parametrize("n", [1, 2, 399, 16000, 479999, 480000, 480001, 600000])
def test_clip_lengths(n):
    """Empty-ish, shorter than one window, exactly 30 s, longer than 30 s (truncated, transcribe.rs:49-52)."""
    pcm = synth(n, n)
    want = mn.frame_clip(pcm, 80)
    got = wb.MelSpectrogram(80, DEV).compute_log(torch.from_numpy(pcm).to(DEV)).cpu().numpy()
    assert float(np.abs(got - want).max()) < TOL


def test_batch_of_different_clips_and_silence():
    clips = np.stack([synth(240000, 1), synth(240000, 2) * 0.001, np.zeros(240000, np.float32)])
    got = wb.MelSpectrogram(128, DEV).compute_log(torch.from_numpy(clips).to(DEV)).cpu().numpy()
    for b in range(3):
        want = mn.frame_clip(clips[b], 128)
        assert float(np.abs(got[b] - want).max()) < TOL, b
    # silence: log10(1e-10) = -10 everywhere -> (-10 + 4) / 4 = -1.5 exactly
    assert np.all(got[2] == -1.5)


def test_audio_to_tokens_matches_oracle(golden_dir):
    """Config-4 shape of the flow through the single C-ABI call wq4_transcribe (src/transcribe.rs:34-107): samples -> pad /
    truncate -> GPU log-mel -> encoder -> greedy decode -> ids >= 50257 dropped, for a ragged batch of clips, against the
    oracle running each clip through its own mel restatement, on a small "walk" checkpoint (all 224 ids comparable)."""
    g = np.load(golden_dir / "test_wav_mel.npz")
    pcm = g["pcm_s16"].astype(np.float32) / 32768.0
    cfg = wn.tiny_config(d=256, heads=4, enc_layers=1, dec_layers=2, n_audio_ctx=1500)
    _, blob, _ = wn.plain_and_walk_gguf(cfg, seed=17)
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, wb.WhisperConfig(**cfg.__dict__), DEV)
    state = wb.InferenceState(model, torch.device(DEV), cfg.n_mels)
    clips = [pcm, synth(50000, 3), synth(480000 + 1234, 4)]          # 1.92 s, 3.1 s, longer than 30 s (truncated)
    for lang in ("en", None):
        text, raw, ms = wb.transcribe(state, clips, lang)
        for b, clip in enumerate(clips):
            want, _, margins = oracle.transcribe(mn.frame_clip(clip, cfg.n_mels)[None], None if lang is None else wb.lang_token(lang),
                                                 return_margins=True)
            rel = np.array(margins) / np.array(oracle.last_maxabs)
            assert rel.min() > 1e-2
            assert raw[b] == want, (lang, b)
            assert text[b] == [t for t in want if 0 <= t < 50257]
        print(f"lang {lang}: 3 clips x 224 tokens identical, model part {ms:.1f} ms")
