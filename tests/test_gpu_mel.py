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
    """Config-4 shape of the flow: samples -> GPU log-mel -> encoder -> greedy decode, against the oracle running the same
    clip through its own mel restatement, on a tiny random-init model."""
    g = np.load(golden_dir / "test_wav_mel.npz")
    pcm = g["pcm_s16"].astype(np.float32) / 32768.0
    cfg = wn.tiny_config(d=256, heads=4, enc_layers=1, dec_layers=2, n_audio_ctx=1500)
    blob = wn.state_to_gguf(wn.random_state(cfg, seed=17))
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, wb.WhisperConfig(**cfg.__dict__), DEV)
    model.reserve(1)
    mel_gpu = wb.MelSpectrogram(cfg.n_mels, DEV).compute_log(torch.from_numpy(pcm).to(DEV).unsqueeze(0))
    ids, langs = model.transcribe(mel_gpu, "en")
    want, wlang, margins = oracle.transcribe(mn.frame_clip(pcm, cfg.n_mels)[None], wb.lang_token("en"), return_margins=True)
    n = len(want)
    for i, m in enumerate(margins):
        if m < 5e-3:
            n = min(n, i)
            break
    print(f"compared {n} of {len(want)} tokens")
    assert n >= 1 and ids[0][:n] == want[:n]
