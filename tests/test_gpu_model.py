"""GPU parity of the callers around the Q4 path (encoder, decoder pass, greedy transcribe) against the
numpy oracle on seeded random-init Q4_0 models.  f16 tensor-core operands and f16 KV caches are used on
the GPU; the oracle is f32 throughout, so tolerances here are looser than the per-GEMM 1e-3."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import whisper_burn_b200 as wb
from oracle import whisper_np as wn

DEV = "cuda:0"


def rel_err(got, want):
    return float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))


@pytest.fixture(autouse=True)
def _flag():
    yield
    torch.cuda.synchronize()
    assert wb.lib().wq4_device_error_flag(0) == 0


def _cfg_pair(cfg):
    return wb.WhisperConfig(**cfg.__dict__)


@pytest.fixture(scope="module")
def tiny():
    cfg = wn.tiny_config(d=256, heads=4, enc_layers=2, dec_layers=2, n_audio_ctx=1500)
    blob = wn.state_to_gguf(wn.random_state(cfg, seed=7))
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, _cfg_pair(cfg), DEV)
    model.reserve(3)
    return cfg, oracle, model


@pytest.fixture(scope="module")
def tiny_walk():
    """Same layers, "walk" embedding tables (oracle/whisper_np.walk_embeddings): the greedy path visits ~220 distinct
    tokens with decision margins of several per cent of the largest logit, so EVERY token id is comparable."""
    cfg = wn.tiny_config(d=256, heads=4, enc_layers=2, dec_layers=2, n_audio_ctx=1500)
    _, blob, psi = wn.plain_and_walk_gguf(cfg, seed=7)
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, _cfg_pair(cfg), DEV)
    return cfg, oracle, model, psi


def test_encoder_matches_oracle(tiny):
    cfg, oracle, model = tiny
    mel = wn.synthetic_mel(2, cfg.n_mels, 3000, seed=11)
    want = oracle.encode(mel)
    got = model.encode(torch.from_numpy(mel).to(DEV)).cpu().numpy()
    assert got.shape == want.shape == (2, 1500, 256)
    e = rel_err(got, want)
    print("encoder rel err", e)
    assert e < 5e-3


@pytest.mark.parametrize("t_enc,batch", [(64, 2), (100, 1), (100, 3), (129, 1), (193, 2), (256, 1), (300, 3)])
def test_encoder_ragged_row_counts(t_enc, batch):
    """Row counts around the chained launch's 128-row blocks (rows = t_enc x batch: exactly one block, one row more, ragged
    last blocks, several clips per block) and below its threshold (100 rows: one launch per GEMM), against the oracle."""
    cfg = wn.tiny_config(d=256, heads=4, enc_layers=2, dec_layers=1, n_audio_ctx=t_enc)
    blob = wn.state_to_gguf(wn.random_state(cfg, seed=31, n_vocab=300))
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, _cfg_pair(cfg), DEV)
    mel = wn.synthetic_mel(batch, cfg.n_mels, 2 * t_enc, seed=32 + t_enc)
    want = oracle.encode(mel)
    got = model.encode(torch.from_numpy(mel).to(DEV)).cpu().numpy()
    assert got.shape == want.shape == (batch, t_enc, 256)
    e = rel_err(got, want)
    print(f"t_enc {t_enc} batch {batch}: encoder rel err {e:.3e}")
    assert e < 5e-3


def test_encode_replays_its_graph_with_fresh_inputs():
    """wq4_whisper_encode captures the encoder into a CUDA graph on the second call per batch size and replays it: different
    inputs on every call, a larger batch in between (re-reservation drops the graphs of the old workspaces), on a side stream."""
    cfg = wn.tiny_config(d=256, heads=4, enc_layers=2, dec_layers=1, n_audio_ctx=300)
    blob = wn.state_to_gguf(wn.random_state(cfg, seed=5, n_vocab=300))
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, _cfg_pair(cfg), DEV)
    counts = []
    with torch.cuda.stream(torch.cuda.Stream()):
        for i, b in enumerate((2, 2, 2, 2, 3, 3, 3, 2, 2, 2)):
            mel = wn.synthetic_mel(b, cfg.n_mels, 600, seed=40 + i)
            wb.reset_launch_count()
            got = model.encode(torch.from_numpy(mel).to(DEV)).cpu().numpy()
            counts.append(wb.launch_count())
            assert rel_err(got, oracle.encode(mel)) < 5e-3, i
    assert counts[2] == counts[3] == counts[9] and counts[2] > 2      # replayed passes report the launches of the graph


def test_encoder_short_context_reference_written_gguf(golden_dir):
    """Loads the GGUF the reference's converter wrote (tests/golden) and encodes: T_enc = 32 (one
    partial attention tile), vocab 64."""
    blob = (golden_dir / "tiny_whisper.gguf").read_bytes()
    cfg = wn.WhisperConfig(80, 32, 256, 4, 1, 16, 256, 4, 1, 64, 99)
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, _cfg_pair(cfg), DEV)
    assert model.vocab == 64 and model.d_model == 256 and model.ffn == 512
    mel = wn.synthetic_mel(3, 80, 64, seed=5)
    want = oracle.encode(mel)
    got = model.encode(torch.from_numpy(mel).to(DEV)).cpu().numpy()
    assert rel_err(got, want) < 5e-3
    # decoder passes on the same model: prompt of 3 tokens then two cached steps (teacher forced)
    enc0 = want[:1]
    cache = oracle.new_cache()
    lg = oracle.forward_prompt([1, 5, 9], enc0, cache)
    got_lg = model.decoder_pass([[1, 5, 9]] * 3, 0, 0, recompute_cross=True).cpu().numpy()
    assert rel_err(got_lg[0], lg) < 1e-2
    lg2 = oracle.decode_step(7, 3, enc0, cache)
    got2 = model.decoder_pass([[7]] * 3, 3, 3).cpu().numpy()
    assert rel_err(got2[0], lg2) < 1e-2
    lg3 = oracle.decode_step(11, 4, enc0, cache)
    got3 = model.decoder_pass([[11]] * 3, 4, 4).cpu().numpy()
    assert rel_err(got3[0], lg3) < 1e-2


def test_cuda_path_against_hf_second_opinion(golden_dir):
    """The CUDA path held directly to the Hugging Face outputs of scripts/make_golden_hf.py (an implementation independent
    of oracle/whisper_np.py) on the reference-written GGUF: encoder output and the logits of 12 positions, prompt pass then
    incremental steps over the f16 KV cache.  Tolerance as for the tiny-model oracle tests (f16 operands / caches)."""
    fx = np.load(golden_dir / "hf_second_opinion.npz")
    blob = (golden_dir / "tiny_whisper.gguf").read_bytes()
    cfg = wn.WhisperConfig(80, 32, 256, 4, 1, 16, 256, 4, 1, 64, 99)
    model = wb.load_whisper_from_gguf(blob, _cfg_pair(cfg), DEV)
    got = model.encode(torch.from_numpy(fx["mel"]).to(DEV)).cpu().numpy()
    e = rel_err(got, fx["enc_out"])
    print("encoder vs HF rel err", e)
    assert e < 5e-3
    toks = fx["tokens"]
    B = toks.shape[0]
    lg = model.decoder_pass([[int(t) for t in toks[b, :3]] for b in range(B)], 0, 0, recompute_cross=True).cpu().numpy()
    worst = max(rel_err(lg[b], fx["logits"][b, 2]) for b in range(B))
    for p in range(3, toks.shape[1]):
        lg = model.decoder_pass([[int(toks[b, p])] for b in range(B)], p, p).cpu().numpy()
        worst = max(worst, max(rel_err(lg[b], fx["logits"][b, p]) for b in range(B)))
    print("logits vs HF worst rel err", worst)
    assert worst < 1e-2


def test_decoder_logits_teacher_forced(tiny):
    cfg, oracle, model = tiny
    mel = wn.synthetic_mel(1, cfg.n_mels, 3000, seed=12)
    enc = oracle.encode(mel)
    model.encode(torch.from_numpy(mel).to(DEV), want_output=False)
    cache = oracle.new_cache()
    prompt = [wn.SOT, 50259, cfg.transcribe_token(), cfg.no_timestamps_token()]
    want = oracle.forward_prompt(prompt, enc, cache)
    got = model.decoder_pass([prompt], 0, 0, recompute_cross=True).cpu().numpy()[0]
    print("prompt logits rel err", rel_err(got, want))
    assert rel_err(got, want) < 1e-2
    assert int(got.argmax()) == int(want.argmax())
    tok, pos = int(want.argmax()), 4
    for step in range(6):
        want = oracle.decode_step(tok, pos, enc, cache)
        got = model.decoder_pass([[tok]], pos, pos).cpu().numpy()[0]
        assert rel_err(got, want) < 1e-2, step
        s = np.sort(want)
        if s[-1] - s[-2] > 1e-2 * np.abs(want).max():
            assert int(got.argmax()) == int(want.argmax())
        tok, pos = int(want.argmax()), pos + 1


@pytest.mark.parametrize("lang", ["en", None])
def test_transcribe_tokens_match_oracle(tiny_walk, lang):
    """Config-4 style token parity in both prompt modes (explicit language and auto-detect with the
    reference's cache-overwrite quirk), batch of 3 different clips vs the oracle clip by clip: all 224 ids."""
    cfg, oracle, model, _ = tiny_walk
    mel = wn.synthetic_mel(3, cfg.n_mels, 3000, seed=21)
    ids, langs = model.transcribe(torch.from_numpy(mel).to(DEV), lang)
    for b in range(3):
        want, wlang, margins = oracle.transcribe(mel[b:b + 1], None if lang is None else wb.lang_token(lang),
                                                 return_margins=True)
        rel = np.array(margins) / np.array(oracle.last_maxabs)
        print(f"clip {b}: {len(want)} tokens, {len(set(want))} distinct, min margin / max|logit| {rel.min():.4f}")
        assert langs[b] == wlang
        assert len(want) == 224 and len(set(want)) >= 50 and rel.min() > 1e-2
        assert ids[b] == want


def test_transcribe_token_budgets_and_compaction(tiny_walk):
    """Per-clip token budgets (wq4_whisper_transcribe_ex): clips retire at different steps, the active set is compacted
    on the device and the launches shrink (5 rows -> 3 -> 2 -> 1: skinny path down to the one-kernel step).  Every clip
    must produce exactly the oracle's tokens up to its budget."""
    cfg, oracle, model, _ = tiny_walk
    budgets = [224, 0, 9, 1, 40]
    mel = wn.synthetic_mel(5, cfg.n_mels, 3000, seed=33)
    ids, _ = model.transcribe(torch.from_numpy(mel).to(DEV), "en", max_new_tokens=budgets)
    for b, n in enumerate(budgets):
        want, _ = oracle.transcribe(mel[b:b + 1], wb.lang_token("en"), max_new_tokens=n)
        assert len(want) == n
        assert ids[b] == want, b
    # a second call with a LARGER batch than any before: reserve() grows (the reference's model has no batch ceiling)
    mel7 = wn.synthetic_mel(7, cfg.n_mels, 3000, seed=34)
    ids7, _ = model.transcribe(torch.from_numpy(mel7).to(DEV), "en", max_new_tokens=[3] * 7)
    want, _ = oracle.transcribe(mel7[6:7], wb.lang_token("en"), max_new_tokens=3)
    assert ids7[6] == want and all(len(t) == 3 for t in ids7)


def test_transcribe_stops_at_eot():
    """EOT handling of the batched loop (whisper.rs:105-108): a walk checkpoint whose path reaches EOT at step 20 (EOT
    takes over the code of the token the walk visits there): the clips stop with 20 tokens, EOT itself is not emitted."""
    cfg = wn.tiny_config(d=256, heads=4, enc_layers=1, dec_layers=2, n_audio_ctx=1500)
    psi = wn.walk_position_codes(cfg.n_text_ctx, 5)
    prev, seq = cfg.no_timestamps_token(), []
    for i in range(30):
        prev, _ = wn.walk_next(prev, 3 + i, psi)
        seq.append(prev)
    assert len(set(seq)) == 30
    _, blob, _ = wn.plain_and_walk_gguf(cfg, seed=5, eot_alias_token=seq[20])
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, _cfg_pair(cfg), DEV)
    mel = wn.synthetic_mel(4, cfg.n_mels, 3000, seed=35)
    ids, _ = model.transcribe(torch.from_numpy(mel).to(DEV), "en")
    for b in range(4):
        want, _ = oracle.transcribe(mel[b:b + 1], wb.lang_token("en"))
        assert want == seq[:20]
        assert ids[b] == want


def test_compacted_rows_logits_tiny(tiny8):
    """decoder_pass_slots on the small model at every path boundary: rows 1 (one-kernel step), 2..16 (skinny), 17+ (tcgen05)
    taken as shuffled subsets of 26 clips must reproduce the full-batch logits of those clips."""
    cfg, oracle, model = tiny8
    nb = 26
    mel = wn.synthetic_mel(nb, cfg.n_mels, 3000, seed=77)
    model.encode(torch.from_numpy(mel).to(DEV), want_output=False)
    prompts = [[wn.SOT, 50259 + b, cfg.transcribe_token(), cfg.no_timestamps_token()] for b in range(nb)]
    model.decoder_pass(prompts, 0, 0, recompute_cross=True)
    toks = [wn.teacher_token(b, 1) for b in range(nb)]
    full = model.decoder_pass([[t] for t in toks], 4, 4).clone()
    rng = np.random.default_rng(0)
    for rows in (1, 2, 3, 8, 9, 16, 17, 25):
        slots = rng.permutation(nb)[:rows].tolist()
        got = model.decoder_pass_slots([toks[c] for c in slots], slots, 4, 4)
        for r, c in enumerate(slots):
            e = float((got[r] - full[c]).abs().max() / full[c].abs().max())
            assert e < 2e-3, (rows, c, e)


@pytest.fixture(scope="module")
def tiny8():
    cfg = wn.tiny_config(d=256, heads=4, enc_layers=1, dec_layers=2, n_audio_ctx=1500)
    blob = wn.state_to_gguf(wn.random_state(cfg, seed=9))
    oracle = wn.WhisperOracle(blob, cfg)
    model = wb.load_whisper_from_gguf(blob, _cfg_pair(cfg), DEV)
    model.reserve(26)
    return cfg, oracle, model


@pytest.mark.parametrize("batch", [1, 2, 4, 5, 8, 11, 19, 26])
def test_decode_step_kernel_every_batch_size(tiny8, batch):
    """Decode steps at every batch size against the oracle clip by clip: different audio and different tokens per clip, so a
    wrong batch index, cache offset or split-KV merge shows.  Batch 1 runs the one-kernel decode step (csrc/wq4_decode.cu,
    template R = 1); 2, 4, 5, 8 and 11 the per-op path on the skinny linear (fused LayerNorm, cache append in the q|k|v
    epilogue, csrc/wq4_skinny.cu); 19 and 26 the per-op path on the tcgen05 GEMM; all per-op batches use the key-per-lane
    attention kernel.  The R = 2, 4 and 8 templates and the grouped launch of the one-kernel step are covered by the child-process test below."""
    cfg, oracle, model = tiny8
    mel = wn.synthetic_mel(batch, cfg.n_mels, 3000, seed=40 + batch)
    model.encode(torch.from_numpy(mel).to(DEV), want_output=False)
    encs = [oracle.encode(mel[b:b + 1]) for b in range(batch)]
    caches = [oracle.new_cache() for _ in range(batch)]
    prompts = [[wn.SOT, 50259 + b, cfg.transcribe_token(), cfg.no_timestamps_token()] for b in range(batch)]
    want = [oracle.forward_prompt(prompts[b], encs[b], caches[b]) for b in range(batch)]
    got = model.decoder_pass(prompts, 0, 0, recompute_cross=True).cpu().numpy()
    for b in range(batch):
        assert rel_err(got[b], want[b]) < 1e-2, ("prompt", b)
    toks = [int(w.argmax()) for w in want]
    pos = 4
    worst = 0.0
    for step in range(10):
        want = [oracle.decode_step(toks[b], pos, encs[b], caches[b]) for b in range(batch)]
        got = model.decoder_pass([[t] for t in toks], pos, pos).cpu().numpy()
        for b in range(batch):
            e = rel_err(got[b], want[b])
            worst = max(worst, e)
            assert e < 1e-2, (step, b, e)
        toks = [(int(w.argmax()) + 17 * b) % 50000 for b, w in enumerate(want)]     # distinct, deterministic
        pos += 1
    print(f"batch {batch}: worst logit rel err over 10 steps {worst:.2e}")


def test_load_from_path_equals_load_from_memory(golden_dir):
    """wq4_whisper_load_gguf (mmap of the file) and wq4_whisper_load_gguf_memory give the same model."""
    path = golden_dir / "tiny_whisper.gguf"
    cfg = wn.WhisperConfig(80, 32, 256, 4, 1, 16, 256, 4, 1, 64, 99)
    a = wb.load_whisper_from_gguf(str(path), _cfg_pair(cfg), DEV)
    b = wb.load_whisper_from_gguf(path.read_bytes(), _cfg_pair(cfg), DEV)
    mel = torch.from_numpy(wn.synthetic_mel(2, 80, 64, seed=3)).to(DEV)
    assert torch.equal(a.encode(mel), b.encode(mel))
    with pytest.raises(wb.Wq4Error):
        wb.load_whisper_from_gguf(str(golden_dir / "does_not_exist.gguf"), _cfg_pair(cfg), DEV)


def test_grouped_decode_kernel_in_child_process():
    """Batches 2, 4, 5, 8 (templates R = 2, 4 and 8 of the single-kernel step) and 11, 19 (the kernel run once per group of 8
    clips) with WQ4_DECODE_GROUP_MAX=24, which is read once per process; by default anything beyond 1 clip takes the per-op
    path, which the in-process cases above cover."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    if os.environ.get("WQ4_DECODE_GROUP_MAX"):
        pytest.skip("already inside the child")
    env = dict(os.environ, WQ4_DECODE_GROUP_MAX="24")
    r = subprocess.run([sys.executable, "-m", "pytest", str(Path(__file__)), "-q", "-x", "-m", "gpu", "-k",
                        "every_batch_size and (2 or 4 or 5 or 8 or 11 or 19)"], env=env, capture_output=True, text=True, timeout=900)
    print(r.stdout[-600:], r.stderr[-600:])
    import re
    m = re.search(r"(\d+) passed", r.stdout)          # "-k 2" also selects batch 26: at least the six listed cases
    assert r.returncode == 0 and m and int(m.group(1)) >= 6 and "failed" not in r.stdout


def test_per_op_decode_without_skinny_kernel_in_child_process():
    """WQ4_SKINNY=0 (read once per process) keeps the previous per-op decode path -- separate LayerNorm and cache-append
    launches, tcgen05 GEMM at 9..64 rows -- as the A/B arm of the skinny linear; batches 11 and 26 against the oracle."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    if os.environ.get("WQ4_SKINNY") or os.environ.get("WQ4_DECODE_GROUP_MAX"):
        pytest.skip("already inside a child")
    env = dict(os.environ, WQ4_SKINNY="0")
    r = subprocess.run([sys.executable, "-m", "pytest", str(Path(__file__)), "-q", "-x", "-m", "gpu", "-k",
                        "every_batch_size and (11 or 26)"], env=env, capture_output=True, text=True, timeout=900)
    print(r.stdout[-600:], r.stderr[-600:])
    assert r.returncode == 0 and "2 passed" in r.stdout


def _child(env_extra, kexpr, timeout=900):
    import os
    import subprocess
    import sys
    from pathlib import Path
    env = dict(os.environ, **env_extra)
    r = subprocess.run([sys.executable, "-m", "pytest", str(Path(__file__)), "-q", "-x", "-m", "gpu", "-k", kexpr],
                       env=env, capture_output=True, text=True, timeout=timeout)
    print(r.stdout[-600:], r.stderr[-600:])
    return r


@pytest.mark.parametrize("env", [{"WQ4_CHAIN": "0"}, {"WQ4_CHAIN_SPLITK": "1"}, {"WQ4_CHAIN_TILE_M": "192"}])
def test_encoder_arms_in_child_process(env):
    """The encoder's A/B arms (settings are read once per process): one launch per GEMM and LayerNorm (WQ4_CHAIN=0, the
    round-1 path), the chained launch with fc2 split along K through an epilogue dependency, and the 192-row tile of the
    chained launch at 1500 rows -- each against the oracle and the Hugging Face outputs."""
    import os
    if any(os.environ.get(k) for k in ("WQ4_CHAIN", "WQ4_CHAIN_SPLITK", "WQ4_CHAIN_TILE_M", "WQ4_DECODE_CHAIN")):
        pytest.skip("already inside a child")
    r = _child(env, "encoder_matches_oracle or hf_second_opinion")
    assert r.returncode == 0 and "2 passed" in r.stdout


@pytest.mark.parametrize("mode", ["1", "2"])
def test_chained_decode_steps_in_child_process(mode):
    """WQ4_DECODE_CHAIN=1 / 2: decode steps of 17+ clips on the chain kernel (LayerNorm folded into the consumer's epilogue;
    one launch per linear / one launch per group of linears), off by default because measured slower; batches 19 and 26
    against the oracle."""
    import os
    if any(os.environ.get(k) for k in ("WQ4_CHAIN", "WQ4_CHAIN_SPLITK", "WQ4_CHAIN_TILE_M", "WQ4_DECODE_CHAIN")):
        pytest.skip("already inside a child")
    r = _child({"WQ4_DECODE_CHAIN": mode, "WQ4_DECODE_CHAIN_FC2": "2"}, "every_batch_size and (19 or 26)")
    assert r.returncode == 0 and "2 passed" in r.stdout
