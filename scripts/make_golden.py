#!/usr/bin/env python3
"""Generate tests/golden/* by RUNNING the reference's own Python code in the build container.

The reference's only importable component is scripts/convert_whisper.py.  This script imports
it from /root/reference (read-only, absent on the GPU box) and records:

  q4_converter_golden.npz  inputs (f32) and the exact bytes quantize_q4_0
                           (convert_whisper.py:33-74) returns for them: seeded normals plus
                           edge blocks (zeros, constants, +/-amax, exact .5 ties, huge, tiny,
                           subnormal-scale values).
  names_golden.json        hf_name_to_gguf (:224-275) and should_quantize (:77-96) on the
                           parameter names/shapes of a HF Whisper model.
  tiny_whisper.gguf        a GGUF written by the reference's convert_model (:123-221), unmodified,
                           from a tiny random-init HF WhisperForConditionalGeneration (seed 0)
                           injected in place of the network download.
  tiny_whisper_state.npz   f32 originals of two of the quantised tensors (what went into
                           quantize_q4_0); sha256 of every f32 tensor is in names_golden.json.

Run:  python scripts/make_golden.py     (only in the build container)
"""
import hashlib
import importlib.util
import json
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
REF = Path("/root/reference/scripts/convert_whisper.py")
OUT = REPO / "tests" / "golden"


def load_reference_converter():
    spec = importlib.util.spec_from_file_location("ref_convert_whisper", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def quantiser_cases():
    rng = np.random.default_rng(0)
    cases = {}
    cases["normal_0p02"] = (rng.standard_normal(64 * 32) * 0.02).astype(np.float32)
    cases["normal_1"] = rng.standard_normal(32 * 32).astype(np.float32)
    cases["uniform"] = rng.uniform(-3, 3, 16 * 32).astype(np.float32)
    cases["zeros"] = np.zeros(32, np.float32)
    cases["const_half"] = np.full(32, 0.5, np.float32)
    cases["const_neg"] = np.full(64, -0.1, np.float32)
    cases["ramp"] = ((np.arange(32, dtype=np.float32) - 15.5) / 15.5).astype(np.float32)
    cases["ramp_x100"] = ((np.arange(32, dtype=np.float32) - 15.5) * 100.0).astype(np.float32)
    # exact ties: amax = 7 -> d = 1 -> values k + 0.5 round half-to-even
    ties = np.array([7.0] + [k + 0.5 for k in range(-7, 7)] + [-(k + 0.5) for k in range(0, 7)]
                    + [0.0] * 10, dtype=np.float32)[:32]
    cases["ties"] = ties
    cases["neg_amax"] = np.array([-7.0] + [1.0] * 31, dtype=np.float32)
    cases["huge"] = (rng.standard_normal(32) * 1e4).astype(np.float32)      # d = amax/7 fits f16
    cases["overflow_scale"] = (rng.standard_normal(32) * 1e6).astype(np.float32)  # f16(d) = inf
    cases["tiny"] = (rng.standard_normal(64) * 1e-6).astype(np.float32)     # subnormal f16 scale
    cases["underflow_scale"] = (rng.standard_normal(32) * 1e-9).astype(np.float32)  # f16(d) = 0
    cases["weights_cos"] = (np.cos(np.arange(128 * 64, dtype=np.float32) * np.float32(0.0007))
                            * np.float32(0.05)).astype(np.float32)
    return cases


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    ref = load_reference_converter()

    # --- quantiser goldens ---------------------------------------------------------------
    blob = {}
    with np.errstate(all="ignore"):
        for name, arr in quantiser_cases().items():
            blob[f"in_{name}"] = arr
            blob[f"out_{name}"] = np.frombuffer(ref.quantize_q4_0(arr), dtype=np.uint8)
    np.savez_compressed(OUT / "q4_converter_golden.npz", **blob)
    print("quantiser cases:", len(blob) // 2)

    # --- tiny HF whisper through the reference's convert_model ---------------------------
    import torch
    from transformers import WhisperConfig, WhisperForConditionalGeneration

    torch.manual_seed(0)
    cfg = WhisperConfig(
        vocab_size=64, num_mel_bins=80, d_model=256, encoder_layers=1, decoder_layers=1,
        encoder_attention_heads=4, decoder_attention_heads=4, encoder_ffn_dim=512,
        decoder_ffn_dim=512, max_source_positions=32, max_target_positions=16,
        pad_token_id=0, bos_token_id=1, eos_token_id=2, decoder_start_token_id=1,
    )
    model = WhisperForConditionalGeneration(cfg).float().eval()

    # The only change to the reference's environment: the network download is replaced by the
    # in-memory model above.  convert_model itself runs unmodified.
    real = WhisperForConditionalGeneration.from_pretrained
    WhisperForConditionalGeneration.from_pretrained = staticmethod(lambda name, **kw: model)
    try:
        ref.convert_model("tiny-random-whisper-seed0", str(OUT / "tiny_whisper.gguf"))
    finally:
        WhisperForConditionalGeneration.from_pretrained = real

    sd = model.state_dict()
    names = {}
    state = {}
    for hf_name, t in sd.items():
        g = ref.hf_name_to_gguf(hf_name)
        names[hf_name] = {"gguf": g, "shape": list(t.shape),
                          "q4": bool(g is not None and ref.should_quantize(g, tuple(t.shape)))}
        if g is not None:
            arr = t.float().numpy()
            names[hf_name]["f32_sha256"] = hashlib.sha256(arr.tobytes()).hexdigest()
            # keep the f32 originals of two quantised tensors (what went INTO quantize_q4_0)
            if g in ("decoder.blocks.0.attn.query.weight", "encoder.blocks.0.mlp.0.weight"):
                state[g] = arr
    (OUT / "names_golden.json").write_text(json.dumps(names, indent=1, sort_keys=True))
    np.savez_compressed(OUT / "tiny_whisper_state.npz", **state)
    print("gguf bytes:", (OUT / "tiny_whisper.gguf").stat().st_size, "tensors:", len(state))


if __name__ == "__main__":
    sys.exit(main())
