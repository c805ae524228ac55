#!/usr/bin/env python3
"""Independent second opinion for the rows the reference leaves unpinned (Conv1D stem, LayerNorm, SDPA, KV cache, logits).

The reference has no test or golden output for src/model/{layers,attention,encoder,decoder}.rs, so oracle/whisper_np.py is a
restatement nothing checks.  Hugging Face `transformers` is NOT the reference, but it is an independent implementation of
the same architecture, and tests/golden/tiny_whisper.gguf is the file the reference's own converter wrote from a HF model.
This script, run in the build container only:

  1. rebuilds that HF model (seed 0, scripts/make_golden.py) and checks the sha256 of its f32 tensors against
     names_golden.json, i.e. that it is the model the GGUF was converted from;
  2. replaces every tensor that made it into the GGUF by the value the GGUF holds (Q4_0 linears dequantised,
     F16 tensors widened), so HF and the oracle compute with identical weights;
  3. runs HF (float32, tanh-GELU as src/model/layers.rs:35-41 -- HF's stem hard-codes the erf form, patched here) on a
     seeded mel and a fixed token sequence, full-sequence (no cache);
  4. stores mel, tokens, encoder output and the logits of every position in tests/golden/hf_second_opinion.npz.

tests/test_oracle.py::test_oracle_matches_hf_second_opinion then asserts that the oracle's encode / forward_prompt /
decode_step (incremental, with its KV cache) reproduce them from the GGUF alone.
"""
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
OUT = REPO / "tests" / "golden"


def main():
    import torch
    import torch.nn.functional as F
    from transformers import WhisperConfig, WhisperForConditionalGeneration
    from oracle import gguf_np, q4

    torch.manual_seed(0)
    cfg = WhisperConfig(
        vocab_size=64, num_mel_bins=80, d_model=256, encoder_layers=1, decoder_layers=1,
        encoder_attention_heads=4, decoder_attention_heads=4, encoder_ffn_dim=512,
        decoder_ffn_dim=512, max_source_positions=32, max_target_positions=16,
        pad_token_id=0, bos_token_id=1, eos_token_id=2, decoder_start_token_id=1,
        activation_function="gelu_pytorch_tanh", attn_implementation="eager",
    )
    model = WhisperForConditionalGeneration(cfg).float().eval()
    names = json.loads((OUT / "names_golden.json").read_text())
    g = gguf_np.GgufFile((OUT / "tiny_whisper.gguf").read_bytes())
    sd = model.state_dict()
    replaced = 0
    for hf_name, info in names.items():
        if info["gguf"] is None:
            continue
        arr = sd[hf_name].float().numpy()
        assert hashlib.sha256(arr.tobytes()).hexdigest() == info["f32_sha256"], f"{hf_name}: not the converted model"
        if info["gguf"] not in g.tensors:
            continue
        ti = g.tensors[info["gguf"]]
        if ti.dtype == gguf_np.GGML_Q4_0:      # the pinned dequantiser (tensor.rs:96-109; bit-exact tests in tests/test_oracle.py)
            w = q4.dequantize_q4_0(g.tensor_data(info["gguf"]), arr.size).reshape(arr.shape)
        else:
            w = g.tensor_f32(info["gguf"]).reshape(arr.shape)
        sd[hf_name].copy_(torch.from_numpy(np.ascontiguousarray(w)))
        replaced += 1
    model.load_state_dict(sd)
    model.proj_out.weight = model.model.decoder.embed_tokens.weight      # decoder.rs:338-343: logits against token_embedding

    rng = np.random.default_rng(21)
    mel = np.clip(rng.standard_normal((2, 80, 64)).astype(np.float32) * 0.5, -1.0, 1.5)
    tokens = np.array([[1, 5, 9, 7, 11, 3, 40, 63, 22, 8, 17, 2], [1, 33, 12, 60, 4, 4, 19, 50, 6, 31, 2, 45]], dtype=np.int64)

    real_gelu = F.gelu
    F.gelu = lambda x, approximate="none": real_gelu(x, approximate="tanh")   # the stem's hard-coded erf-GELU -> tanh form
    try:
        with torch.no_grad():
            enc = model.model.encoder(torch.from_numpy(mel)).last_hidden_state
            out = model(input_features=torch.from_numpy(mel), decoder_input_ids=torch.from_numpy(tokens), use_cache=False)
    finally:
        F.gelu = real_gelu
    np.savez_compressed(OUT / "hf_second_opinion.npz", mel=mel, tokens=tokens.astype(np.int32),
                        enc_out=enc.numpy().astype(np.float32), logits=out.logits.numpy().astype(np.float32))
    print("tensors replaced by GGUF values:", replaced, "enc", tuple(enc.shape), "logits", tuple(out.logits.shape),
          "max|logit|", float(out.logits.abs().max()))


if __name__ == "__main__":
    sys.exit(main())
