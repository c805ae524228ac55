"""GGUF v3 writer for Whisper Q4_0 checkpoints (the on-disk format of the drop-in surface,
src/gguf/reader.rs:13-48; same byte sequence as scripts/convert_whisper.py:188-217), plus a generator
of random-init checkpoints with the reference's tensor names for benchmarks (no network here)."""
from __future__ import annotations

import struct

import numpy as np

GGUF_MAGIC = 0x46554747
ALIGNMENT = 32
GGML_F32, GGML_F16, GGML_Q4_0 = 0, 1, 2


def _wstr(buf: bytearray, s: str) -> None:
    e = s.encode("utf-8")
    buf += struct.pack("<Q", len(e))
    buf += e


def write_gguf(tensors, name: str = "whisper") -> bytes:
    """tensors: iterable of (name, row-major shape, ggml dtype, raw bytes / uint8 array)."""
    tensors = list(tensors)
    buf = bytearray()
    buf += struct.pack("<IIQQ", GGUF_MAGIC, 3, len(tensors), 2)
    for key, val in (("general.architecture", "whisper"), ("general.name", name)):
        _wstr(buf, key)
        buf += struct.pack("<I", 8)
        _wstr(buf, val)
    cur, pads = 0, []
    for tname, shape, dtype, raw in tensors:
        aligned = (cur + ALIGNMENT - 1) // ALIGNMENT * ALIGNMENT
        pads.append(aligned - cur)
        _wstr(buf, tname)
        buf += struct.pack("<I", len(shape))
        for d in reversed(shape):          # GGUF stores dims reversed (loader.rs:126-131)
            buf += struct.pack("<Q", d)
        buf += struct.pack("<IQ", dtype, aligned)
        cur = aligned + len(raw)
    buf += b"\x00" * ((-len(buf)) % ALIGNMENT)
    out = [bytes(buf)]
    for pad, (_, _, _, raw) in zip(pads, tensors):
        out.append(b"\x00" * pad)
        out.append(raw.tobytes() if isinstance(raw, np.ndarray) else bytes(raw))
    return b"".join(out)


def random_q4_bytes(rng, n: int, k: int, std: float = 0.02) -> np.ndarray:
    """Random Q4_0 blocks for an [n,k] weight: uniform nibbles, f16 scales chosen so that the
    dequantised weights have roughly the given std (nibble-8 has std ~4.6)."""
    nb = n * k // 32
    raw = rng.integers(0, 256, (nb, 18), dtype=np.uint8)
    d = (rng.uniform(0.6, 1.4, nb) * (std / 4.6)).astype(np.float16)
    raw[:, :2] = d.view(np.uint8).reshape(nb, 2)
    return raw.reshape(-1)


def random_whisper_gguf(cfg, seed: int = 0, n_vocab: int = 51866, with_decoder: bool = True) -> bytes:
    """Random-init checkpoint with the tensor names of loader.rs:160-377 / convert_whisper.py:224-275."""
    rng = np.random.default_rng(seed)
    D, Dt = cfg.n_audio_state, cfg.n_text_state
    t = []
    f32 = lambda name, a: t.append((name, a.shape, GGML_F32, np.ascontiguousarray(a, dtype=np.float32)))
    q4 = lambda name, n, k: t.append((name, (n, k), GGML_Q4_0, random_q4_bytes(rng, n, k, 0.03 * np.sqrt(1024.0 / k))))
    nrm = lambda shape, std: rng.standard_normal(shape, dtype=np.float32) * np.float32(std)

    def ln(p, d):
        f32(p + ".weight", np.ones(d, np.float32))
        f32(p + ".bias", np.zeros(d, np.float32))

    def lin(p, n, k, bias=True):
        q4(p + ".weight", n, k)
        if bias:
            f32(p + ".bias", nrm((n,), 0.02))

    f32("encoder.conv1.weight", nrm((D, cfg.n_mels, 3), 1.0 / np.sqrt(3 * cfg.n_mels)))
    f32("encoder.conv1.bias", nrm((D,), 0.02))
    f32("encoder.conv2.weight", nrm((D, D, 3), 1.0 / np.sqrt(3 * D)))
    f32("encoder.conv2.bias", nrm((D,), 0.02))
    inc = np.log(10000.0) / (D // 2 - 1)
    tt = np.arange(cfg.n_audio_ctx)[:, None] * np.exp(-inc * np.arange(D // 2))[None, :]
    f32("encoder.positional_embedding", np.concatenate([np.sin(tt), np.cos(tt)], axis=1).astype(np.float32))
    for i in range(cfg.n_audio_layer):
        p = f"encoder.blocks.{i}"
        ln(p + ".attn_ln", D)
        lin(p + ".attn.query", D, D); lin(p + ".attn.key", D, D, False); lin(p + ".attn.value", D, D); lin(p + ".attn.out", D, D)
        ln(p + ".mlp_ln", D)
        lin(p + ".mlp.0", 4 * D, D); lin(p + ".mlp.2", D, 4 * D)
    ln("encoder.ln_post", D)
    nv = n_vocab if with_decoder else 8
    f32("decoder.token_embedding.weight", nrm((nv, Dt), 0.02))
    f32("decoder.positional_embedding", nrm((cfg.n_text_ctx, Dt), 0.02))
    for i in range(cfg.n_text_layer if with_decoder else 0):
        p = f"decoder.blocks.{i}"
        ln(p + ".attn_ln", Dt)
        lin(p + ".attn.query", Dt, Dt); lin(p + ".attn.key", Dt, Dt, False); lin(p + ".attn.value", Dt, Dt); lin(p + ".attn.out", Dt, Dt)
        ln(p + ".cross_attn_ln", Dt)
        lin(p + ".cross_attn.query", Dt, Dt); lin(p + ".cross_attn.key", Dt, Dt, False)
        lin(p + ".cross_attn.value", Dt, Dt); lin(p + ".cross_attn.out", Dt, Dt)
        ln(p + ".mlp_ln", Dt)
        lin(p + ".mlp.0", 4 * Dt, Dt); lin(p + ".mlp.2", Dt, 4 * Dt)
    ln("decoder.ln", Dt)
    return write_gguf(sorted(t, key=lambda e: e[0]), name="random-init")
