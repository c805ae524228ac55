"""numpy restatement of the reference's model-level callers of the Q4 path (TEST INFRASTRUCTURE ONLY).

Everything is f32 and follows the reference's operation order; weights are the dequantised Q4_0
tensors (Q4Tensor::dequantize, src/gguf/tensor.rs:96-109), matmuls are plain f32 `x @ W.T`
(numpy/BLAS stands in for Burn's f32 matmul, which is third-party code outside /root/reference:
burn 0.20.1 -> cubek-matmul 0.1.1, Cargo.lock:386,1554).

Parity status: the Q4 dequantiser / quantiser underneath is pinned (tests/test_oracle.py); the
model-level flow (Conv1D, LayerNorm, SDPA, KV cache, greedy loop) is a restatement of the cited
source lines that NO reference test pins -- "parity unpinned" for those rows.

  Conv1D.forward            src/model/layers.rs:77-132
  gelu / LayerNorm          src/model/layers.rs:23-41
  scaled_dot_product_attention   src/model/attention.rs:243-298
  WhisperEncoder.forward    src/model/encoder.rs:37-49,87-115
  WhisperDecoder            src/model/decoder.rs:77-167,251-348
  WhisperModel.transcribe   src/model/whisper.rs:51-138
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from . import gguf_np, q4

SOT, EOT = 50258, 50257          # whisper.rs:15-16
MAX_TOKENS, MIN_TOKENS = 224, 3  # whisper.rs:20,97


@dataclass
class WhisperConfig:             # src/model/config.rs:5-28
    n_mels: int
    n_audio_ctx: int
    n_audio_state: int
    n_audio_head: int
    n_audio_layer: int
    n_text_ctx: int
    n_text_state: int
    n_text_head: int
    n_text_layer: int
    n_vocab: int
    n_lang: int

    @staticmethod
    def large_v3():              # config.rs:32-46
        return WhisperConfig(128, 1500, 1280, 20, 32, 448, 1280, 20, 32, 51865, 100)

    @staticmethod
    def medium():                # config.rs:49-63
        return WhisperConfig(80, 1500, 1024, 16, 24, 448, 1024, 16, 24, 51865, 99)

    def transcribe_token(self):  # config.rs:66-70
        return 50260 + self.n_lang

    def no_timestamps_token(self):  # config.rs:73-75
        return self.transcribe_token() + 4

    def lang_token_range(self):  # config.rs:78-80
        return range(50259, 50259 + self.n_lang)


def gelu(x):                     # layers.rs:35-41
    x = x.astype(np.float32)
    c = np.float32(np.sqrt(np.float32(2.0) / np.float32(np.pi)))
    x3 = x * x * x
    inner = (x + x3 * np.float32(0.044715)) * c
    return x * np.float32(0.5) * (np.tanh(inner) + np.float32(1.0))


def layernorm(x, w, b):          # layers.rs:23-31
    mean = x.mean(axis=-1, keepdims=True, dtype=np.float32)
    c = x - mean
    var = (c * c).mean(axis=-1, keepdims=True, dtype=np.float32)
    return (c / np.sqrt(var + np.float32(1e-5)) * w + b).astype(np.float32)


def conv1d(x, w, bias, padding, stride):   # layers.rs:77-132, im2col exactly as written
    B, C, T = x.shape
    O, C2, K = w.shape
    t_out = (T + 2 * padding - K) // stride + 1
    if padding:
        z = np.zeros((B, C, padding), np.float32)
        x = np.concatenate([z, x, z], axis=2)
    cols = []
    for k in range(K):
        if stride == 1:
            cols.append(x[:, :, k:k + t_out])
        else:
            needed = t_out * stride
            cols.append(x[:, :, k:k + needed].reshape(B, C, t_out, stride)[:, :, :, 0])
    col = np.concatenate(cols, axis=1).transpose(0, 2, 1)          # [B, t_out, K*C] kernel-major
    w_flat = w.transpose(0, 2, 1).reshape(O, K * C)                # [O, K, C] -> [O, K*C]
    out = col.astype(np.float32) @ w_flat.T.astype(np.float32) + bias
    return np.ascontiguousarray(out.transpose(0, 2, 1), dtype=np.float32)


def sdpa(q, k, v, n_heads, causal):        # attention.rs:243-298, head_dim = 64
    B, Tq, D = q.shape
    Tk = k.shape[1]
    hd = 64
    qh = q.reshape(B, Tq, n_heads, hd).transpose(0, 2, 1, 3)
    kh = k.reshape(B, Tk, n_heads, hd).transpose(0, 2, 1, 3)
    vh = v.reshape(B, Tk, n_heads, hd).transpose(0, 2, 1, 3)
    scores = (qh @ kh.transpose(0, 1, 3, 2)) / np.float32(np.sqrt(np.float32(hd)))
    if causal and Tq > 1:
        mask = np.zeros((Tq, Tk), np.float32)
        for i in range(Tq):
            mask[i, i + 1:] = -np.inf
        scores = scores + mask
    scores = scores - scores.max(axis=-1, keepdims=True)
    e = np.exp(scores)
    p = e / e.sum(axis=-1, keepdims=True)
    out = p.astype(np.float32) @ vh
    return np.ascontiguousarray(out.transpose(0, 2, 1, 3).reshape(B, Tq, D), dtype=np.float32)


class Linear:                    # Q4Linear, src/gguf/linear.rs:17-40
    def __init__(self, g: gguf_np.GgufFile, prefix: str):
        info = g.tensors[prefix + ".weight"]
        assert info.dtype == gguf_np.GGML_Q4_0, f"Expected Q4_0 for weight '{prefix}.weight'"   # loader.rs:132-134
        k, n = info.dims                                                    # loader.rs:126-131
        self.w = q4.dequantize_q4_0(g.tensor_data(prefix + ".weight"), n * k).reshape(n, k)
        self.b = g.tensor_f32(prefix + ".bias") if (prefix + ".bias") in g.tensors else None   # loader.rs:137-145

    def __call__(self, x):
        y = x.astype(np.float32) @ self.w.T
        return y + self.b if self.b is not None else y


class WhisperOracle:
    def __init__(self, gguf_bytes: bytes, cfg: WhisperConfig, threads: int | None = None):
        q4.set_threads(threads or q4.host_cores())
        g = gguf_np.GgufFile(gguf_bytes)
        self.cfg = cfg
        f = g.tensor_f32
        self.conv1 = (f("encoder.conv1.weight"), f("encoder.conv1.bias"))   # [out,in,k] (loader.rs:273-276)
        self.conv2 = (f("encoder.conv2.weight"), f("encoder.conv2.bias"))
        self.enc_pos = f("encoder.positional_embedding")
        ln = lambda p: (f(p + ".weight"), f(p + ".bias"))
        self.enc = []
        for i in range(cfg.n_audio_layer):
            p = f"encoder.blocks.{i}"
            self.enc.append(dict(attn_ln=ln(p + ".attn_ln"), q=Linear(g, p + ".attn.query"), k=Linear(g, p + ".attn.key"),
                                 v=Linear(g, p + ".attn.value"), o=Linear(g, p + ".attn.out"), mlp_ln=ln(p + ".mlp_ln"),
                                 fc1=Linear(g, p + ".mlp.0"), fc2=Linear(g, p + ".mlp.2")))
        self.ln_post = ln("encoder.ln_post")
        self.tok_emb = f("decoder.token_embedding.weight")
        self.dec_pos = f("decoder.positional_embedding")
        self.dec = []
        for i in range(cfg.n_text_layer):
            p = f"decoder.blocks.{i}"
            self.dec.append(dict(attn_ln=ln(p + ".attn_ln"), q=Linear(g, p + ".attn.query"), k=Linear(g, p + ".attn.key"),
                                 v=Linear(g, p + ".attn.value"), o=Linear(g, p + ".attn.out"),
                                 cross_ln=ln(p + ".cross_attn_ln"), cq=Linear(g, p + ".cross_attn.query"),
                                 ck=Linear(g, p + ".cross_attn.key"), cv=Linear(g, p + ".cross_attn.value"),
                                 co=Linear(g, p + ".cross_attn.out"), mlp_ln=ln(p + ".mlp_ln"),
                                 fc1=Linear(g, p + ".mlp.0"), fc2=Linear(g, p + ".mlp.2")))
        self.ln = ln("decoder.ln")

    # ---- encoder (encoder.rs:87-115) ----------------------------------------------------------
    def encode(self, mel):
        c = self.cfg
        x = gelu(conv1d(mel.astype(np.float32), *self.conv1, 1, 1))
        x = gelu(conv1d(x, *self.conv2, 1, 2))
        x = x.transpose(0, 2, 1)
        x = x + self.enc_pos[:x.shape[1], :x.shape[2]]
        for L in self.enc:                                                   # encoder.rs:37-49
            r = x
            h = layernorm(x, *L["attn_ln"])
            a = sdpa(L["q"](h), L["k"](h), L["v"](h), c.n_audio_head, False)
            x = r + L["o"](a)
            r = x
            h = layernorm(x, *L["mlp_ln"])
            x = r + L["fc2"](gelu(L["fc1"](h)))
        return layernorm(x, *self.ln_post)

    # ---- decoder (decoder.rs) -----------------------------------------------------------------
    def new_cache(self):
        return [dict(self_k=None, self_v=None, cross_k=None, cross_v=None) for _ in self.dec]

    def _block(self, L, x, enc, cache, mode):
        c = self.cfg
        r = x
        h = layernorm(x, *L["attn_ln"])
        q, k, v = L["q"](h), L["k"](h), L["v"](h)
        if mode == "init":                                                   # forward_init_cache, decoder.rs:140-167
            a = sdpa(q, k, v, c.n_text_head, True)
            cache["self_k"], cache["self_v"] = k, v                          # overwrites (decoder.rs:149-151)
        else:                                                                # forward_with_cache, decoder.rs:77-111
            if cache["self_k"] is not None:
                k = np.concatenate([cache["self_k"], k], axis=1)             # attention.rs:106-113
                v = np.concatenate([cache["self_v"], v], axis=1)
            cache["self_k"], cache["self_v"] = k, v
            a = sdpa(q, k, v, c.n_text_head, False)
        x = r + L["o"](a)
        r = x
        h = layernorm(x, *L["cross_ln"])
        q = L["cq"](h)
        if mode == "init" or cache["cross_k"] is None:                       # attention.rs:185-186 / 219-226
            cache["cross_k"], cache["cross_v"] = L["ck"](enc), L["cv"](enc)
        a = sdpa(q, cache["cross_k"], cache["cross_v"], c.n_text_head, False)
        x = r + L["co"](a)
        r = x
        h = layernorm(x, *L["mlp_ln"])
        return r + L["fc2"](gelu(L["fc1"](h)))

    def forward_prompt(self, tokens, enc, cache):                            # decoder.rs:251-296
        T = len(tokens)
        x = (self.tok_emb[np.asarray(tokens)] + self.dec_pos[:T])[None].astype(np.float32)
        for L, cb in zip(self.dec, cache):
            x = self._block(L, x, enc, cb, "init")
        x = layernorm(x, *self.ln)
        return (x[0, T - 1:T] @ self.tok_emb.T)[0]

    def decode_step(self, token, position, enc, cache):                      # decoder.rs:306-348
        x = (self.tok_emb[token:token + 1] + self.dec_pos[position:position + 1])[None].astype(np.float32)
        for L, cb in zip(self.dec, cache):
            x = self._block(L, x, enc, cb, "step")
        x = layernorm(x, *self.ln)
        return (x[0] @ self.tok_emb.T)[0]

    @staticmethod
    def argmax(logits):                                                      # whisper.rs:131-138: last maximum wins
        m = logits.max()
        return int(np.flatnonzero(logits == m)[-1])

    def transcribe(self, mel, lang_token: int | None, enc=None, return_margins=False, max_new_tokens=None):   # whisper.rs:51-128
        # max_new_tokens: per-clip token budget <= MAX_TOKENS (a batching extension of the C ABI; None = the reference's 224)
        c = self.cfg
        if enc is None:
            enc = self.encode(mel)
        cache = self.new_cache()
        position = 0
        margins = []
        self.last_maxabs = []            # max |logit| of every step whose margin is recorded (same order as margins)
        if lang_token is not None:
            prompt = [SOT, lang_token, c.transcribe_token(), c.no_timestamps_token()]
        else:
            logits = self.decode_step(SOT, 0, enc, cache)
            position = 1
            rng_ = list(c.lang_token_range())
            sub = logits[rng_[0]:rng_[-1] + 1]
            lang_token = rng_[0] + int(np.flatnonzero(sub == sub.max())[-1])  # max_by: last max
            s2 = np.sort(sub)
            margins.append(float(s2[-1] - s2[-2]))
            self.last_maxabs.append(float(np.abs(logits).max()))
            prompt = [lang_token, c.transcribe_token(), c.no_timestamps_token()]
        logits = self.forward_prompt(prompt, enc, cache)
        position += len(prompt)
        logits[EOT] = -np.inf
        nxt = self.argmax(logits)
        generated = []
        limit = MAX_TOKENS if max_new_tokens is None else max(0, min(MAX_TOKENS, int(max_new_tokens)))
        for step in range(limit):
            fin = logits[np.isfinite(logits)]
            s2 = np.partition(fin, -2)[-2:]
            margins.append(float(s2[1] - s2[0]))
            self.last_maxabs.append(float(np.abs(fin).max()))
            if nxt == EOT:
                break
            generated.append(nxt)
            if step == limit - 1:
                break
            logits = self.decode_step(nxt, position, enc, cache)
            position += 1
            if step + 1 < MIN_TOKENS:
                logits[EOT] = -np.inf
            nxt = self.argmax(logits)
        if return_margins:
            return generated, lang_token, margins
        return generated, lang_token


# ------------------------------------------------------------------------------------------------
# Random-init models (there is no network for checkpoints): seeded f32 state with the reference's
# tensor names, quantised with the converter rule (scripts/convert_whisper.py:33-96).
def sinusoids(length, channels, max_timescale=10000.0):
    inc = np.log(max_timescale) / (channels // 2 - 1)
    inv = np.exp(-inc * np.arange(channels // 2))
    t = np.arange(length)[:, None] * inv[None, :]
    return np.concatenate([np.sin(t), np.cos(t)], axis=1).astype(np.float32)


def random_state(cfg: WhisperConfig, seed: int, n_vocab: int = 51866, lin_std=0.02):
    rng = np.random.default_rng(seed)
    D, Dt = cfg.n_audio_state, cfg.n_text_state
    st = {}
    nrm = lambda shape, std: (rng.standard_normal(shape, dtype=np.float32) * np.float32(std))
    st["encoder.conv1.weight"] = nrm((D, cfg.n_mels, 3), 1.0 / np.sqrt(3 * cfg.n_mels))
    st["encoder.conv1.bias"] = nrm((D,), 0.02)
    st["encoder.conv2.weight"] = nrm((D, D, 3), 1.0 / np.sqrt(3 * D))
    st["encoder.conv2.bias"] = nrm((D,), 0.02)
    st["encoder.positional_embedding"] = sinusoids(cfg.n_audio_ctx, D)

    def ln(p, d):
        st[p + ".weight"] = (1.0 + nrm((d,), 0.05)).astype(np.float32)
        st[p + ".bias"] = nrm((d,), 0.05)

    def lin(p, n, k, bias=True, std=None):
        st[p + ".weight"] = nrm((n, k), std if std is not None else lin_std * np.sqrt(1024.0 / k) * 1.6)
        if bias:
            st[p + ".bias"] = nrm((n,), 0.02)

    for i in range(cfg.n_audio_layer):
        p = f"encoder.blocks.{i}"
        ln(p + ".attn_ln", D)
        lin(p + ".attn.query", D, D); lin(p + ".attn.key", D, D, bias=False); lin(p + ".attn.value", D, D); lin(p + ".attn.out", D, D)
        ln(p + ".mlp_ln", D)
        lin(p + ".mlp.0", 4 * D, D); lin(p + ".mlp.2", D, 4 * D)
    ln("encoder.ln_post", D)
    # mildly heavy-tailed row norms widen the top-1 / top-2 logit margin without collapsing the greedy
    # path onto one token (SURVEY 7 hard part G); chosen by scanning (sigma, scale) on the tiny config
    g = np.exp(rng.standard_normal(n_vocab).astype(np.float32) * 0.35)
    st["decoder.token_embedding.weight"] = (nrm((n_vocab, Dt), 0.01) * g[:, None]).astype(np.float32)
    st["decoder.positional_embedding"] = nrm((cfg.n_text_ctx, Dt), 0.1)
    for i in range(cfg.n_text_layer):
        p = f"decoder.blocks.{i}"
        ln(p + ".attn_ln", Dt)
        lin(p + ".attn.query", Dt, Dt); lin(p + ".attn.key", Dt, Dt, bias=False); lin(p + ".attn.value", Dt, Dt); lin(p + ".attn.out", Dt, Dt)
        ln(p + ".cross_attn_ln", Dt)
        lin(p + ".cross_attn.query", Dt, Dt); lin(p + ".cross_attn.key", Dt, Dt, bias=False)
        lin(p + ".cross_attn.value", Dt, Dt); lin(p + ".cross_attn.out", Dt, Dt)
        ln(p + ".mlp_ln", Dt)
        lin(p + ".mlp.0", 4 * Dt, Dt); lin(p + ".mlp.2", Dt, 4 * Dt)
    ln("decoder.ln", Dt)
    return st


# ------------------------------------------------------------------------------------------------
# "Walk" embeddings: a random-init model whose greedy path is NOT a fixed point.
#
# With plain random weights the residual stream carries token_embedding[tok] straight into the tied logits matrix, so
# arg-max = the input token and a 224-token transcription is one token repeated (round-1 finding).  Here the token and
# positional embeddings carry a 16-bit code on a few blocks of coordinates (on top of the random part the layers see):
#   token v < 2^15            : code = the bits of v, "special" bit (bit 15) = -1
#   token v >= 2^15 (specials): code = the bits of v - 2^15, special bit = +1
#   position p                : +-2 on a position-dependent subset of the 15 code bits ("forced"), 0 on the others
#                               ("copied" from the previous token), -2 on the special bit
# The final LayerNorm keeps signs, so logit_w ~ sum_i s_i * code_i(w) with s = code(prev token) + code(position): the
# winner is the token whose bits are sign(s), i.e. every bit is either forced by the position or copied from the
# previous token.  The greedy path therefore walks the vocabulary, depends on the fed-back token at every step (a wrong
# token breaks the copied bits), never lands on an id >= 2^15 (so never on EOT unless `eot_alias_step` asks for it),
# and the top-1 / top-2 margin is 2*min|s_i| out of sum|s_i| -- several per cent of the largest logit, far above the
# 1e-3 numeric error.  The 32 random layers still add their (audio-dependent) contribution to every coordinate; the code
# amplitude `c` is chosen so that contribution is a perturbation of the margin, not its source.
WALK_BITS = 15


def _walk_code(ids):
    ids = np.asarray(ids, dtype=np.int64)
    special = ids >= (1 << WALK_BITS)
    low = np.where(special, ids - (1 << WALK_BITS), ids)
    bits = ((low[:, None] >> np.arange(WALK_BITS)[None, :]) & 1) * 2 - 1
    return np.concatenate([bits, np.where(special, 1, -1)[:, None]], axis=1).astype(np.float32)   # [n, 16]


def walk_position_codes(n_ctx: int, seed: int, forced_frac: float = 0.6):
    rng = np.random.default_rng(seed + 7919)
    forced = rng.random((n_ctx, WALK_BITS)) < forced_frac
    sign = rng.integers(0, 2, (n_ctx, WALK_BITS)) * 2 - 1
    psi = np.where(forced, 2 * sign, 0)
    return np.concatenate([psi, np.full((n_ctx, 1), -2)], axis=1).astype(np.float32)             # [ctx, 16]


def walk_next(prev_token: int, position: int, psi) -> tuple[int, float]:
    """The structural prediction: next token and (2*min|s|) / sum|s| for prev_token fed at `position`."""
    s = _walk_code([prev_token])[0] + psi[position]
    bits = (s[:WALK_BITS] > 0).astype(np.int64)
    return int((bits << np.arange(WALK_BITS)).sum()), float(2 * np.abs(s).min() / np.abs(s).sum())


def walk_embeddings(state: dict, cfg: WhisperConfig, seed: int, c: float = 100.0, eot_alias_token: int | None = None):
    """Replaces decoder.token_embedding.weight / decoder.positional_embedding of `state` (in place) by their walk form.
    eot_alias_token: a regular token id whose code EOT takes over (and which gets EOT's): the walk then ends there."""
    tok = state["decoder.token_embedding.weight"]
    pos = state["decoder.positional_embedding"]
    n_vocab, D = tok.shape
    nb = WALK_BITS + 1
    bs = max(1, min(16, D // (2 * nb)))
    amp = np.float32(c / np.sqrt(bs))
    ids = np.arange(n_vocab)
    if eot_alias_token is not None:
        ids[EOT], ids[eot_alias_token] = eot_alias_token, EOT
    code = _walk_code(ids)
    psi = walk_position_codes(cfg.n_text_ctx, seed)
    tok[:, :nb * bs] += np.repeat(code, bs, axis=1) * amp
    pos[:, :nb * bs] += np.repeat(psi, bs, axis=1) * amp
    return psi


def should_quantize(name: str, shape) -> bool:      # scripts/convert_whisper.py:77-96
    if len(shape) != 2 or min(shape) < 256:
        return False
    for s in ("bias", "ln", "layer_norm", "positional_embedding", "token_embedding", "conv"):
        if s in name:
            return False
    return True


def state_to_gguf(state: dict) -> bytes:
    """Same tensor order (sorted names), dtypes and layout as convert_model (convert_whisper.py:154-217)."""
    q4.set_threads(q4.host_cores())
    tensors = []
    for name in sorted(state):
        a = np.ascontiguousarray(state[name], dtype=np.float32)
        if should_quantize(name, a.shape):
            tensors.append((name, a.shape, gguf_np.GGML_Q4_0, q4.quantize_q4_0_converter(a).tobytes()))
        else:
            tensors.append((name, a.shape, gguf_np.GGML_F32, a.tobytes()))
    return gguf_np.write_gguf(tensors, metadata=[("general.architecture", 8, "whisper"),
                                                 ("general.name", 8, "random-init")])


def state_to_tensors(state: dict):
    q4.set_threads(q4.host_cores())
    tensors = []
    for name in sorted(state):
        a = np.ascontiguousarray(state[name], dtype=np.float32)
        if should_quantize(name, a.shape):
            tensors.append((name, a.shape, gguf_np.GGML_Q4_0, q4.quantize_q4_0_converter(a).tobytes()))
        else:
            tensors.append((name, a.shape, gguf_np.GGML_F32, a.tobytes()))
    return tensors


def plain_and_walk_gguf(cfg: WhisperConfig, seed: int, eot_alias_token: int | None = None):
    """Two checkpoints that share every weight except the two decoder embedding tables: the plain random-init model
    (logits fully determined by the layers: the teacher-forced logits checks) and its walk form (token-parity checks).
    Quantises once.  Returns (plain_gguf, walk_gguf, psi)."""
    state = random_state(cfg, seed)
    tensors = state_to_tensors(state)
    meta = [("general.architecture", 8, "whisper"), ("general.name", 8, "random-init")]
    plain = gguf_np.write_gguf(tensors, metadata=meta)
    psi = walk_embeddings(state, cfg, seed, eot_alias_token=eot_alias_token)
    swap = {n: np.ascontiguousarray(state[n], dtype=np.float32) for n in ("decoder.token_embedding.weight", "decoder.positional_embedding")}
    tensors = [(n, sh, dt, swap[n].tobytes() if n in swap else raw) for n, sh, dt, raw in tensors]
    return plain, gguf_np.write_gguf(tensors, metadata=meta), psi


def teacher_token(clip: int, step: int) -> int:
    """Deterministic, varied teacher-forced token for (clip, step) -- regular vocabulary ids only."""
    return (1000 + 7919 * step + 2311 * clip + 17 * step * step) % 50000


def tiny_config(d=256, heads=4, enc_layers=2, dec_layers=2, n_mels=80, n_lang=99, n_audio_ctx=1500):
    return WhisperConfig(n_mels, n_audio_ctx, d, heads, enc_layers, 448, d, heads, dec_layers, 51865, n_lang)


def synthetic_mel(batch, n_mels, frames, seed):
    """Log-mel-like input: N(0,1) clipped to the range compute_log produces after (x+4)/4
    (src/audio/mel.rs:152), as SURVEY 8(d) config 2 specifies."""
    rng = np.random.default_rng(seed)
    return np.clip(rng.standard_normal((batch, n_mels, frames)).astype(np.float32) * 0.5, -1.0, 1.5)
