"""Host-side mirror of src/model/{config,whisper}.rs, src/gguf/loader.rs and src/transcribe.rs over
the model-level C ABI (wq4_whisper_*).  All compute runs in libwq4.so; torch only supplies device
buffers and the stream."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from .gguf import _as_device, _check_f32_cuda, _stream

# src/lib.rs:64-165 -- language code -> token id (50259 + index); unknown codes fall back to "en"
# exactly like WhisperTokenizer::lang_token (src/tokenizer.rs:38-44).
LANG_CODES = ("en zh de es ru ko fr ja pt tr pl ca nl ar sv it id hi fi vi he uk el ms cs ro da hu ta no th ur hr bg lt "
              "la mi ml cy sk te fa lv bn sr az sl kn et mk br eu is hy ne mn bs kk sq sw gl mr pa si km sn yo so af oc "
              "ka be tg sd gu am yi lo uz fo ht ps tk nn mt sa lb my bo tl mg as tt haw ln ha ba jw su").split()
MAX_TOKENS = 224   # src/model/whisper.rs:20


def lang_token(code: str) -> int:
    return 50259 + LANG_CODES.index(code) if code in LANG_CODES else 50259


class _CConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_mels", "n_audio_ctx", "n_audio_state", "n_audio_head", "n_audio_layer",
                                         "n_text_ctx", "n_text_state", "n_text_head", "n_text_layer", "n_vocab", "n_lang")]


@dataclass
class WhisperConfig:   # src/model/config.rs:5-28
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
    def large_v3():
        return WhisperConfig(128, 1500, 1280, 20, 32, 448, 1280, 20, 32, 51865, 100)

    @staticmethod
    def medium():
        return WhisperConfig(80, 1500, 1024, 16, 24, 448, 1024, 16, 24, 51865, 99)

    def transcribe_token(self) -> int:
        return 50260 + self.n_lang

    def no_timestamps_token(self) -> int:
        return self.transcribe_token() + 4

    def _c(self) -> _CConfig:
        return _CConfig(*[getattr(self, f[0]) for f in _CConfig._fields_])


def _bind(L):
    vp, i32, i64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_size_t
    if getattr(L, "_model_bound", False):
        return
    L.wq4_whisper_config_preset.argtypes = [C.c_char_p, C.POINTER(_CConfig)]
    L.wq4_whisper_load_gguf.argtypes = [C.c_char_p, C.POINTER(_CConfig), i32, C.POINTER(vp)]
    L.wq4_whisper_load_gguf_memory.argtypes = [vp, sz, C.POINTER(_CConfig), i32, C.POINTER(vp)]
    L.wq4_whisper_destroy.argtypes = [vp]
    L.wq4_whisper_reserve.argtypes = [vp, i64]
    L.wq4_whisper_info.argtypes = [vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]
    L.wq4_whisper_encode.argtypes = [vp, vp, i64, vp, vp]
    L.wq4_whisper_transcribe.argtypes = [vp, vp, i64, vp, vp, vp, vp, vp]
    L.wq4_whisper_transcribe_ex.argtypes = [vp, vp, i64, vp, vp, vp, vp, vp, vp]
    L.wq4_transcribe.argtypes = [vp, vp, vp, i64, vp, i64, vp, vp, vp, vp, vp, vp, vp]
    L.wq4_whisper_decoder_pass.argtypes = [vp, vp, i64, i64, i64, i64, i32, vp, vp]
    L.wq4_whisper_decode_step_device.argtypes = [vp, i64, i64, i64, vp]
    L.wq4_whisper_decoder_pass_slots.argtypes = [vp, vp, vp, i64, i64, i64, vp, vp]
    for fn in ("wq4_whisper_config_preset", "wq4_whisper_load_gguf", "wq4_whisper_load_gguf_memory", "wq4_whisper_destroy",
               "wq4_whisper_reserve", "wq4_whisper_info", "wq4_whisper_encode", "wq4_whisper_transcribe",
               "wq4_whisper_decoder_pass", "wq4_whisper_decode_step_device", "wq4_whisper_transcribe_ex", "wq4_transcribe", "wq4_whisper_decoder_pass_slots"):
        getattr(L, fn).restype = i32
    L._model_bound = True


class WhisperModel:
    """WhisperModel (src/model/whisper.rs:22-128) with batched entry points."""

    def __init__(self, handle, config: WhisperConfig, device: torch.device):
        self._h = C.c_void_p(handle)
        self.config = config
        self.device = device
        L = _lib.lib()
        d, f, v, wb = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        _lib.check(L.wq4_whisper_info(self._h, C.byref(d), C.byref(f), C.byref(v), C.byref(wb)))
        self.d_model, self.ffn, self.vocab, self.weight_bytes = d.value, f.value, v.value, wb.value

    def reserve(self, max_batch: int) -> None:
        _lib.check(_lib.lib().wq4_whisper_reserve(self._h, int(max_batch)))

    def _check_mel(self, mel: torch.Tensor) -> torch.Tensor:
        mel = _check_f32_cuda(mel, "mel", self.device)
        c = self.config
        if mel.dim() != 3 or mel.shape[0] < 1 or mel.shape[1] != c.n_mels or mel.shape[2] != 2 * c.n_audio_ctx:
            raise _lib.Wq4Error(2, f"mel must be [B, {c.n_mels}, {2 * c.n_audio_ctx}], got {list(mel.shape)}")
        return mel

    def encode(self, mel: torch.Tensor, want_output: bool = True):
        """mel [B, n_mels, 3000] f32 (device) -> [B, 1500, D] f32 (whisper.rs:41-43)."""
        mel = self._check_mel(mel)
        c = self.config
        b = mel.shape[0]
        out = torch.empty((b, c.n_audio_ctx, self.d_model), dtype=torch.float32, device=self.device) if want_output else None
        _lib.check(_lib.lib().wq4_whisper_encode(self._h, C.c_void_p(mel.data_ptr()), b,
                                                 None if out is None else C.c_void_p(out.data_ptr()),
                                                 C.c_void_p(_stream(self.device))))
        return out

    def transcribe(self, mel: torch.Tensor, lang_code: str | None, max_new_tokens=None):
        """Greedy decoding (whisper.rs:51-128).  mel [B, n_mels, 3000]; one language (or None = auto)
        for the whole batch.  max_new_tokens: optional per-clip token budgets (<= 224).
        Returns (list of token-id lists, list of language tokens)."""
        mel = self._check_mel(mel)
        b = mel.shape[0]
        langs = np.full(b, -1 if lang_code is None else lang_token(lang_code), dtype=np.int32)
        toks = np.zeros((b, MAX_TOKENS), dtype=np.int32)
        counts = np.zeros(b, dtype=np.int32)
        det = np.zeros(b, dtype=np.int32)
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        budget = None
        if max_new_tokens is not None:
            budget = np.ascontiguousarray(np.broadcast_to(np.asarray(max_new_tokens, dtype=np.int32), (b,)))
        _lib.check(_lib.lib().wq4_whisper_transcribe_ex(self._h, C.c_void_p(mel.data_ptr()), b, vp(langs),
                                                        None if budget is None else vp(budget), vp(toks),
                                                        vp(counts), vp(det), C.c_void_p(_stream(self.device))))
        return [toks[i, :counts[i]].tolist() for i in range(b)], det.tolist()

    def decoder_pass(self, tokens, emb_pos0: int, kv_pos0: int, recompute_cross: bool = False) -> torch.Tensor:
        """Test hook: tokens [B][q_len] -> logits [B, vocab] of the last position."""
        t = np.ascontiguousarray(np.asarray(tokens, dtype=np.int32))
        b, q = t.shape
        out = torch.empty((b, self.vocab), dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().wq4_whisper_decoder_pass(self._h, t.ctypes.data_as(C.c_void_p), b, q, emb_pos0, kv_pos0,
                                                       1 if recompute_cross else 0, C.c_void_p(out.data_ptr()),
                                                       C.c_void_p(_stream(self.device))))
        return out

    def decoder_pass_slots(self, tokens, slots, emb_pos: int, kv_pos: int) -> torch.Tensor:
        """Test hook: one single-query step over the clips `slots` (row r = clip slots[r]) -> logits [rows, vocab]."""
        t = np.ascontiguousarray(np.asarray(tokens, dtype=np.int32).reshape(-1))
        sl = np.ascontiguousarray(np.asarray(slots, dtype=np.int32).reshape(-1))
        assert t.size == sl.size
        out = torch.empty((t.size, self.vocab), dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().wq4_whisper_decoder_pass_slots(self._h, t.ctypes.data_as(C.c_void_p), sl.ctypes.data_as(C.c_void_p),
                                                             t.size, emb_pos, kv_pos, C.c_void_p(out.data_ptr()),
                                                             C.c_void_p(_stream(self.device))))
        return out

    def decode_step_device(self, batch: int, position: int, kv_len: int) -> None:
        _lib.check(_lib.lib().wq4_whisper_decode_step_device(self._h, batch, position, kv_len,
                                                             C.c_void_p(_stream(self.device))))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None and h.value:
            try:
                _lib.lib().wq4_whisper_destroy(h)
            except Exception:
                pass


def load_whisper_from_gguf(path_or_bytes, config: WhisperConfig, device="cuda") -> WhisperModel:
    """load_whisper_from_gguf (src/gguf/loader.rs:26-46); also accepts the file's bytes."""
    L = _lib.lib()
    _bind(L)
    dev = _as_device(device)
    cc = config._c()
    h = C.c_void_p()
    if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
        buf = np.frombuffer(path_or_bytes, dtype=np.uint8)
        _lib.check(L.wq4_whisper_load_gguf_memory(buf.ctypes.data_as(C.c_void_p), buf.size, C.byref(cc), dev.index, C.byref(h)))
    else:
        _lib.check(L.wq4_whisper_load_gguf(str(path_or_bytes).encode(), C.byref(cc), dev.index, C.byref(h)))
    return WhisperModel(h.value, config, dev)


@dataclass
class InferenceState:   # src/transcribe.rs:26-31 (tokenizer omitted: parity is on token ids)
    model: WhisperModel
    device: torch.device
    n_mels: int


def transcribe(state: InferenceState, samples, language: str | None, mel=None):
    """transcribe() (src/transcribe.rs:34-107) for a batch of utterances through the single C-ABI call wq4_transcribe:
    samples = list of 1-D f32 arrays / tensors (16 kHz mono) -> (text-token ids per clip [ids >= 50257 dropped],
    raw ids per clip, elapsed ms of the model part)."""
    from .audio import MelSpectrogram
    mel = mel or MelSpectrogram(state.n_mels, state.device)
    clips = [torch.as_tensor(s, dtype=torch.float32).reshape(-1) for s in samples]
    b = len(clips)
    lens = np.array([int(c.numel()) for c in clips], dtype=np.int64)
    stride = max(1, int(lens.max()))
    pcm = torch.zeros((b, stride), dtype=torch.float32, device=state.device)
    for i, c in enumerate(clips):
        pcm[i, :c.numel()] = c.to(state.device)
    langs = np.full(b, -1 if language is None else lang_token(language), dtype=np.int32)
    text = np.zeros((b, MAX_TOKENS), dtype=np.int32)
    raw = np.zeros((b, MAX_TOKENS), dtype=np.int32)
    tc, rc = np.zeros(b, dtype=np.int32), np.zeros(b, dtype=np.int32)
    ms = C.c_double()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    _lib.check(_lib.lib().wq4_transcribe(state.model._h, mel._h, C.c_void_p(pcm.data_ptr()), stride, vp(lens), b, vp(langs),
                                         vp(text), vp(tc), vp(raw), vp(rc), C.byref(ms), C.c_void_p(_stream(state.device))))
    return [text[i, :tc[i]].tolist() for i in range(b)], [raw[i, :rc[i]].tolist() for i in range(b)], ms.value


def transcribe_mels(state: InferenceState, mel: torch.Tensor, language: str | None):
    """The GPU part of transcribe() (src/transcribe.rs:84-104) for a batch of log-mel clips
    [B, n_mels, 3000]: returns (token ids with ids >= 50257 dropped, raw ids, elapsed ms)."""
    import time
    torch.cuda.synchronize(state.device)
    t0 = time.perf_counter()
    ids, _ = state.model.transcribe(mel, language)
    ms = (time.perf_counter() - t0) * 1e3
    return [[t for t in seq if 0 <= t < 50257] for seq in ids], ids, ms
