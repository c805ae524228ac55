"""ctypes front-end of oracle/q4_oracle.c (TEST INFRASTRUCTURE ONLY, see the C header).

Every wrapper takes/returns numpy arrays and cites the reference lines its C twin follows.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None

Q4_BLOCK_SIZE = 32   # src/gguf/tests.rs:18
Q4_BLOCK_BYTES = 18  # src/gguf/tests.rs:19


def build(force: bool = False) -> Path:
    """Compile the C restatement with gcc (oracle/Makefile)."""
    so = _HERE / "libq4oracle.so"
    src = _HERE / "q4_oracle.c"
    if force or not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE), "-s"], check=True)
    return so


def lib() -> C.CDLL:
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(str(build()))
        f32p, u8p, sz = C.POINTER(C.c_float), C.POINTER(C.c_uint8), C.c_size_t
        L = _LIB
        L.orc_set_threads.argtypes = [C.c_int]
        L.orc_max_threads.restype = C.c_int
        L.orc_f32_to_f16.argtypes = [C.c_float]
        L.orc_f32_to_f16.restype = C.c_uint16
        L.orc_f16_to_f32.argtypes = [C.c_uint16]
        L.orc_f16_to_f32.restype = C.c_float
        L.orc_quantize_q4_0_tests.argtypes = [f32p, sz, u8p]
        L.orc_quantize_q4_0_converter.argtypes = [f32p, sz, u8p]
        L.orc_dequantize_q4_0.argtypes = [u8p, sz, f32p]
        L.orc_reference_matmul.argtypes = [f32p, f32p, sz, sz, sz, f32p]
        L.orc_q4_matmul_fused.argtypes = [f32p, u8p, sz, sz, sz, sz, f32p]
        L.orc_q4_linear.argtypes = [f32p, u8p, f32p, sz, sz, sz, sz, f32p]
        L.orc_gelu.argtypes = [f32p, sz]
        L.orc_q4_ffn.argtypes = [f32p, u8p, f32p, u8p, f32p, sz, sz, sz, sz, f32p, f32p]
        L.orc_layernorm.argtypes = [f32p, f32p, f32p, sz, sz, f32p]
        L.orc_conv1d.argtypes = [f32p, f32p, f32p, sz, sz, sz, sz, sz, sz, sz, f32p]
        L.orc_dequant_then_matmul.argtypes = [f32p, u8p, sz, sz, sz, f32p, f32p]
    return _LIB


def _f(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _b(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _cf(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def _cb(a) -> np.ndarray:
    if isinstance(a, (bytes, bytearray, memoryview)):
        a = np.frombuffer(a, dtype=np.uint8)
    return np.ascontiguousarray(a, dtype=np.uint8)


def set_threads(t: int) -> None:
    lib().orc_set_threads(int(t))


def max_threads() -> int:
    return int(lib().orc_max_threads())


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        return os.cpu_count() or 1


def quantize_q4_0_tests(data) -> np.ndarray:
    """src/gguf/tests.rs:24-57 (round-half-up, `as u8` saturation)."""
    d = _cf(data).ravel()
    assert d.size % Q4_BLOCK_SIZE == 0
    out = np.empty(d.size // Q4_BLOCK_SIZE * Q4_BLOCK_BYTES, dtype=np.uint8)
    lib().orc_quantize_q4_0_tests(_f(d), d.size, _b(out))
    return out


def quantize_q4_0_converter(data) -> np.ndarray:
    """scripts/convert_whisper.py:33-74 (round-half-even, & 0xF)."""
    d = _cf(data).ravel()
    assert d.size % Q4_BLOCK_SIZE == 0
    out = np.empty(d.size // Q4_BLOCK_SIZE * Q4_BLOCK_BYTES, dtype=np.uint8)
    lib().orc_quantize_q4_0_converter(_f(d), d.size, _b(out))
    return out


def quantize_q4_0_converter_np(data) -> np.ndarray:
    """Vectorised numpy twin of scripts/convert_whisper.py:33-74, used to cross-check the C
    version; asserted byte-identical to the reference's own function in tests/test_oracle.py."""
    flat = np.asarray(data, dtype=np.float32).ravel()
    blocks = flat.reshape(-1, Q4_BLOCK_SIZE)
    amax = np.max(np.abs(blocks), axis=1)
    d = np.where(amax > 0, amax / np.float32(7.0), np.float32(0.0)).astype(np.float32)
    safe = np.where(d > 0, d, np.float32(1.0)).astype(np.float32)
    q = np.where(d[:, None] > 0, np.round(blocks / safe[:, None]), 0).astype(np.int8)
    lo = (q[:, :16].astype(np.int32) + 8) & 0x0F
    hi = (q[:, 16:].astype(np.int32) + 8) & 0x0F
    out = np.empty((blocks.shape[0], Q4_BLOCK_BYTES), dtype=np.uint8)
    out[:, :2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:] = (lo | (hi << 4)).astype(np.uint8)
    return out.ravel()


def dequantize_q4_0(q4, n_elements: int) -> np.ndarray:
    """Q4Tensor::dequantize, src/gguf/tensor.rs:96-109 == src/gguf/tests.rs:60-87."""
    q = _cb(q4)
    assert n_elements % Q4_BLOCK_SIZE == 0
    assert q.size == n_elements // Q4_BLOCK_SIZE * Q4_BLOCK_BYTES
    out = np.empty(n_elements, dtype=np.float32)
    lib().orc_dequantize_q4_0(_b(q), n_elements, _f(out))
    return out


def reference_matmul(a, b_t, m: int, k: int, n: int) -> np.ndarray:
    """src/gguf/tests.rs:172-184."""
    a = _cf(a).ravel()
    b_t = _cf(b_t).ravel()
    assert a.size == m * k and b_t.size == n * k
    out = np.empty(m * n, dtype=np.float32)
    lib().orc_reference_matmul(_f(a), _f(b_t), m, k, n, _f(out))
    return out.reshape(m, n)


def q4_matmul(x, q4, n: int) -> np.ndarray:
    """CPU twin of q4_matmul / the WGSL kernel (src/gguf/op.rs:47, shader.wgsl:51-92).
    x [B,M,K] -> [B,M,N]."""
    x = _cf(x)
    assert x.ndim == 3, "Input must be 3D [B, M, K]"  # op.rs:53
    B, M, K = x.shape
    q = _cb(q4)
    assert q.size == n * K // 32 * Q4_BLOCK_BYTES
    out = np.empty((B, M, n), dtype=np.float32)
    lib().orc_q4_matmul_fused(_f(x), _b(q), B, M, K, n, _f(out))
    return out


def q4_linear(x, q4, bias, n: int) -> np.ndarray:
    """Q4Linear::forward, src/gguf/linear.rs:34-40."""
    x = _cf(x)
    B, M, K = x.shape
    q = _cb(q4)
    out = np.empty((B, M, n), dtype=np.float32)
    bp = None
    if bias is not None:
        bias = _cf(bias)
        bp = _f(bias)
    lib().orc_q4_linear(_f(x), _b(q), bp, B, M, K, n, _f(out))
    return out


def gelu(x) -> np.ndarray:
    """src/model/layers.rs:35-41 (tanh approximation)."""
    out = _cf(x).copy()
    lib().orc_gelu(_f(out), out.size)
    return out


def q4_ffn(x, w1, b1, w2, b2, f: int) -> np.ndarray:
    """Q4FFN::forward, src/model/layers.rs:54-58."""
    x = _cf(x)
    B, M, D = x.shape
    w1, w2 = _cb(w1), _cb(w2)
    hidden = np.empty((B, M, f), dtype=np.float32)
    out = np.empty((B, M, D), dtype=np.float32)
    b1 = None if b1 is None else _cf(b1)
    b2 = None if b2 is None else _cf(b2)
    lib().orc_q4_ffn(_f(x), _b(w1), None if b1 is None else _f(b1), _b(w2),
                     None if b2 is None else _f(b2), B, M, D, f, _f(hidden), _f(out))
    return out


def layernorm(x, w, b) -> np.ndarray:
    """LayerNorm::forward, src/model/layers.rs:23-31."""
    x = _cf(x)
    C_ = x.shape[-1]
    out = np.empty_like(x)
    w, b = _cf(w), _cf(b)
    lib().orc_layernorm(_f(x), _f(w), _f(b), x.size // C_, C_, _f(out))
    return out


def conv1d(x, w, bias, padding: int, stride: int) -> np.ndarray:
    """Conv1D::forward, src/model/layers.rs:77-132.  x [B,C_in,T], w [C_out,C_in,ks]."""
    x, w, bias = _cf(x), _cf(w), _cf(bias)
    B, C_in, T = x.shape
    C_out, C_in2, ks = w.shape
    assert C_in == C_in2
    T_out = (T + 2 * padding - ks) // stride + 1
    out = np.empty((B, C_out, T_out), dtype=np.float32)
    lib().orc_conv1d(_f(x), _f(w), _f(bias), B, C_in, T, C_out, ks, padding, stride, _f(out))
    return out


def dequant_then_matmul(x, q4, m: int, k: int, n: int) -> np.ndarray:
    """The two-pass CPU path of BASELINE.md section 3 (tensor.rs:96-109 then tests.rs:172-184)."""
    x = _cf(x).ravel()
    q = _cb(q4)
    wdeq = np.empty(n * k, dtype=np.float32)
    out = np.empty(m * n, dtype=np.float32)
    lib().orc_dequant_then_matmul(_f(x), _b(q), m, k, n, _f(wdeq), _f(out))
    return out.reshape(m, n)
