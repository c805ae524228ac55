"""Host-side mirror of the reference's src/gguf/{tensor,op,linear}.rs over the C ABI.

Same names, argument meaning and error behaviour as the reference:
  Q4Tensor.from_q4_bytes(raw_bytes, [n, k], device)   src/gguf/tensor.rs:35
  Q4Tensor.shape() / num_blocks() / dequantize()       src/gguf/tensor.rs:74,79,88
  q4_matmul(input[B,M,K], weights) -> [B,M,N]          src/gguf/op.rs:47
  Q4Linear(weights, bias).forward(x)                   src/gguf/linear.rs:26,34
torch supplies device memory and the current stream only; all arithmetic runs in libwq4.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

_WS: dict[tuple[int, int], torch.Tensor] = {}


def _stream(device: torch.device) -> int:
    if not torch.cuda.is_available():   # let the C ABI report "no device" itself
        return 0
    return int(torch.cuda.current_stream(device).cuda_stream)


def workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Per-(device, stream) scratch buffer, grown on demand (torch owns the memory)."""
    key = (device.index or 0, _stream(device))
    buf = _WS.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _WS[key] = buf
    return buf


def _as_device(device) -> torch.device:
    d = torch.device(device) if not isinstance(device, torch.device) else device
    if d.type != "cuda":
        raise _lib.Wq4Error(4, f"device {d} is not a CUDA device; there is no CPU fallback")
    if d.index is not None:
        return d
    return torch.device("cuda", torch.cuda.current_device() if torch.cuda.is_available() else 0)


def _check_f32_cuda(x: torch.Tensor, name: str, device: torch.device | None = None) -> torch.Tensor:
    if not x.is_cuda:
        raise _lib.Wq4Error(4, f"{name} must live on a CUDA device; there is no CPU fallback")
    if x.dtype != torch.float32:
        raise _lib.Wq4Error(1, f"{name} must be float32, got {x.dtype}")
    if device is not None and x.device != device:      # the library dereferences the pointer on the weights' device
        raise _lib.Wq4Error(1, f"{name} lives on {x.device}, the weights on {device}")
    return x.contiguous()  # op.rs:50 into_contiguous


class Q4Tensor:
    """A Q4_0 quantised weight [N, K] living on the GPU (src/gguf/tensor.rs:21-27)."""

    def __init__(self, handle: int, shape, device: torch.device):
        self._h = C.c_void_p(handle)
        self._shape = [int(shape[0]), int(shape[1])]
        self.device = device

    @classmethod
    def from_q4_bytes(cls, raw_bytes, shape, device) -> "Q4Tensor":
        n, k = int(shape[0]), int(shape[1])
        if isinstance(raw_bytes, torch.Tensor):
            raw_bytes = raw_bytes.detach().cpu().numpy()
        raw = np.ascontiguousarray(np.frombuffer(raw_bytes, dtype=np.uint8)
                                   if isinstance(raw_bytes, (bytes, bytearray, memoryview))
                                   else np.asarray(raw_bytes, dtype=np.uint8))
        dev = _as_device(device)
        out = C.c_void_p()
        L = _lib.lib()
        _lib.check(L.wq4_tensor_create(raw.ctypes.data_as(C.c_void_p), raw.size, n, k, dev.index,
                                       C.c_void_p(_stream(dev)), C.byref(out)))
        return cls(out.value, (n, k), dev)

    def shape(self):
        return list(self._shape)

    def num_blocks(self) -> int:
        return int(_lib.lib().wq4_tensor_num_blocks(self._h))

    def device_bytes(self) -> int:
        return int(_lib.lib().wq4_tensor_device_bytes(self._h))

    def dequantize(self) -> torch.Tensor:
        """[N, K] float32 on the device, bit-identical to Q4Tensor::dequantize (tensor.rs:88-113)."""
        n, k = self._shape
        out = torch.empty((n, k), dtype=torch.float32, device=self.device)
        _lib.check(_lib.lib().wq4_tensor_dequantize(self._h, C.c_void_p(out.data_ptr()),
                                                    C.c_void_p(_stream(self.device))))
        return out

    def export_gguf_bytes(self) -> torch.Tensor:
        """The original GGUF Q4_0 bytes reconstructed from the device layout (uint8 tensor)."""
        out = torch.empty(self.num_blocks() * 18, dtype=torch.uint8, device=self.device)
        _lib.check(_lib.lib().wq4_tensor_export_gguf(self._h, C.c_void_p(out.data_ptr()),
                                                     C.c_void_p(_stream(self.device))))
        return out

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None and h.value:
            try:
                _lib.lib().wq4_tensor_destroy(h)
            except Exception:
                pass


def _linear_ex(weights: Q4Tensor, bias, x: torch.Tensor, flags: int, residual=None) -> torch.Tensor:
    x = _check_f32_cuda(x, "input", weights.device)
    if x.dim() != 3:
        raise _lib.Wq4Error(2, "Input must be 3D [B, M, K]")          # op.rs:53
    b, m, k = x.shape
    n, wk = weights.shape()
    if k != wk:
        raise _lib.Wq4Error(2, f"K dimension mismatch: input has {k}, weights have {wk}")  # op.rs:58-61
    L = _lib.lib()
    dev = weights.device
    out = torch.empty((b, m, n), dtype=torch.float32, device=dev)
    nbytes = int(L.wq4_linear_workspace_bytes(weights._h, b, m))
    ws = workspace(dev, nbytes)
    bp = None
    if bias is not None:
        bias = _check_f32_cuda(bias, "bias", dev)
        bp = C.c_void_p(bias.data_ptr())
    rp = None
    if residual is not None:
        residual = _check_f32_cuda(residual, "residual", dev)
        rp = C.c_void_p(residual.data_ptr())
    _lib.check(L.wq4_linear_forward_ex(weights._h, bp, C.c_void_p(x.data_ptr()), None,
                                       C.c_void_p(out.data_ptr()), None, rp, b * m, flags,
                                       C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(_stream(dev))))
    return out


def q4_matmul(input: torch.Tensor, weights: Q4Tensor, flags: int = 0) -> torch.Tensor:
    """output[B,M,N] = input[B,M,K] x weights[N,K]^T (src/gguf/op.rs:47-116)."""
    return _linear_ex(weights, None, input, flags)


class Q4Linear:
    """x @ W^T + bias with Q4_0 weights (src/gguf/linear.rs:17-40); the bias add is fused."""

    def __init__(self, weights: Q4Tensor, bias: torch.Tensor | None):
        self.weights = weights
        self.bias = None if bias is None else _check_f32_cuda(bias, "bias")

    def forward(self, x: torch.Tensor, flags: int = 0) -> torch.Tensor:
        return _linear_ex(self.weights, self.bias, x, flags)
