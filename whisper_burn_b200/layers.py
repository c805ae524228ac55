"""Host-side mirror of the reference's src/model/layers.rs (Q4FFN, Conv1D) over the C ABI."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .gguf import Q4Linear, _check_f32_cuda, _as_device, _stream, workspace


class Q4FFN:
    """fc1 -> tanh-GELU -> fc2 (src/model/layers.rs:44-59), one fused call."""

    def __init__(self, fc1: Q4Linear, fc2: Q4Linear):
        self.fc1, self.fc2 = fc1, fc2

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = _check_f32_cuda(x, "input")
        if x.dim() != 3:
            raise _lib.Wq4Error(2, "Input must be 3D [B, M, K]")
        b, m, k = x.shape
        f, k1 = self.fc1.weights.shape()
        d, k2 = self.fc2.weights.shape()
        if k != k1:
            raise _lib.Wq4Error(2, f"K dimension mismatch: input has {k}, weights have {k1}")
        L = _lib.lib()
        dev = self.fc1.weights.device
        out = torch.empty((b, m, d), dtype=torch.float32, device=dev)
        ws = workspace(dev, int(L.wq4_ffn_workspace_bytes(self.fc1.weights._h, self.fc2.weights._h, b, m)))
        p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        _lib.check(L.wq4_ffn_forward(self.fc1.weights._h, p(self.fc1.bias), self.fc2.weights._h, p(self.fc2.bias),
                                     C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), b, m,
                                     C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(_stream(dev))))
        return out

    def forward_static(self, x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """wq4_ffn_forward_static: <= 8 rows, hidden block in the library's static device array; the caller passes the
        output buffer and nothing is allocated (no counterpart in the reference; include/wq4.h)."""
        x = _check_f32_cuda(x, "input")
        if x.dim() != 3:
            raise _lib.Wq4Error(2, "Input must be 3D [B, M, K]")
        b, m, k = x.shape
        k1 = self.fc1.weights.shape()[1]
        d = self.fc2.weights.shape()[0]
        if k != k1:
            raise _lib.Wq4Error(2, f"K dimension mismatch: input has {k}, weights have {k1}")
        if out.dtype != torch.float32 or not out.is_contiguous() or tuple(out.shape) != (b, m, d) or out.device != x.device:
            raise _lib.Wq4Error(1, f"output must be a contiguous f32 [{b}, {m}, {d}] tensor on {x.device}")
        L = _lib.lib()
        p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        _lib.check(L.wq4_ffn_forward_static(self.fc1.weights._h, p(self.fc1.bias), self.fc2.weights._h, p(self.fc2.bias),
                                            C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), b, m,
                                            C.c_void_p(_stream(self.fc1.weights.device))))
        return out


class Conv1D:
    """1-D convolution as an implicit-im2col tensor-core GEMM (src/model/layers.rs:61-133).
    weight [out_channels, in_channels, kernel_size] f32, bias [out_channels] f32 (host or device)."""

    def __init__(self, weight, bias, device="cuda"):
        w = weight.detach().cpu().numpy() if isinstance(weight, torch.Tensor) else np.asarray(weight)
        b = bias.detach().cpu().numpy() if isinstance(bias, torch.Tensor) else np.asarray(bias)
        w = np.ascontiguousarray(w, dtype=np.float32)
        b = np.ascontiguousarray(b, dtype=np.float32)
        if w.ndim != 3 or b.shape != (w.shape[0],):
            raise _lib.Wq4Error(2, "Conv1D weight must be [out, in, k] and bias [out]")
        if isinstance(weight, torch.Tensor) and weight.is_cuda:
            device = weight.device
        self.device = _as_device(device)
        self.out_channels, self.in_channels, self.kernel_size = (int(s) for s in w.shape)
        h = C.c_void_p()
        _lib.check(_lib.lib().wq4_conv1d_create(w.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p),
                                                self.out_channels, self.in_channels, self.kernel_size,
                                                self.device.index, C.c_void_p(_stream(self.device)), C.byref(h)))
        self._h = h

    def forward(self, x: torch.Tensor, padding: int, stride: int, flags: int = 0) -> torch.Tensor:
        x = _check_f32_cuda(x, "input")
        if x.dim() != 3 or x.shape[1] != self.in_channels:
            raise _lib.Wq4Error(2, f"Conv1D input must be [B, {self.in_channels}, T]")
        b, _, t = x.shape
        t_out = (t + 2 * padding - self.kernel_size) // stride + 1   # layers.rs:80
        L = _lib.lib()
        out = torch.empty((b, self.out_channels, t_out), dtype=torch.float32, device=self.device)
        ws = workspace(self.device, int(L.wq4_conv1d_workspace_bytes(self._h, b, t, padding, stride)))
        _lib.check(L.wq4_conv1d_forward(self._h, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), b, t,
                                        padding, stride, flags, C.c_void_p(ws.data_ptr()), ws.numel(),
                                        C.c_void_p(_stream(self.device))))
        return out

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None and h.value:
            try:
                _lib.lib().wq4_conv1d_destroy(h)
            except Exception:
                pass
