#!/usr/bin/env python3
"""Runs one GEMM shape a few times (target of `ncu`)."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from whisper_burn_b200.gguf import _stream

rows, n, k, iters = (int(v) for v in (sys.argv[1:5] + ["12000", "1280", "1280", "6"][len(sys.argv) - 1:]))
DEV = torch.device("cuda:0")
L = wb.lib()
rng = np.random.default_rng(0)
raw = rng.integers(0, 256, n * k // 32 * 18, dtype=np.uint8).reshape(-1, 18)
raw[:, 0:2] = np.frombuffer(np.float16(rng.uniform(0.001, 0.01, raw.shape[0])).tobytes(), np.uint8).reshape(-1, 2)
w = wb.Q4Tensor.from_q4_bytes(raw.ravel(), [n, k], DEV)
x16 = torch.randn((rows, k), device=DEV, dtype=torch.float16)
y = torch.empty((rows, n), device=DEV, dtype=torch.float32)
for _ in range(iters):
    wb._lib.check(L.wq4_linear_forward_ex(w._h, None, None, C.c_void_p(x16.data_ptr()), C.c_void_p(y.data_ptr()),
                                          None, None, rows, wb.WQ4_FORCE_GEMM, None, 0, C.c_void_p(_stream(DEV))))
torch.cuda.synchronize()
print("ok", float(y.abs().mean()), "errflag", L.wq4_device_error_flag(0))
