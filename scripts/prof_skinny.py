#!/usr/bin/env python3
"""A few skinny-linear launches for an ncu capture (rows, N, K from argv; default 32 1280 1280)."""
import ctypes as C
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from microbench import DEV, L, make_weights
from whisper_burn_b200.gguf import _stream

rows, n, k = (int(v) for v in (sys.argv[1:4] if len(sys.argv) >= 4 else (32, 1280, 1280)))
ws = make_weights(n, k, 4)
x16 = torch.randn((rows, k), device=DEV).half()
y = torch.zeros((rows, n), device=DEV)
bias = torch.randn(n, device=DEV)
p = lambda t: C.c_void_p(t.data_ptr())
for i in range(8):
    wb._lib.check(L.wq4_linear_forward_ex(ws[i % 4]._h, p(bias), None, p(x16), p(y), None, None, rows, wb.WQ4_FORCE_SKINNY, None, 0,
                                          C.c_void_p(_stream(DEV))))
torch.cuda.synchronize()
print("ok", L.wq4_device_error_flag(0))
