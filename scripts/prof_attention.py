#!/usr/bin/env python3
"""Runs the encoder attention kernel a few times on one shape (target of `ncu`): prof_attention.py B T H iters"""
import ctypes as C
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb

B, T, H, iters = (int(v) for v in (sys.argv[1:5] + ["8", "1500", "20", "5"][len(sys.argv) - 1:]))
D = H * 64
L = wb.lib()
L.wq4_debug_enc_attention.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
L.wq4_debug_enc_attention.restype = C.c_int32
qkv = torch.randn((B * T, 3 * D), device="cuda:0").to(torch.float16)
out = torch.empty((B * T, D), device="cuda:0", dtype=torch.float16)
for _ in range(iters):
    wb._lib.check(L.wq4_debug_enc_attention(C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), B, T, D, H, None))
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()), "errflag", L.wq4_device_error_flag(0))
