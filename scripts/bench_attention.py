#!/usr/bin/env python3
"""Encoder attention kernel alone (wq4_debug_enc_attention): CUDA-graph replay, device time per call.
Set WQ4_ATTN_IMPL=1 for the mma.sync kernel it replaced."""
import ctypes as C
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb

DEV = torch.device("cuda:0")
L = wb.lib()
L.wq4_debug_enc_attention.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
L.wq4_debug_enc_attention.restype = C.c_int32

for name, B, T, H in (("medium b1", 1, 1500, 16), ("large-v3 b1", 1, 1500, 20), ("large-v3 b8", 8, 1500, 20)):
    D = H * 64
    qkv = torch.randn((B * T, 3 * D), device=DEV).to(torch.float16)
    out = torch.empty((B * T, D), device=DEV, dtype=torch.float16)
    s = torch.cuda.Stream()
    reps = 20
    with torch.cuda.stream(s):
        run = lambda: wb._lib.check(L.wq4_debug_enc_attention(C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), B, T, D, H,
                                                               C.c_void_p(s.cuda_stream)))
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                run()
        g.replay()
        torch.cuda.synchronize()
        best = 1e9
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(5):
            e0.record(s)
            g.replay()
            e1.record(s)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / reps * 1e-3)
    flop = 4.0 * B * H * T * T * 64
    exps = B * H * T * T
    print(f"{name:12s}: {best * 1e6:8.1f} us  {flop / best / 1e12:7.1f} TFLOP/s  {exps / best / 1e9:7.1f} Gexp/s "
          f"(148 SMs x 16/clk x 1.965 GHz = 4653)", flush=True)
print("error flag", L.wq4_device_error_flag(0))
