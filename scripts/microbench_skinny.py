#!/usr/bin/env python3
"""Skinny Q4 linear (csrc/wq4_skinny.cu) against the tcgen05 GEMM and the GEMV at decode-batch row counts: CUDA-graph replays
over weight copies that exceed L2 (scripts/microbench.py harness).  Bytes = Q4 weight bytes + x + y."""
import ctypes as C
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from microbench import DEV, L, graph_time, make_weights
from whisper_burn_b200.gguf import _stream


def run_one(rows, n, k, flags, xkind, reps=48):
    copies = max(2, min(reps, int(300e6 // (n * k * 0.5625)) + 1))
    ws = make_weights(n, k, copies)
    x32 = torch.randn((rows, k), device=DEV)
    x16 = x32.half()
    y = torch.zeros((rows, n), device=DEV)
    bias = torch.randn(n, device=DEV)
    wsb = torch.empty(rows * k * 2 + 256, dtype=torch.uint8, device=DEV)
    p = lambda t: C.c_void_p(t.data_ptr())

    def run(i):
        w = ws[i % copies]
        xa = (p(x32), None) if xkind == "f32" else (None, p(x16))
        wb._lib.check(L.wq4_linear_forward_ex(w._h, p(bias), *xa, p(y), None, None, rows, flags, p(wsb), wsb.numel(),
                                              C.c_void_p(_stream(DEV))))
    return graph_time(run, reps)


if __name__ == "__main__":
    peak = 6455.0
    print("rows  N     K     kernel            us      GB/s (Q4 bytes + x + y)")
    for n, k in ((1280, 1280), (3840, 1280), (5120, 1280), (1280, 5120)):
        for rows in ([int(v) for v in sys.argv[1].split(',')] if len(sys.argv) > 1 else (4, 8, 16, 32, 64)):
            byts = n * k // 32 * 18 + rows * k * 2 + rows * n * 4 + n * 4
            arms = [("skinny f16x", wb.WQ4_FORCE_SKINNY, "f16"), ("tcgen05 gemm", wb.WQ4_FORCE_GEMM, "f16")]
            if rows <= 8:
                arms.append(("gemv f32", wb.WQ4_FORCE_GEMV, "f32"))
            else:
                arms.append(("skinny f32x", wb.WQ4_FORCE_SKINNY, "f32"))
            if len(sys.argv) > 2:
                arms = arms[:1]
            for name, fl, xk in arms:
                dt = run_one(rows, n, k, fl, xk)
                print(f"{rows:4d} {n:5d} {k:5d}  {name:14s} {dt * 1e6:8.2f}  {byts / dt / 1e9:8.1f}  ({100 * byts / dt / 1e9 / peak:4.1f}% of {peak:.0f})", flush=True)
    print("error flag", L.wq4_device_error_flag(0))
