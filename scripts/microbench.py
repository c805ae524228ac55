#!/usr/bin/env python3
"""Kernel-only timings with launch overhead removed: each measurement replays a CUDA graph of
`reps` launches that rotate over enough distinct weight copies to exceed the 126 MB L2."""
import argparse
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from whisper_burn_b200.gguf import _stream

DEV = torch.device("cuda:0")
L = wb.lib()


def make_weights(n, k, copies, seed=0):
    rng = np.random.default_rng(seed)
    out = []
    for c in range(copies):
        raw = rng.integers(0, 256, n * k // 32 * 18, dtype=np.uint8).reshape(-1, 18)
        raw[:, 0:2] = np.frombuffer(np.float16(rng.uniform(0.001, 0.01, raw.shape[0])).tobytes(), np.uint8).reshape(-1, 2)
        out.append(wb.Q4Tensor.from_q4_bytes(raw.ravel(), [n, k], DEV))
    return out


def graph_time(fn_of_i, reps, iters=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3):
            fn_of_i(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for i in range(reps):
                fn_of_i(i)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(iters):
            e0.record(s)
            g.replay()
            e1.record(s)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / reps * 1e-3)
    return best


def bench_gemm(rows, n, k, flags=0, reps=20, mode="f32"):
    """mode: f32 (q4_matmul drop-in), f16 (bias -> f16, fused QKV), gelu (bias+GELU -> f16, fc1), res (bias + residual
    -> f32 in place, out-proj / fc2): the four epilogues the encoder uses."""
    copies = max(2, min(reps, int(300e6 // (n * k * 0.5625)) + 1))
    ws = make_weights(n, k, copies)
    x16 = torch.randn((rows, k), device=DEV, dtype=torch.float16)
    y = torch.zeros((rows, n), device=DEV, dtype=torch.float32)
    y16 = torch.empty((rows, n), device=DEV, dtype=torch.float16)
    bias = torch.randn(n, device=DEV)
    p = lambda t: C.c_void_p(t.data_ptr())
    if mode == "gelu":
        flags |= wb.WQ4_EPI_GELU

    def run(i):
        w = ws[i % copies]
        b_, y32_, y16_, res_ = {"f32": (None, p(y), None, None), "f16": (p(bias), None, p(y16), None),
                                "gelu": (p(bias), None, p(y16), None), "res": (p(bias), p(y), None, p(y))}[mode]
        wb._lib.check(L.wq4_linear_forward_ex(w._h, b_, None, p(x16), y32_, y16_, res_, rows, flags | wb.WQ4_FORCE_GEMM, None, 0,
                                              C.c_void_p(_stream(DEV))))
    dt = graph_time(run, reps)
    tile = C.c_int32()
    L.wq4_linear_plan(ws[0]._h, rows, wb.WQ4_FORCE_GEMM, C.byref(tile))
    return dt, tile.value


def bench_gemv(rows, n, k, reps=64):
    copies = max(2, min(reps, int(300e6 // (n * k * 0.5625)) + 1))
    ws = make_weights(n, k, copies)
    x = torch.randn((rows, k), device=DEV)
    bias = torch.randn(n, device=DEV)
    y = torch.empty((rows, n), device=DEV)

    def run(i):
        w = ws[i % copies]
        wb._lib.check(L.wq4_linear_forward_ex(w._h, C.c_void_p(bias.data_ptr()), C.c_void_p(x.data_ptr()), None,
                                              C.c_void_p(y.data_ptr()), None, None, rows, wb.WQ4_FORCE_GEMV, None, 0,
                                              C.c_void_p(_stream(DEV))))
    return graph_time(run, reps)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--rows", default="", help="comma-separated GEMM row counts (default 1500,6000,48000)")
    ap.add_argument("--no-gemv", action="store_true")
    ap.add_argument("--mode", default="f32", choices=["f32", "f16", "gelu", "res"])
    a = ap.parse_args()
    shapes = [(1280, 1280), (5120, 1280), (1280, 5120), (3840, 1280), (1024, 1024), (4096, 1024), (1024, 4096)]
    print("-- GEMM, f16 activations resident, CUDA-graph replay, weights rotate through >L2 --")
    for rows in (tuple(int(r) for r in a.rows.split(',')) if a.rows else (1500,) if a.quick else (1500, 6000, 48000)):
        for n, k in shapes:
            dt, tile = bench_gemm(rows, n, k, reps=20 if rows < 20000 else 6, mode=a.mode)
            print(f"rows {rows:6d} N {n:5d} K {k:5d} tile_m {tile:3d}: {dt*1e6:9.1f} us {2*rows*n*k/dt/1e12:8.1f} TFLOP/s "
                  f"({2*rows*n*k/dt/1e12/1679*100:5.1f}% of measured 1679)", flush=True)
    print("-- GEMV, f32, CUDA-graph replay --")
    for rows in (() if a.no_gemv else (1, 4, 8)):
        for n, k in shapes:
            dt = bench_gemv(rows, n, k)
            by = n * k // 32 * 18 + 4 * k * rows + 4 * n + 4 * n * rows
            print(f"rows {rows} N {n:5d} K {k:5d}: {dt*1e6:8.2f} us {by/dt/1e9:8.1f} GB/s ({by/dt/1e9/6455*100:5.1f}% of measured 6455)", flush=True)
    print("error flag", L.wq4_device_error_flag(0))
