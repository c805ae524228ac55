#!/usr/bin/env python3
"""First-contact diagnostics on a B200: runs each kernel family once with verbose error reports
so that a wrong descriptor / layout shows up as a pattern, not just a failed assert."""
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import whisper_burn_b200 as wb
from oracle import q4

DEV = "cuda:0"
L = wb.lib()
print("devices", L.wq4_device_count(), torch.cuda.get_device_name(0))
q4.set_threads(q4.host_cores())


def report(name, got, want):
    err = np.abs(got - want)
    rel = err.max() / max(np.abs(want).max(), 1e-30)
    flag = L.wq4_device_error_flag(0)
    print(f"{name:44s} rel {rel:.3e} absmax {err.max():.3e} nan {int(np.isnan(got).sum())} errflag {flag}", flush=True)
    if rel > 1e-3 or flag:
        bad = np.argwhere(err > 1e-3 * np.abs(want).max())
        print("   first bad idx:", bad[:6].tolist(), "count", len(bad), "of", err.size)
        r0 = bad[0] if len(bad) else (0,) * got.ndim
        sl = tuple(int(v) for v in r0[:-1])
        print("   got ", got[sl][:8], "\n   want", want[sl][:8])
    return rel


rng = np.random.default_rng(0)
n, k = 1280, 1280
w = (rng.standard_normal(n * k) * 0.02).astype(np.float32)
qb = q4.quantize_q4_0_converter(w)
qt = wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV)
deq = qt.dequantize().cpu().numpy().ravel()
want = q4.dequantize_q4_0(qb, n * k)
print("dequant bit-exact:", np.array_equal(deq.view(np.uint32), want.view(np.uint32)),
      " export exact:", np.array_equal(qt.export_gguf_bytes().cpu().numpy(), qb), flush=True)

for rows in (1, 2, 4, 8):
    x = rng.standard_normal((1, rows, k)).astype(np.float32)
    got = wb.q4_matmul(torch.from_numpy(x).to(DEV), qt, flags=wb.WQ4_FORCE_GEMV).cpu().numpy()
    report(f"gemv rows={rows} 1280x1280", got, q4.q4_matmul(x, qb, n))

# GEMM: start with structured inputs that localise layout bugs
for rows, tag in ((128, "ones"), (128, "rand"), (32, "rand"), (256, "rand"), (1500, "rand")):
    x = np.ones((1, rows, k), np.float32) if tag == "ones" else rng.standard_normal((1, rows, k)).astype(np.float32)
    torch.cuda.synchronize()
    t0 = time.time()
    got = wb.q4_matmul(torch.from_numpy(x).to(DEV), qt, flags=wb.WQ4_FORCE_GEMM).cpu().numpy()
    dt = time.time() - t0
    report(f"gemm rows={rows} {tag} 1280x1280 ({dt*1e3:.1f} ms)", got, q4.q4_matmul(x, qb, n))

for tile in (32, 64, 128, 256):
    import os
    # tile override is read once per process; exercised separately by the test-suite via env
    pass

# conv
x = rng.standard_normal((1, 80, 3000)).astype(np.float32)
wc = (rng.standard_normal((1024, 80, 3)) * 0.06).astype(np.float32)
bc = rng.standard_normal(1024).astype(np.float32) * 0.1
conv = wb.Conv1D(wc, bc, device=DEV)
got = conv.forward(torch.from_numpy(x).to(DEV), 1, 1).cpu().numpy()
report("conv1 80->1024 T=3000 s1", got, q4.conv1d(x, wc, bc, 1, 1))
x2 = rng.standard_normal((1, 256, 600)).astype(np.float32)
wc2 = (rng.standard_normal((256, 256, 3)) * 0.04).astype(np.float32)
bc2 = rng.standard_normal(256).astype(np.float32) * 0.1
conv2 = wb.Conv1D(wc2, bc2, device=DEV)
got = conv2.forward(torch.from_numpy(x2).to(DEV), 1, 2).cpu().numpy()
report("conv2 256->256 T=600 s2", got, q4.conv1d(x2, wc2, bc2, 1, 2))
torch.cuda.synchronize()
print("bringup done; error flag", L.wq4_device_error_flag(0))

# ---- quick timings (CUDA events on the current stream) -------------------------------------------
def time_op(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


import ctypes as C
from whisper_burn_b200.gguf import workspace, _stream

tensors = {}
for (n_, k_) in ((1280, 1280), (5120, 1280), (1280, 5120), (1024, 1024), (4096, 1024), (1024, 4096)):
    w_ = (rng.standard_normal(n_ * k_) * 0.02).astype(np.float32)
    tensors[(n_, k_)] = wb.Q4Tensor.from_q4_bytes(q4.quantize_q4_0_converter(w_), [n_, k_], DEV)

print("\n-- GEMM (f16 activations resident, kernel only) --")
dev = torch.device(DEV)
for rows in (1500, 12000, 48000):
    for (n_, k_), qt_ in tensors.items():
        x16 = torch.randn((rows, k_), device=DEV, dtype=torch.float16)
        y = torch.empty((rows, n_), device=DEV, dtype=torch.float32)
        def run():
            wb._lib.check(L.wq4_linear_forward_ex(qt_._h, None, None, C.c_void_p(x16.data_ptr()), C.c_void_p(y.data_ptr()),
                                                  None, None, rows, 0, None, 0, C.c_void_p(_stream(dev))))
        dt = time_op(run, iters=20)
        tile = C.c_int32()
        L.wq4_linear_plan(qt_._h, rows, wb.WQ4_FORCE_GEMM, C.byref(tile))
        print(f"rows {rows:6d} N {n_:5d} K {k_:5d} tile_m {tile.value:3d}: {dt*1e6:9.1f} us  {2*rows*n_*k_/dt/1e12:8.1f} TFLOP/s", flush=True)

print("\n-- GEMV (M=1, f32) --")
for (n_, k_), qt_ in tensors.items():
    x = torch.randn((1, 1, k_), device=DEV)
    bias = torch.randn(n_, device=DEV)
    lin = wb.Q4Linear(qt_, bias)
    dt = time_op(lambda: lin.forward(x), iters=50)
    by = n_ * k_ // 32 * 18 + 4 * k_ + 8 * n_
    print(f"N {n_:5d} K {k_:5d}: {dt*1e6:8.2f} us  {by/dt/1e9:8.1f} GB/s (incl. launch + torch.empty)", flush=True)
print("error flag", L.wq4_device_error_flag(0))
