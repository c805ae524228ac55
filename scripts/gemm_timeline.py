#!/usr/bin/env python3
"""Per-stage clock64 stamps of CTA 0 of one Q4 GEMM launch (wq4_debug_gemm_timeline hook).
Needs a library built with the stamps compiled in:  bash whisper_burn_b200/csrc/build.sh -DWQ4_GEMM_TIMELINE
Usage: gemm_timeline.py rows N K"""
import ctypes as C, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent)); sys.path.insert(0, str(Path(__file__).resolve().parent))
import whisper_burn_b200 as wb
import microbench as mb
rows, n, k = (int(v) for v in sys.argv[1:4]); mode = sys.argv[4] if len(sys.argv) > 4 else "f16"
L = wb.lib(); L.wq4_debug_gemm_timeline.argtypes = [C.c_void_p]; L.wq4_debug_gemm_timeline.restype = None
ws = mb.make_weights(n, k, 2)
x16 = torch.randn((rows, k), device=mb.DEV, dtype=torch.float16); y16 = torch.empty((rows, n), device=mb.DEV, dtype=torch.float16)
bias = torch.randn(n, device=mb.DEV)
def run(w):
    wb._lib.check(L.wq4_linear_forward_ex(w._h, C.c_void_p(bias.data_ptr()), None, C.c_void_p(x16.data_ptr()), None, C.c_void_p(y16.data_ptr()), None, rows, wb.WQ4_FORCE_GEMM, None, 0, None))
for _ in range(3): run(ws[0])
torch.cuda.synchronize()
tl = torch.zeros(4 * 64 * 4, dtype=torch.int64, device=mb.DEV)
L.wq4_debug_gemm_timeline(C.c_void_p(tl.data_ptr())); run(ws[1]); torch.cuda.synchronize(); L.wq4_debug_gemm_timeline(None)
t = tl.cpu().numpy().reshape(4, 64, 4); t0 = t[t > 0].min()
nst = int((t[1, :, 1] > 0).sum())
print(f"rows {rows} N {n} K {k}: {nst} stages recorded for CTA 0; clocks relative to first stamp")
print("stage | X: empty-wait done | MMA: wait-start full-seen issued | DQ: loop-top raw-seen math-done slot-free st-done")
r = lambda v: int(v - t0) if v > 0 else -1
for c in range(min(nst, 40)):
    print(f"{c:3d} | {r(t[0,c,0]):6d} | {r(t[1,c,0]):6d} {r(t[1,c,1]):6d} {r(t[1,c,2]):6d} | {r(t[2,c,0]):6d} {r(t[2,c,1]):6d} {r(t[2,c,2]):6d} {r(t[2,c,3]):6d} {r(t[3,c,0]):6d}")
f = t[1, :nst, 1]
if nst > 8: print("MMA full-seen period (clk/stage), stages 8..end:", round(float(np.diff(f[8:]).mean()), 1))
