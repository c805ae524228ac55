"""The skinny Q4 linear (csrc/wq4_skinny.cu: 1..64 activation rows, mma.sync with the weight rows as the MMA M dimension)
through the C ABI, every row-tile count, both activation types and every epilogue mode, against the f64 product of the
oracle's dequantised weights (oracle/q4.py, tensor.rs:96-109 twin) with the f16-rounded activations; tolerance 1e-3 of the
output scale (src/gguf/tests.rs:406).  Also: default dispatch for 9..16 rows, and run-to-run determinism."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import whisper_burn_b200 as wb
from whisper_burn_b200.gguf import _stream
from oracle import q4

DEV = torch.device("cuda:0")
GELU = lambda h: 0.5 * h * (1 + np.tanh(0.7978845608028654 * (h + 0.044715 * h ** 3)))


def _p(a):
    return None if a is None else C.c_void_p(a.data_ptr())


def _call(t, bias, x32, x16, y32, y16, res, rows, flags):
    ws = torch.empty(rows * t.shape()[1] * 2 + 256, dtype=torch.uint8, device=DEV)     # f16 copy of an f32 x
    wb._lib.check(wb.lib().wq4_linear_forward_ex(t._h, _p(bias), _p(x32), _p(x16), _p(y32), _p(y16), _p(res), rows, flags,
                                                 _p(ws), ws.numel(), C.c_void_p(_stream(DEV))))
    torch.cuda.synchronize()
    assert wb.lib().wq4_device_error_flag(0) == 0


# (N, K): Large-V3 shapes, N not a multiple of 16 / of 128, K = one block, K not a multiple of 8 blocks
SHAPES = [(1280, 1280), (5120, 1280), (1280, 5120), (200, 96), (40, 32), (3840, 1280), (136, 2080)]


@pytest.mark.parametrize("rows", [1, 8, 9, 15, 16, 17, 31, 32, 33, 48, 64])
@pytest.mark.parametrize("n,k", SHAPES)
def test_skinny_all_epilogues(rows, n, k):
    if rows not in (9, 32, 64) and (n, k) in ((5120, 1280), (1280, 5120), (3840, 1280)):
        pytest.skip("large shapes on three row counts")
    rng = np.random.default_rng(rows * 131 + n + k)
    qb = q4.quantize_q4_0_converter((rng.standard_normal(n * k) * 0.05).astype(np.float32))
    wd = q4.dequantize_q4_0(qb, n * k).reshape(n, k).astype(np.float64)
    t = wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV)
    x = rng.standard_normal((rows, k)).astype(np.float32)
    x32 = torch.from_numpy(x).to(DEV)
    x16 = x32.to(torch.float16)
    base = x16.float().cpu().numpy().astype(np.float64) @ wd.T
    bias = rng.standard_normal(n).astype(np.float32)
    res = rng.standard_normal((rows, n)).astype(np.float32)
    bias_d = torch.from_numpy(bias).to(DEV)
    F = wb.WQ4_FORCE_SKINNY
    for xkind in ("f16", "f32"):
        xa = (None, x16) if xkind == "f16" else (x32, None)
        # plain f32 output, no bias
        y32 = torch.full((rows, n), 7.0, dtype=torch.float32, device=DEV)
        _call(t, None, *xa, y32, None, None, rows, F)
        scale = np.abs(base).max()
        assert np.abs(y32.cpu().numpy() - base).max() / scale < 1e-3, (xkind, "f32")
        # bias + GELU -> f16 output only
        y16 = torch.zeros((rows, n), dtype=torch.float16, device=DEV)
        _call(t, bias_d, *xa, None, y16, None, rows, F | wb.WQ4_EPI_GELU)
        want = GELU(base + bias)
        assert np.abs(y16.float().cpu().numpy() - want).max() / np.abs(want).max() < 2e-3, (xkind, "gelu16")
        # bias + residual in place, both outputs
        y32 = torch.from_numpy(res.copy()).to(DEV)
        _call(t, bias_d, *xa, y32, y16, y32, rows, F)
        want = base + bias + res
        assert np.abs(y32.cpu().numpy() - want).max() / np.abs(want).max() < 1e-3, (xkind, "res")
        assert np.abs(y16.float().cpu().numpy() - want).max() / np.abs(want).max() < 2e-3, (xkind, "res16")


def test_default_dispatch_and_determinism():
    n, k = 1280, 1280
    rng = np.random.default_rng(1)
    qb = q4.quantize_q4_0_converter((rng.standard_normal(n * k) * 0.02).astype(np.float32))
    t = wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV)
    L = wb.lib()
    tile = C.c_int32()
    assert L.wq4_linear_plan(t._h, 2, 0, C.byref(tile)) == 0            # 1-2 rows: f32 GEMV
    assert L.wq4_linear_plan(t._h, 8, wb.WQ4_FORCE_GEMV, C.byref(tile)) == 0
    assert L.wq4_linear_plan(t._h, 3, 0, C.byref(tile)) == 2 and tile.value == 16   # 3..16 rows: skinny kernel
    assert L.wq4_linear_plan(t._h, 9, 0, C.byref(tile)) == 2 and tile.value == 16
    assert L.wq4_linear_plan(t._h, 16, 0, C.byref(tile)) == 2
    assert L.wq4_linear_plan(t._h, 17, 0, C.byref(tile)) == 1          # beyond 16 rows the tcgen05 GEMM is faster
    assert L.wq4_linear_plan(t._h, 64, wb.WQ4_FORCE_SKINNY, C.byref(tile)) == 2
    assert L.wq4_linear_plan(t._h, 65, wb.WQ4_FORCE_SKINNY, C.byref(tile)) == 1
    assert L.wq4_linear_plan(t._h, 12, wb.WQ4_FORCE_GEMM, C.byref(tile)) == 1
    x = torch.from_numpy(rng.standard_normal((1, 14, k)).astype(np.float32)).to(DEV)
    bias = torch.from_numpy(rng.standard_normal(n).astype(np.float32)).to(DEV)
    lin = wb.Q4Linear(t, bias)
    a = lin.forward(x)
    b = lin.forward(x)
    assert torch.equal(a, b)                                    # fixed-order K reduction: bit-identical run to run
    g = lin.forward(x, flags=wb.WQ4_FORCE_GEMM)                 # same operand precision as the tcgen05 kernel
    assert float((a - g).abs().max() / g.abs().max()) < 2e-4
    q4.set_threads(q4.host_cores())
    want = q4.q4_linear(x.cpu().numpy(), qb, bias.cpu().numpy(), n)
    assert float(np.abs(a.cpu().numpy() - want).max() / np.abs(want).max()) < 1e-3


def test_skinny_rejects_more_than_64_rows():
    n, k = 64, 64
    qb = q4.quantize_q4_0_converter(np.ones(n * k, dtype=np.float32))
    t = wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV)
    x = torch.zeros((1, 65, k), dtype=torch.float32, device=DEV)
    # FORCE_SKINNY is ignored beyond its row limit: the call falls through to the tensor-core GEMM and succeeds
    y = wb.q4_matmul(x, t, flags=wb.WQ4_FORCE_SKINNY)
    assert y.shape == (1, 65, n) and float(y.abs().max()) == 0.0
