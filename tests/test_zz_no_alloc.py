"""wq4_ffn_forward_static (no counterpart in the reference; include/wq4.h): the allocation-free form of Q4FFN::forward for
<= 8 rows, whose hidden block lives in a static __device__ array of the library.  Checked through the C ABI against the
existing wq4_ffn_forward on two calls in a row (the shared scratch keeps the first call's contents).  Collected last on
purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import whisper_burn_b200 as wb
from oracle import q4

DEV = "cuda:0"
REL_TOL = 1e-3   # the bar of tests/test_gpu_parity.py: max|y - y_ref| / max|y_ref|


def _ffn(d, f, seed):
    rng = np.random.default_rng(seed)
    w1 = q4.quantize_q4_0_converter((rng.standard_normal(f * d) * 0.05).astype(np.float32))
    w2 = q4.quantize_q4_0_converter((rng.standard_normal(d * f) * 0.05).astype(np.float32))
    b1 = torch.from_numpy(rng.standard_normal(f).astype(np.float32) * 0.1).to(DEV)
    b2 = torch.from_numpy(rng.standard_normal(d).astype(np.float32) * 0.1).to(DEV)
    fc1 = wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(w1, [f, d], DEV), b1)
    fc2 = wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(w2, [d, f], DEV), b2)
    return wb.Q4FFN(fc1, fc2), (w1, w2, b1.cpu().numpy(), b2.cpu().numpy())


def _rel(a, b):
    return float((a - b).abs().max() / b.abs().max())


def test_ffn_static_matches_ffn_forward_two_calls_in_a_row():
    d, f = 1280, 5120
    ffn, (w1, w2, b1, b2) = _ffn(d, f, 11)
    rng = np.random.default_rng(12)
    # first call: 8 rows fills the whole scratch block; second call: 1 row over the stale contents
    for call, rows in enumerate((8, 1)):
        x = torch.from_numpy(rng.standard_normal((1, rows, d)).astype(np.float32)).to(DEV)
        want = ffn.forward(x)                       # existing entry point (caller workspace)
        out = torch.full((1, rows, d), float("nan"), dtype=torch.float32, device=DEV)
        got = ffn.forward_static(x, out)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        rel = _rel(got, want)
        print(f"call {call}: rows={rows} static vs wq4_ffn_forward rel {rel:.3e}")
        assert rel < REL_TOL, (call, rows, rel)
        if rows == 1:
            # and against the CPU oracle, like the other FFN parity tests
            ref = torch.from_numpy(q4.q4_ffn(x.cpu().numpy(), w1, b1, w2, b2, f))
            assert _rel(got.cpu(), ref) < REL_TOL
    # a repeat of the same input is bit-identical (fixed-order reductions, scratch fully overwritten before it is read)
    again = ffn.forward_static(x, torch.empty_like(out))
    torch.cuda.synchronize()
    assert torch.equal(again, got)
    assert wb.lib().wq4_device_error_flag(0) == 0


def test_ffn_static_rejects_sizes_beyond_the_scratch():
    L = wb.lib()
    mr, mh = C.c_int64(), C.c_int64()
    L.wq4_ffn_static_limits(C.byref(mr), C.byref(mh))
    assert (mr.value, mh.value) == (8, 8192)
    ffn, _ = _ffn(64, 128, 13)
    x = torch.zeros((1, mr.value + 1, 64), dtype=torch.float32, device=DEV)
    with pytest.raises(wb.Wq4Error) as e:
        ffn.forward_static(x, torch.empty((1, mr.value + 1, 64), dtype=torch.float32, device=DEV))
    assert e.value.status == 6 and "beyond the static scratch" in e.value.message
    wide, _ = _ffn(32, mh.value + 32, 14)
    x = torch.zeros((1, 1, 32), dtype=torch.float32, device=DEV)
    with pytest.raises(wb.Wq4Error) as e:
        wide.forward_static(x, torch.empty((1, 1, 32), dtype=torch.float32, device=DEV))
    assert e.value.status == 6
