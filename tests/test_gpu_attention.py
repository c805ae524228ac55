"""Encoder self-attention kernel (tcgen05, wq4_attn.cu) against a torch fp32 SDPA of the same f16 inputs
(src/model/attention.rs:30-58: softmax(q k^T / sqrt(64)) v per head, no mask)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

import whisper_burn_b200 as wb

DEV = "cuda:0"


def run_attention(qkv, B, T, D, H):
    L = wb.lib()
    L.wq4_debug_enc_attention.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    L.wq4_debug_enc_attention.restype = C.c_int32
    out = torch.full((B * T, D), float("nan"), device=DEV, dtype=torch.float16)
    st = L.wq4_debug_enc_attention(C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), B, T, D, H,
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert st == 0, wb.last_error() if hasattr(wb, "last_error") else st
    torch.cuda.synchronize()
    assert L.wq4_device_error_flag(0) == 0
    return out


def reference(qkv, B, T, D, H):
    x = qkv.float().view(B, T, 3, H, 64)
    q, k, v = (x[:, :, i].permute(0, 2, 1, 3) for i in range(3))          # [B,H,T,64]
    o = torch.nn.functional.scaled_dot_product_attention(q, k, v)
    return o.permute(0, 2, 1, 3).reshape(B * T, D)


@pytest.mark.parametrize("B,T,H", [(2, 1500, 4), (1, 32, 2), (1, 128, 1), (1, 129, 2), (1, 1, 1), (3, 300, 20), (1, 257, 3),
                                   (1, 1500, 16), (1, 1500, 20), (1, 193, 2)])
@pytest.mark.parametrize("spread", [1.0, 4.0])
def test_attention_matches_sdpa(B, T, H, spread):
    D = H * 64
    g = torch.Generator(device="cpu").manual_seed(B * 1000 + T + H)
    qkv = (torch.randn((B * T, 3 * D), generator=g) * spread).to(torch.float16).to(DEV)
    got = run_attention(qkv, B, T, D, H).float()
    want = reference(qkv, B, T, D, H)
    assert torch.isfinite(got).all()
    err = float((got - want).abs().max() / want.abs().max())
    print(f"B{B} T{T} H{H} spread {spread}: rel err {err:.2e}")
    assert err < 2e-3          # f16 P and f16 output rounding (2^-11) dominate


def test_attention_rows_are_convex_combinations():
    """Size-independent property at the full L-V3 shape: with v == 1 everywhere every output must be exactly 1."""
    B, T, H = 2, 1500, 20
    D = H * 64
    qkv = torch.randn((B * T, 3 * D), device=DEV).to(torch.float16)
    qkv[:, 2 * D:] = 1.0
    got = run_attention(qkv, B, T, D, H).float()
    assert float((got - 1.0).abs().max()) < 2e-3

