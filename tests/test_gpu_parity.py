"""GPU parity tests: the CUDA path (through the C ABI, via the host mirror of the reference's
operator interface) against the CPU oracle on the same inputs.  Each test cites the reference test
it restates (src/gguf/tests.rs).  Bar: dequantisation bit-exact; matmul within the reference's
absolute tolerances AND within 1e-3 relative error (max|y - y_ref| / max|y_ref|)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import whisper_burn_b200 as wb
from oracle import q4

DEV = "cuda:0"
REL_TOL = 1e-3   # north_star: "within 1e-3 relative error (fp32 accumulate)"


@pytest.fixture(autouse=True)
def _device_error_flag_clean():
    yield
    torch.cuda.synchronize()
    assert wb.lib().wq4_device_error_flag(0) == 0, "a kernel reported a pipeline timeout"


def gen_act(n):   # tests.rs:432-434
    i = np.arange(n, dtype=np.float32)
    return (np.sin(i * np.float32(0.001)) * np.float32(0.1)).astype(np.float32)


def gen_w(n):     # tests.rs:436-438
    i = np.arange(n, dtype=np.float32)
    return (np.cos(i * np.float32(0.0007)) * np.float32(0.05)).astype(np.float32)


def rel_err(got, want):
    return float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))


def t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def test_q4_dequantize_gpu():  # tests.rs:330-363 (16x16: blocks straddle rows)
    rows = cols = 16
    i = np.arange(rows * cols, dtype=np.float32)
    orig = (np.sin(i * np.float32(0.05) - np.float32(6.4)) * np.float32(0.3)).astype(np.float32)
    qb = q4.quantize_q4_0_tests(orig)
    want = q4.dequantize_q4_0(qb, rows * cols)
    qt = wb.Q4Tensor.from_q4_bytes(qb, [rows, cols], DEV)
    got = qt.dequantize()
    assert list(got.shape) == [rows, cols]
    assert np.array_equal(got.cpu().numpy().ravel().view(np.uint32), want.view(np.uint32))


def test_q4_roundtrip_small():  # tests.rs:661-706
    rows, cols = 32, 64
    i = np.arange(rows * cols, dtype=np.float32)
    orig = (np.sin(i * np.float32(0.003) - np.float32(3.0)) * np.float32(0.5)).astype(np.float32)
    qb = q4.quantize_q4_0_tests(orig)
    want = q4.dequantize_q4_0(qb, rows * cols)
    qt = wb.Q4Tensor.from_q4_bytes(qb, [rows, cols], DEV)
    got = qt.dequantize().cpu().numpy().ravel()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.abs(got - orig).max() < 0.1
    assert qt.num_blocks() == rows * cols // 32 and qt.shape() == [rows, cols]
    assert np.array_equal(qt.export_gguf_bytes().cpu().numpy(), qb)


@pytest.mark.parametrize("n,k,seed", [(1280, 1280, 0), (5120, 1280, 1), (200, 96, 2), (1, 32, 3), (384, 4096, 4)])
def test_dequant_bit_exact_random_bytes(n, k, seed):
    """Arbitrary bit patterns (every nibble value, subnormal/huge/negative f16 scales)."""
    rng = np.random.default_rng(seed)
    qb = rng.integers(0, 256, n * k // 32 * 18, dtype=np.uint8)
    sc = qb.reshape(-1, 18)[:, :2].view(np.uint16)
    sc[(sc & 0x7C00) == 0x7C00] = 0x3C00          # keep scales finite (NaN payloads compare unequal)
    want = q4.dequantize_q4_0(qb, n * k)
    qt = wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV)
    got = qt.dequantize().cpu().numpy().ravel()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(qt.export_gguf_bytes().cpu().numpy(), qb)


def test_q4_matmul_small():  # tests.rs:369-410
    k = n = 32
    i = np.arange(n * k, dtype=np.float32)
    qb = q4.quantize_q4_0_tests((np.sin(i * np.float32(0.1)) * np.float32(0.5)).astype(np.float32))
    act = (np.arange(k, dtype=np.float32) * np.float32(0.1)).astype(np.float32)
    expected = q4.reference_matmul(act, q4.dequantize_q4_0(qb, n * k), 1, k, n)
    out = wb.q4_matmul(t(act.reshape(1, 1, k)), wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV))
    assert list(out.shape) == [1, 1, n]
    assert np.abs(out.cpu().numpy().ravel() - expected.ravel()).max() < 1e-3


SHAPES = [  # tests.rs:418-424, 603-655, 712-764, plus M=1500
    (1, 1, 128, 64, 1e-2), (1, 1, 1280, 1280, 1e-2), (1, 10, 1280, 1280, 1e-2),
    (1, 1, 1280, 5120, 1e-2), (1, 1, 5120, 1280, 1e-2), (4, 10, 128, 64, 1e-3),
    (1, 32, 1280, 1280, 1e-2), (1, 1500, 1280, 1280, 1e-2),
]


@pytest.mark.parametrize("batch,seq,k,n,tol", SHAPES)
def test_q4_matmul_shapes(batch, seq, k, n, tol):
    act = gen_act(batch * seq * k).reshape(batch, seq, k)
    qb = q4.quantize_q4_0_tests(gen_w(n * k))
    q4.set_threads(q4.host_cores())
    want = q4.q4_matmul(act, qb, n)
    out = wb.q4_matmul(t(act), wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV)).cpu().numpy()
    assert out.shape == (batch, seq, n)
    assert np.abs(out - want).max() < tol
    assert rel_err(out, want) < REL_TOL


@pytest.mark.parametrize("rows", [1, 2, 3, 4, 5, 7, 8])
@pytest.mark.parametrize("k,n", [(1280, 1280), (1280, 5120), (5120, 1280), (64, 32), (96, 200)])
def test_gemv_family_all_row_counts(rows, k, n):
    rng = np.random.default_rng(rows * 7 + k + n)
    w = (rng.standard_normal(n * k) * 0.02).astype(np.float32)
    qb = q4.quantize_q4_0_converter(w)
    x = rng.standard_normal((1, rows, k)).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    q4.set_threads(q4.host_cores())
    want = q4.q4_linear(x, qb, bias, n)
    lin = wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV), t(bias))
    out = lin.forward(t(x), flags=wb.WQ4_FORCE_GEMV).cpu().numpy()
    assert rel_err(out, want) < 1e-5      # f32 arithmetic: only the summation order differs


@pytest.mark.parametrize("rows,k,n", [(1, 64, 32), (8, 128, 64), (9, 1280, 1280), (31, 1280, 1280), (33, 96, 200),
                                      (64, 1280, 5120), (100, 5120, 1280), (129, 1280, 1280),
                                      (257, 1280, 1280), (300, 1024, 4096), (1500, 1024, 1024)])
@pytest.mark.parametrize("tile", [0])
def test_gemm_family_tensor_cores(rows, k, n, tile):
    rng = np.random.default_rng(rows + k + n)
    w = (rng.standard_normal(n * k) * 0.02).astype(np.float32)
    qb = q4.quantize_q4_0_converter(w)
    x = rng.standard_normal((1, rows, k)).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    q4.set_threads(q4.host_cores())
    want = q4.q4_linear(x, qb, bias, n)
    lin = wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV), t(bias))
    out = lin.forward(t(x), flags=wb.WQ4_FORCE_GEMM).cpu().numpy()
    assert rel_err(out, want) < REL_TOL


def test_gemm_bf16_operands_documented_precision():
    """north_star names bf16 operands; they do not meet 1e-3 (SURVEY 7 hard part B), which is why
    f16 operands are the default.  bf16 stays available behind WQ4_OPERANDS_BF16 and is bounded here."""
    rows, k, n = 256, 1280, 1280
    rng = np.random.default_rng(5)
    qb = q4.quantize_q4_0_converter((rng.standard_normal(n * k) * 0.02).astype(np.float32))
    x = rng.standard_normal((1, rows, k)).astype(np.float32)
    q4.set_threads(q4.host_cores())
    want = q4.q4_matmul(x, qb, n)
    qt = wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV)
    e16 = rel_err(wb.q4_matmul(t(x), qt, flags=wb.WQ4_FORCE_GEMM).cpu().numpy(), want)
    eb16 = rel_err(wb.q4_matmul(t(x), qt, flags=wb.WQ4_FORCE_GEMM | wb.WQ4_OPERANDS_BF16).cpu().numpy(), want)
    print(f"rel err f16 operands {e16:.3e}  bf16 operands {eb16:.3e}")
    assert e16 < REL_TOL
    assert eb16 < 2e-2


def test_q4_linear_forward_shape():  # tests.rs:486-506
    fin, fout = 128, 64
    i = np.arange(fout * fin, dtype=np.float32)
    qb = q4.quantize_q4_0_tests((np.sin(i * np.float32(0.001)) * np.float32(0.1)).astype(np.float32))
    lin = wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(qb, [fout, fin], DEV), None)
    out = lin.forward(torch.zeros((2, 5, fin), device=DEV))
    assert list(out.shape) == [2, 5, fout]
    assert torch.all(out == 0)


def test_q4_linear_forward_with_bias():  # tests.rs:508-564
    fin, fout = 64, 32
    i = np.arange(fout * fin, dtype=np.float32)
    qb = q4.quantize_q4_0_tests((np.sin(i * np.float32(0.001)) * np.float32(0.1)).astype(np.float32))
    bias = (np.arange(fout, dtype=np.float32) * np.float32(0.01)).astype(np.float32)
    act = (np.arange(fin, dtype=np.float32) * np.float32(0.1)).astype(np.float32)
    expected = q4.reference_matmul(act, q4.dequantize_q4_0(qb, fout * fin), 1, fin, fout).ravel() + bias
    lin = wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(qb, [fout, fin], DEV), t(bias))
    out = lin.forward(t(act.reshape(1, 1, fin)))
    assert list(out.shape) == [1, 1, fout]
    assert np.abs(out.cpu().numpy().ravel() - expected).max() < 1e-3


def test_k_mismatch_is_an_error():  # op.rs:58-61
    qb = q4.quantize_q4_0_tests(gen_w(64 * 64))
    qt = wb.Q4Tensor.from_q4_bytes(qb, [64, 64], DEV)
    with pytest.raises(wb.Wq4Error) as e:
        wb.q4_matmul(torch.zeros((1, 1, 32), device=DEV), qt)
    assert "K dimension mismatch" in e.value.message
    with pytest.raises(wb.Wq4Error):
        wb.q4_matmul(torch.zeros((1, 64), device=DEV), qt)          # op.rs:53
    flat = wb.Q4Tensor.from_q4_bytes(q4.quantize_q4_0_tests(gen_w(256)), [16, 16], DEV)
    with pytest.raises(wb.Wq4Error) as e:
        wb.q4_matmul(torch.zeros((1, 1, 16), device=DEV), flat)
    assert e.value.status == 5


def _mk_lin(rows, cols, scale=0.05, bias=False, seed=None):
    if seed is None:
        i = np.arange(rows * cols, dtype=np.float32)
        w = (np.sin(i * np.float32(0.001)) * np.float32(scale)).astype(np.float32)
        qb = q4.quantize_q4_0_tests(w)
        b = None
    else:
        rng = np.random.default_rng(seed)
        qb = q4.quantize_q4_0_converter((rng.standard_normal(rows * cols) * scale).astype(np.float32))
        b = rng.standard_normal(rows).astype(np.float32) * 0.1 if bias else None
    return qb, b


def test_q4_ffn_forward_shape():  # tests.rs:570-597
    d, f = 64, 256
    q1, _ = _mk_lin(f, d)
    q2, _ = _mk_lin(d, f)
    ffn = wb.Q4FFN(wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(q1, [f, d], DEV), None),
                   wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(q2, [d, f], DEV), None))
    out = ffn.forward(torch.zeros((1, 4, d), device=DEV))
    assert list(out.shape) == [1, 4, d] and torch.all(out == 0)


@pytest.mark.parametrize("b,m,d,f", [(1, 1, 1280, 5120), (1, 4, 64, 256), (2, 3, 256, 1024), (1, 200, 1024, 4096),
                                     (1, 1500, 1280, 5120)])
def test_q4_ffn_values(b, m, d, f):
    q1, b1 = _mk_lin(f, d, 0.03, True, seed=10)
    q2, b2 = _mk_lin(d, f, 0.03, True, seed=11)
    rng = np.random.default_rng(12)
    x = rng.standard_normal((b, m, d)).astype(np.float32)
    q4.set_threads(q4.host_cores())
    want = q4.q4_ffn(x, q1, b1, q2, b2, f)
    ffn = wb.Q4FFN(wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(q1, [f, d], DEV), t(b1)),
                   wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(q2, [d, f], DEV), t(b2)))
    out = ffn.forward(t(x)).cpu().numpy()
    assert rel_err(out, want) < REL_TOL


@pytest.mark.parametrize("b,cin,t_in,cout,stride", [(2, 5, 40, 6, 1), (2, 5, 40, 6, 2), (1, 80, 3000, 1024, 1),
                                                    (1, 128, 300, 1280, 1), (2, 256, 300, 256, 2),
                                                    (1, 1024, 3000, 1024, 2),
                                                    # Large-V3 stem at full size (encoder.rs:89,93): (3000, 384, 1280) and (1500, 3840, 1280)
                                                    (1, 128, 3000, 1280, 1), (1, 1280, 3000, 1280, 2)])
def test_conv1d_matches_reference_layer(b, cin, t_in, cout, stride):  # layers.rs:77-132
    rng = np.random.default_rng(cin + cout + stride)
    x = rng.standard_normal((b, cin, t_in)).astype(np.float32)
    w = (rng.standard_normal((cout, cin, 3)) * (1.0 / np.sqrt(3 * cin))).astype(np.float32)
    bias = rng.standard_normal(cout).astype(np.float32) * 0.1
    q4.set_threads(q4.host_cores())
    want = q4.conv1d(x, w, bias, 1, stride)
    conv = wb.Conv1D(w, bias, device=DEV)
    out = conv.forward(t(x), 1, stride).cpu().numpy()
    assert out.shape == want.shape
    assert rel_err(out, want) < REL_TOL
    want_g = q4.gelu(want)
    out_g = conv.forward(t(x), 1, stride, flags=wb.WQ4_EPI_GELU).cpu().numpy()
    assert rel_err(out_g, want_g) < REL_TOL


@pytest.mark.parametrize("b,cin,t_in,cout,stride", [(2, 80, 3000, 1024, 1), (1, 1280, 3000, 1280, 2)])
def test_conv1d_against_torch_conv1d(b, cin, t_in, cout, stride):
    """Second opinion for the row the reference leaves unpinned (layers.rs:77-132 has no test): the CUDA stem against
    torch.nn.functional.conv1d in f32 (TF32 off) on the same inputs, an implementation independent of oracle/."""
    rng = np.random.default_rng(7 + cin)
    x = t(rng.standard_normal((b, cin, t_in)).astype(np.float32))
    w = (rng.standard_normal((cout, cin, 3)) * (1.0 / np.sqrt(3 * cin))).astype(np.float32)
    bias = rng.standard_normal(cout).astype(np.float32) * 0.1
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        want = torch.nn.functional.conv1d(x, t(w), t(bias), stride=stride, padding=1)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    got = wb.Conv1D(w, bias, device=DEV).forward(x, 1, stride)
    assert got.shape == want.shape
    e = float((got - want).abs().max() / want.abs().max())
    print("conv1d vs torch rel err", e)
    assert e < REL_TOL


@pytest.mark.parametrize("rows", [17, 33, 64])
@pytest.mark.parametrize("n,k", [(1280, 1280), (1280, 5120), (3840, 1280)])
def test_small_row_counts_one_tile(rows, n, k):
    """17..64 rows (decode steps of that many clips): one 64-row tile per CTA pair at the Large-V3 decoder shapes, against
    the oracle and bit-identical run to run."""
    rng = np.random.default_rng(rows + n + k)
    qb = q4.quantize_q4_0_converter((rng.standard_normal(n * k) * 0.02).astype(np.float32))
    bias = rng.standard_normal(n).astype(np.float32) * 0.1
    lin = wb.Q4Linear(wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV), t(bias))
    x = rng.standard_normal((1, rows, k)).astype(np.float32)
    q4.set_threads(q4.host_cores())
    want = q4.q4_linear(x, qb, bias, n)
    tile = __import__("ctypes").c_int32()
    assert wb.lib().wq4_linear_plan(lin.weights._h, rows, 0, __import__("ctypes").byref(tile)) == 1 and tile.value == 64
    y1 = lin.forward(t(x))
    y2 = lin.forward(t(x))
    y3 = lin.forward(t(x))
    assert torch.equal(y1, y2) and torch.equal(y1, y3)
    assert rel_err(y1.cpu().numpy(), want) < REL_TOL


def test_linearity_and_determinism_at_full_size():
    """Size-independent properties at the Large-V3 encoder shape (no oracle needed):
    f(a*x) == a*f(x) for a power of two (exact in floating point), f(0) == bias, and two runs
    are bit-identical."""
    rows, k, n = 1500, 1280, 5120
    rng = np.random.default_rng(20)
    qb = q4.quantize_q4_0_converter((rng.standard_normal(n * k) * 0.02).astype(np.float32))
    bias = rng.standard_normal(n).astype(np.float32)
    qt = wb.Q4Tensor.from_q4_bytes(qb, [n, k], DEV)
    lin = wb.Q4Linear(qt, t(bias))
    xh = rng.standard_normal((1, rows, k)).astype(np.float32)
    xh = np.where(np.abs(xh) < 1e-2, np.float32(1e-2), xh)   # keep x and 4x out of f16 subnormals
    x = t(xh)
    y1 = wb.q4_matmul(x, qt)
    y2 = wb.q4_matmul(x, qt)
    assert torch.equal(y1, y2)
    assert torch.equal(wb.q4_matmul(x * 4.0, qt), y1 * 4.0)
    z = lin.forward(torch.zeros((1, rows, k), device=DEV))
    assert torch.equal(z, t(bias).expand(1, rows, n))
    # GEMV vs GEMM family agree on the same single row
    x1 = x[:, :1].contiguous()
    a = wb.q4_matmul(x1, qt, flags=wb.WQ4_FORCE_GEMV).cpu().numpy()
    b_ = wb.q4_matmul(x1, qt, flags=wb.WQ4_FORCE_GEMM).cpu().numpy()
    assert rel_err(b_, a) < REL_TOL
