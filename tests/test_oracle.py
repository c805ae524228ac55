"""The oracle against (a) bytes produced by the reference's own Python converter
(tests/golden/, scripts/make_golden.py) and (b) the closed-form generators and tolerances of
the reference's unit tests, src/gguf/tests.rs.  CPU only."""
import json

import numpy as np
import pytest

from oracle import gguf_np, q4


# ------------------------------------------------------------------ (a) converter goldens
def _golden_cases(golden_dir):
    z = np.load(golden_dir / "q4_converter_golden.npz")
    names = sorted(k[3:] for k in z.files if k.startswith("in_"))
    return z, names


def test_converter_quantiser_matches_reference_bytes(golden_dir):
    """C and numpy twins of scripts/convert_whisper.py:33-74 are byte-identical to what the
    reference function returned in the build container."""
    z, names = _golden_cases(golden_dir)
    assert len(names) >= 15
    for n in names:
        x, want = z[f"in_{n}"], z[f"out_{n}"]
        with np.errstate(all="ignore"):
            got_np = q4.quantize_q4_0_converter_np(x)
        got_c = q4.quantize_q4_0_converter(x)
        assert np.array_equal(got_c, want), f"C converter quantiser differs on {n}"
        assert np.array_equal(got_np, want), f"numpy converter quantiser differs on {n}"


def test_dequant_of_golden_bytes_is_finite_and_bounded(golden_dir):
    z, names = _golden_cases(golden_dir)
    for n in names:
        if n in ("overflow_scale",):   # f16(d) = inf in the reference's own output
            continue
        x, qb = z[f"in_{n}"], z[f"out_{n}"]
        deq = q4.dequantize_q4_0(qb, x.size)
        d = np.abs(x.reshape(-1, 32)).max(axis=1) / 7.0
        err = np.abs(deq - x).reshape(-1, 32).max(axis=1)
        # half a quantisation step plus the f16 rounding of the scale times the max level
        assert np.all(err <= d * 0.5 + np.abs(d) * 7 * 2.0 ** -10 + 1e-7), n


def test_f16_conversion_matches_numpy():
    L = q4.lib()
    rng = np.random.default_rng(1)
    vals = np.concatenate([
        rng.standard_normal(2000).astype(np.float32) * 10.0 ** rng.integers(-9, 6, 2000),
        np.array([0.0, -0.0, 65504.0, 65520.0, 65519.99, 1e-8, 5.96e-8, 2.98e-8, 2.9802322e-8,
                  6.1e-5, 6.0975e-5, np.inf, -np.inf, 1.0009765625, 1.00048828125], np.float32)])
    with np.errstate(over="ignore"):
        want = vals.astype(np.float16).view(np.uint16)
    got = np.array([L.orc_f32_to_f16(float(v)) for v in vals], dtype=np.uint16)
    assert np.array_equal(got, want)
    allh = np.arange(65536, dtype=np.uint16)
    back = np.array([L.orc_f16_to_f32(int(h)) for h in allh[::7]], dtype=np.float32)
    ref = allh[::7].view(np.float16).astype(np.float32)
    assert np.array_equal(back.view(np.uint32)[~np.isnan(ref)], ref.view(np.uint32)[~np.isnan(ref)])


# ------------------------------------------------- (b) the reference's unit tests, restated
def test_q4_block_dequant():  # src/gguf/tests.rs:190-226
    original = ((np.arange(32, dtype=np.float32) - np.float32(15.5)) / np.float32(15.5))
    qb = q4.quantize_q4_0_tests(original)
    assert qb.size == 18
    d = qb[:2].view(np.float16).astype(np.float32)[0]
    assert abs(d - np.abs(original).max() / 7.0) < 0.01
    deq = q4.dequantize_q4_0(qb, 32)
    assert np.abs(deq - original).max() < 0.08


def test_q4_block_edge_cases():  # src/gguf/tests.rs:228-273
    z = q4.dequantize_q4_0(q4.quantize_q4_0_tests(np.zeros(32, np.float32)), 32)
    assert np.all(z == 0.0)
    u = np.full(32, 0.5, np.float32)
    assert np.abs(q4.dequantize_q4_0(q4.quantize_q4_0_tests(u), 32) - u).max() < 0.08
    large = (np.arange(32, dtype=np.float32) - np.float32(15.5)) * np.float32(100.0)
    d_large = np.abs(large).max() / 7.0
    err = np.abs(q4.dequantize_q4_0(q4.quantize_q4_0_tests(large), 32) - large).max()
    assert err < d_large / 2.0 + 1.0


def test_tests_quantiser_python_restatement():
    """Independent pure-Python restatement of tests.rs:24-57 on small cases (round-half-up via
    `as u8` truncation of v*id + 8.5, saturation of negatives to 0 and of >15 to 15)."""
    rng = np.random.default_rng(2)
    data = np.concatenate([rng.standard_normal(8 * 32).astype(np.float32),
                           np.array([7.0] + [k + 0.5 for k in range(-8, 8)] + [0.0] * 15, np.float32)])
    got = q4.quantize_q4_0_tests(data)
    out = bytearray()
    for b in range(data.size // 32):
        blk = data[b * 32:(b + 1) * 32]
        amax = np.float32(np.abs(blk).max())
        d = np.float32(amax / np.float32(7.0))
        idv = np.float32(1.0) / d if d != 0 else np.float32(0.0)
        out += np.float16(d).tobytes()
        qs = []
        for v in blk:
            t = np.float32(np.float32(v * idv) + np.float32(8.5))
            qv = 0 if not (t > 0) else min(int(t), 255)
            qs.append(min(qv, 15))
        for i in range(16):
            out.append(qs[i] | (qs[i + 16] << 4))
    assert bytes(got) == bytes(out)


def _gen_act(n):   # tests.rs:432-434
    i = np.arange(n, dtype=np.float32)
    return (np.sin(i * np.float32(0.001)) * np.float32(0.1)).astype(np.float32)


def _gen_w(n):     # tests.rs:436-438
    i = np.arange(n, dtype=np.float32)
    return (np.cos(i * np.float32(0.0007)) * np.float32(0.05)).astype(np.float32)


def test_q4_matmul_small():  # src/gguf/tests.rs:369-410
    k = n = 32
    i = np.arange(n * k, dtype=np.float32)
    w = (np.sin(i * np.float32(0.1)) * np.float32(0.5)).astype(np.float32)
    qb = q4.quantize_q4_0_tests(w)
    wdeq = q4.dequantize_q4_0(qb, n * k)
    act = (np.arange(k, dtype=np.float32) * np.float32(0.1)).astype(np.float32)
    expected = q4.reference_matmul(act, wdeq, 1, k, n)
    out = q4.q4_matmul(act.reshape(1, 1, k), qb, n)
    assert out.shape == (1, 1, n)
    assert np.abs(out.ravel() - expected.ravel()).max() < 1e-3


@pytest.mark.parametrize("batch,seq,k,n,tol", [
    (1, 1, 128, 64, 1e-2), (1, 1, 1280, 1280, 1e-2), (1, 10, 1280, 1280, 1e-2),
    (1, 1, 1280, 5120, 1e-2), (1, 1, 5120, 1280, 1e-2),   # tests.rs:418-424
    (4, 10, 128, 64, 1e-3),                                # tests.rs:603-655
    (1, 32, 1280, 1280, 1e-2),                             # tests.rs:712-764
])
def test_q4_matmul_shapes(batch, seq, k, n, tol):
    act = _gen_act(batch * seq * k)
    qb = q4.quantize_q4_0_tests(_gen_w(n * k))
    wdeq = q4.dequantize_q4_0(qb, n * k)
    # the reference's expected value is a Burn f32 matmul; numpy f32 matmul stands in for it
    expected = act.reshape(batch * seq, k) @ wdeq.reshape(n, k).T
    fused = q4.q4_matmul(act.reshape(batch, seq, k), qb, n).reshape(batch * seq, n)
    two_pass = q4.dequant_then_matmul(act, qb, batch * seq, k, n)
    assert np.abs(fused - expected).max() < tol
    assert np.abs(two_pass - expected).max() < tol
    assert np.abs(fused - two_pass).max() < 1e-4


def test_q4_linear_forward_with_bias():  # src/gguf/tests.rs:508-564
    fin, fout = 64, 32
    i = np.arange(fout * fin, dtype=np.float32)
    w = (np.sin(i * np.float32(0.001)) * np.float32(0.1)).astype(np.float32)
    qb = q4.quantize_q4_0_tests(w)
    wdeq = q4.dequantize_q4_0(qb, fout * fin)
    bias = (np.arange(fout, dtype=np.float32) * np.float32(0.01)).astype(np.float32)
    act = (np.arange(fin, dtype=np.float32) * np.float32(0.1)).astype(np.float32)
    expected = q4.reference_matmul(act, wdeq, 1, fin, fout).ravel() + bias
    out = q4.q4_linear(act.reshape(1, 1, fin), qb, bias, fout)
    assert out.shape == (1, 1, fout)
    assert np.abs(out.ravel() - expected).max() < 1e-3


def test_q4_ffn_forward_shape():  # src/gguf/tests.rs:570-597
    d_model, ffn = 64, 256
    mk = lambda r, c: q4.quantize_q4_0_tests(
        (np.sin(np.arange(r * c, dtype=np.float32) * np.float32(0.001)) * np.float32(0.05)))
    out = q4.q4_ffn(np.zeros((1, 4, d_model), np.float32), mk(ffn, d_model), None,
                    mk(d_model, ffn), None, ffn)
    assert out.shape == (1, 4, d_model)
    assert np.all(out == 0.0)


def test_gelu_layernorm_conv_against_numpy():
    rng = np.random.default_rng(3)
    x = rng.standard_normal((2, 7, 48)).astype(np.float32)
    g = q4.gelu(x)
    want = 0.5 * x.astype(np.float64) * (1 + np.tanh(np.sqrt(2 / np.pi) * (x + 0.044715 * x.astype(np.float64) ** 3)))
    assert np.abs(g - want).max() < 1e-6
    w, b = rng.standard_normal(48).astype(np.float32), rng.standard_normal(48).astype(np.float32)
    ln = q4.layernorm(x, w, b)
    x64 = x.astype(np.float64)
    want = (x64 - x64.mean(-1, keepdims=True)) / np.sqrt(x64.var(-1, keepdims=True) + 1e-5) * w + b
    assert np.abs(ln - want).max() < 1e-5
    import torch
    xc = rng.standard_normal((2, 5, 40)).astype(np.float32)
    wc = rng.standard_normal((6, 5, 3)).astype(np.float32)
    bc = rng.standard_normal(6).astype(np.float32)
    for stride in (1, 2):
        got = q4.conv1d(xc, wc, bc, 1, stride)
        ref = torch.nn.functional.conv1d(torch.from_numpy(xc), torch.from_numpy(wc),
                                         torch.from_numpy(bc), stride=stride, padding=1).numpy()
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() < 1e-5


# ------------------------------------------------------------------ GGUF (reader.rs tests)
def test_gguf_reader_parse_header():  # src/gguf/tests.rs:279-302
    n = 32 * 64
    orig = np.sin(np.arange(n, dtype=np.float32) * np.float32(0.001) - np.float32(1.0)).astype(np.float32)
    qb = q4.quantize_q4_0_tests(orig)
    blob = gguf_np.write_gguf([("test.weight", (64, 32), gguf_np.GGML_Q4_0, qb.tobytes())])
    g = gguf_np.GgufFile(blob)
    assert g.version == 3 and g.tensor_count == 1
    info = g.tensors["test.weight"]
    assert info.dims == [32, 64] and info.dtype == gguf_np.GGML_Q4_0
    assert np.array_equal(g.tensor_data("test.weight"), qb)


def test_gguf_reference_written_file(golden_dir):
    """A GGUF written by the reference's own convert_model parses, its Q4 tensors are the
    converter quantisation of the f32 originals, names follow loader.rs:160-377."""
    blob = (golden_dir / "tiny_whisper.gguf").read_bytes()
    g = gguf_np.GgufFile(blob)
    names = json.loads((golden_dir / "names_golden.json").read_text())
    expect = {v["gguf"]: v for v in names.values() if v["gguf"]}
    assert g.tensor_count == len(expect) == 50
    for name, meta in expect.items():
        info = g.tensors[name]
        assert info.dims == list(reversed(meta["shape"]))
        assert (info.dtype == gguf_np.GGML_Q4_0) == meta["q4"]
    st = np.load(golden_dir / "tiny_whisper_state.npz")
    for name in st.files:
        assert np.array_equal(g.tensor_data(name), q4.quantize_q4_0_converter(st[name]))
    assert g.tensors["decoder.blocks.0.cross_attn.key.weight"].dtype == gguf_np.GGML_Q4_0
    assert "decoder.blocks.0.cross_attn.key.bias" not in g.tensors   # loader.rs:208
    assert g.tensors["encoder.conv1.weight"].dims == [3, 80, 256]     # loader.rs:256-258


def test_oracle_matches_hf_second_opinion(golden_dir):
    """The model-level oracle (Conv1D stem, LayerNorm, SDPA, KV cache, logits: rows the reference leaves unpinned) against an
    INDEPENDENT implementation: Hugging Face Whisper run on the dequantised weights of the GGUF the reference's converter
    wrote (scripts/make_golden_hf.py -> hf_second_opinion.npz).  Measured 2.4e-7 (encoder) / 8.1e-7 (logits)."""
    from oracle import whisper_np as wn
    fx = np.load(golden_dir / "hf_second_opinion.npz")
    cfg = wn.WhisperConfig(80, 32, 256, 4, 1, 16, 256, 4, 1, 64, 99)
    o = wn.WhisperOracle((golden_dir / "tiny_whisper.gguf").read_bytes(), cfg)
    enc = o.encode(fx["mel"])
    assert np.abs(enc - fx["enc_out"]).max() / np.abs(fx["enc_out"]).max() < 5e-6
    for b in range(fx["tokens"].shape[0]):
        toks = [int(t) for t in fx["tokens"][b]]
        cache = o.new_cache()
        lg = o.forward_prompt(toks[:3], enc[b:b + 1], cache)           # decoder.rs:251-296: logits of the last prompt position
        worst = np.abs(lg - fx["logits"][b, 2]).max() / np.abs(fx["logits"][b, 2]).max()
        for p in range(3, len(toks)):                                  # decoder.rs:306-348: incremental steps over the KV cache
            lg = o.decode_step(toks[p], p, enc[b:b + 1], cache)
            worst = max(worst, np.abs(lg - fx["logits"][b, p]).max() / np.abs(fx["logits"][b, p]).max())
            assert int(lg.argmax()) == int(fx["logits"][b, p].argmax())
        assert worst < 5e-6, worst
