"""CPU-only: the C++ GGUF reader (host logic of the drop-in loader) against a file written by the
reference's own converter, the numpy model oracle's building blocks against the C oracle / torch,
and a world_size-2 gloo check of the utterance sharding used by bench.py."""
import ctypes as C
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

import whisper_burn_b200 as wb
from oracle import gguf_np, q4, whisper_np as wn

ROOT = Path(__file__).resolve().parent.parent


def _info(L, buf, name):
    dims = (C.c_uint64 * 4)()
    nd, dt = C.c_uint32(), C.c_uint32()
    off, nb = C.c_uint64(), C.c_uint64()
    L.wq4_gguf_tensor_info.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    st = L.wq4_gguf_tensor_info(buf.ctypes.data_as(C.c_void_p), buf.size, name.encode(), dims, C.byref(nd), C.byref(dt),
                                C.byref(off), C.byref(nb))
    return st, list(dims)[:nd.value], dt.value, off.value, nb.value


def test_cpp_gguf_reader_on_reference_written_file(golden_dir):
    """reader.rs:105-223 restated in C++ parses what scripts/convert_whisper.py wrote."""
    L = wb.lib()
    raw = (golden_dir / "tiny_whisper.gguf").read_bytes()
    buf = np.frombuffer(raw, dtype=np.uint8)
    ver, cnt, doff = C.c_uint32(), C.c_uint64(), C.c_uint64()
    L.wq4_gguf_probe.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    assert L.wq4_gguf_probe(buf.ctypes.data_as(C.c_void_p), buf.size, C.byref(ver), C.byref(cnt), C.byref(doff)) == 0
    g = gguf_np.GgufFile(raw)
    assert (ver.value, cnt.value, doff.value) == (3, 50, g.data_offset)
    names = json.loads((golden_dir / "names_golden.json").read_text())
    for meta in names.values():
        if not meta["gguf"]:
            continue
        st, dims, dt, off, nb = _info(L, buf, meta["gguf"])
        info = g.tensors[meta["gguf"]]
        assert st == 0 and dims == info.dims and dt == info.dtype
        assert off == g.data_offset + info.offset and nb == info.byte_size
    st, *_ = _info(L, buf, "nonexistent")      # tests.rs:323
    assert st != 0 and b"not found" in L.wq4_last_error()
    bad = buf.copy()
    bad[0] = 0
    assert L.wq4_gguf_probe(bad.ctypes.data_as(C.c_void_p), bad.size, None, None, None) != 0
    trunc = buf[:100].copy()
    assert L.wq4_gguf_probe(trunc.ctypes.data_as(C.c_void_p), trunc.size, None, None, None) != 0


def test_gguf_multiple_tensors_and_versions():   # tests.rs:304-324 + reader.rs:118-121
    L = wb.lib()
    mk = lambda v, n: q4.quantize_q4_0_tests(np.full(n, v, np.float32)).tobytes()
    tensors = [("weight_a", (32, 32), 2, mk(0.1, 1024)), ("weight_b", (32, 64), 2, mk(0.2, 2048)),
               ("weight_c", (64, 32), 2, mk(-0.1, 2048))]
    for version in (2, 3):
        blob = np.frombuffer(gguf_np.write_gguf(tensors, version=version), dtype=np.uint8)
        for name, shape, _, raw in tensors:
            st, dims, dt, off, nb = _info(L, blob, name)
            assert st == 0 and dims == list(reversed(shape)) and dt == 2 and nb == len(raw)
            assert bytes(blob[off:off + nb]) == raw
        assert _info(L, blob, "nonexistent")[0] != 0
    bad = np.frombuffer(gguf_np.write_gguf(tensors, version=4), dtype=np.uint8)
    assert L.wq4_gguf_probe(bad.ctypes.data_as(C.c_void_p), bad.size, None, None, None) != 0


def test_numpy_model_blocks_match_c_oracle_and_torch():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((2, 6, 50)).astype(np.float32)
    w = rng.standard_normal((7, 6, 3)).astype(np.float32)
    b = rng.standard_normal(7).astype(np.float32)
    for s in (1, 2):
        assert np.abs(wn.conv1d(x, w, b, 1, s) - q4.conv1d(x, w, b, 1, s)).max() < 1e-5
    h = rng.standard_normal((3, 5, 64)).astype(np.float32)
    assert np.abs(wn.gelu(h) - q4.gelu(h)).max() < 1e-6
    lw, lb = rng.standard_normal(64).astype(np.float32), rng.standard_normal(64).astype(np.float32)
    assert np.abs(wn.layernorm(h, lw, lb) - q4.layernorm(h, lw, lb)).max() < 1e-5
    q, k, v = (rng.standard_normal((2, 9, 128)).astype(np.float32) for _ in range(3))
    for causal in (False, True):
        got = wn.sdpa(q, k, v, 2, causal)
        tq, tk, tv = (torch.from_numpy(a).reshape(2, 9, 2, 64).transpose(1, 2) for a in (q, k, v))
        ref = torch.nn.functional.scaled_dot_product_attention(tq, tk, tv, is_causal=causal)
        assert np.abs(got - ref.transpose(1, 2).reshape(2, 9, 128).numpy()).max() < 1e-5


def test_random_model_gguf_roundtrip_and_quirks():
    cfg = wn.WhisperConfig(80, 16, 256, 4, 1, 12, 256, 4, 1, 51865, 99)
    st = wn.random_state(cfg, seed=3, n_vocab=600)
    blob = wn.state_to_gguf(st)
    g = gguf_np.GgufFile(blob)
    assert g.tensors["encoder.blocks.0.attn.query.weight"].dtype == gguf_np.GGML_Q4_0
    assert g.tensors["encoder.conv1.weight"].dtype == gguf_np.GGML_F32
    assert "encoder.blocks.0.attn.key.bias" not in g.tensors
    o = wn.WhisperOracle(blob, cfg)
    mel = wn.synthetic_mel(1, 80, 32, seed=1)
    enc = o.encode(mel)
    assert enc.shape == (1, 16, 256) and np.isfinite(enc).all()
    # auto-detect quirk: the SOT step's cache entry is overwritten by the 3-token prompt
    cache = o.new_cache()
    o.decode_step(5, 0, enc, cache)
    assert cache[0]["self_k"].shape[1] == 1
    o.forward_prompt([7, 8, 9], enc, cache)
    assert cache[0]["self_k"].shape[1] == 3
    o.decode_step(10, 4, enc, cache)
    assert cache[0]["self_k"].shape[1] == 4
    assert wn.WhisperOracle.argmax(np.array([1.0, 3.0, 3.0, 2.0], np.float32)) == 2   # last max wins


def test_sharding_two_ranks_gloo(tmp_path):
    """bench.py's utterance split: 2 CPU ranks over gloo agree on disjoint, covering shards and on the
    max-over-ranks reduction used for timing."""
    script = tmp_path / "w.py"
    script.write_text(f"""
import sys, torch, torch.distributed as dist
sys.path.insert(0, {str(ROOT)!r})
from bench import shard_clips, max_over_ranks
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
lo, hi = shard_clips(257, r, w)
t = torch.tensor([hi - lo], dtype=torch.int64)
dist.all_reduce(t)
assert t.item() == 257, t
g = [None] * w
dist.all_gather_object(g, (lo, hi))
assert g[0][1] == g[1][0] and g[0][0] == 0 and g[1][1] == 257
assert abs(max_over_ranks(1.0 + r) - 2.0) < 1e-9
dist.barrier()
print("rank", r, "ok")
""")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                       capture_output=True, text=True, env=env, timeout=240)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
