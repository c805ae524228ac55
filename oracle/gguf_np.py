"""GGUF v2/v3 reader and writer in numpy (TEST INFRASTRUCTURE ONLY).

Reader follows src/gguf/reader.rs:105-223 (magic, version 2|3, metadata skipped :131-139,
tensor index :143-175, data section at the next 32-byte boundary :178-179, Q4_0 byte size
= elements/32*18 :38-48).  Writer follows the sequence of scripts/convert_whisper.py:188-217
(header, string/u32 metadata, index with reversed dims :163-166 and 32-byte aligned offsets
:168-178, padding, data).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass

import numpy as np

GGUF_MAGIC = 0x46554747  # reader.rs:13
ALIGNMENT = 32           # reader.rs:14
GGML_F32, GGML_F16, GGML_Q4_0 = 0, 1, 2  # reader.rs:27-35


@dataclass
class TensorInfo:
    name: str
    dims: list[int]   # as stored (reversed w.r.t. the row-major shape)
    dtype: int
    offset: int       # relative to the data section

    @property
    def num_elements(self) -> int:
        n = 1
        for d in self.dims:
            n *= d
        return n

    @property
    def byte_size(self) -> int:  # reader.rs:38-48
        n = self.num_elements
        if self.dtype == GGML_F32:
            return n * 4
        if self.dtype == GGML_F16:
            return n * 2
        if self.dtype == GGML_Q4_0:
            return n // 32 * 18
        raise ValueError(f"Unsupported GGML dtype code: {self.dtype}")


_SCALAR_SIZES = {0: 1, 1: 1, 2: 2, 3: 2, 4: 4, 5: 4, 6: 4, 7: 1, 10: 8, 11: 8, 12: 8}


class GgufFile:
    def __init__(self, data: bytes):
        self.data = memoryview(data)
        pos = 0

        def rd(fmt):
            nonlocal pos
            v = struct.unpack_from(fmt, self.data, pos)
            pos += struct.calcsize(fmt)
            return v[0]

        def rd_str():
            nonlocal pos
            n = rd("<Q")
            s = bytes(self.data[pos:pos + n]).decode("utf-8")
            pos += n
            return s

        def skip_value(t):
            nonlocal pos
            if t in _SCALAR_SIZES:
                pos += _SCALAR_SIZES[t]
            elif t == 8:
                rd_str()
            elif t == 9:
                et = rd("<I")
                cnt = rd("<Q")
                for _ in range(cnt):
                    skip_value(et)
            else:
                raise ValueError(f"Unknown GGUF metadata value type: {t}")

        magic = rd("<I")
        if magic != GGUF_MAGIC:
            raise ValueError(f"Invalid GGUF magic: 0x{magic:08X}")
        self.version = rd("<I")
        if self.version not in (2, 3):
            raise ValueError(f"Unsupported GGUF version: {self.version}")
        self.tensor_count = rd("<Q")
        kv_count = rd("<Q")
        for _ in range(kv_count):
            rd_str()
            skip_value(rd("<I"))
        self.tensors: dict[str, TensorInfo] = {}
        for _ in range(self.tensor_count):
            name = rd_str()
            nd = rd("<I")
            dims = [rd("<Q") for _ in range(nd)]
            dtype = rd("<I")
            if dtype not in (GGML_F32, GGML_F16, GGML_Q4_0):
                raise ValueError(f"Unsupported GGML dtype code: {dtype}")
            off = rd("<Q")
            self.tensors[name] = TensorInfo(name, dims, dtype, off)
        self.data_offset = (pos + ALIGNMENT - 1) // ALIGNMENT * ALIGNMENT

    def tensor_data(self, name: str) -> np.ndarray:
        info = self.tensors[name]
        a = self.data_offset + info.offset
        return np.frombuffer(self.data[a:a + info.byte_size], dtype=np.uint8)

    def tensor_f32(self, name: str) -> np.ndarray:
        """F32/F16 tensor as float32 in row-major shape (dims reversed, loader.rs:99-106)."""
        info = self.tensors[name]
        raw = self.tensor_data(name)
        if info.dtype == GGML_F32:
            v = raw.view(np.float32)
        elif info.dtype == GGML_F16:
            v = raw.view(np.float16).astype(np.float32)
        else:
            raise ValueError(f"Expected F32/F16 for '{name}'")
        return v.reshape(list(reversed(info.dims)))


def _wstr(buf: bytearray, s: str) -> None:
    e = s.encode("utf-8")
    buf += struct.pack("<Q", len(e))
    buf += e


def write_gguf(tensors: list[tuple[str, tuple[int, ...], int, bytes]],
               metadata: list[tuple[str, int, object]] | None = None, version: int = 3) -> bytes:
    """tensors: (name, row-major shape, ggml dtype, raw bytes).  Same byte sequence as
    scripts/convert_whisper.py:188-217 for the same inputs."""
    if metadata is None:
        metadata = [("general.architecture", 8, "whisper")]
    buf = bytearray()
    buf += struct.pack("<I", GGUF_MAGIC)
    buf += struct.pack("<I", version)
    buf += struct.pack("<Q", len(tensors))
    buf += struct.pack("<Q", len(metadata))
    for key, vt, val in metadata:
        _wstr(buf, key)
        buf += struct.pack("<I", vt)
        if vt == 4:
            buf += struct.pack("<I", val)
        elif vt == 8:
            _wstr(buf, val)
        else:
            raise ValueError("writer supports u32 and string metadata only")
    cur = 0
    pads = []
    for name, shape, dtype, raw in tensors:
        aligned = (cur + ALIGNMENT - 1) // ALIGNMENT * ALIGNMENT
        pads.append(aligned - cur)
        _wstr(buf, name)
        buf += struct.pack("<I", len(shape))
        for d in reversed(shape):
            buf += struct.pack("<Q", d)
        buf += struct.pack("<I", dtype)
        buf += struct.pack("<Q", aligned)
        cur = aligned + len(raw)
    buf += b"\x00" * ((-len(buf)) % ALIGNMENT)
    for pad, (_, _, _, raw) in zip(pads, tensors):
        buf += b"\x00" * pad
        buf += bytes(raw)
    return bytes(buf)
