"""ctypes loader for libwq4.so (the C ABI declared in include/wq4.h).

The product path has no CPU fallback: if the shared library is missing or a call fails the
caller gets an exception, never a silently different implementation.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

_PKG = Path(__file__).resolve().parent
SO_PATH = _PKG / "libwq4.so"

WQ4_OK = 0
WQ4_EPI_GELU = 1
WQ4_OPERANDS_BF16 = 16
WQ4_FORCE_GEMM = 32
WQ4_FORCE_GEMV = 64
WQ4_FORCE_SKINNY = 128

_STATUS = {1: "invalid argument", 2: "shape", 3: "byte count", 4: "cuda", 5: "unsupported", 6: "workspace"}


class Wq4Error(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"wq4 error {status} ({_STATUS.get(status, '?')}): {message}")
        self.status = status
        self.message = message


def build(verbose: bool = False, clean: bool = False) -> Path:
    """Compile the CUDA sources for sm_100a in-tree (nvcc cross-compiles without a GPU).  clean=True rebuilds every object."""
    r = subprocess.run([str(_PKG / "csrc" / "build.sh")] + (["--clean"] if clean else []), capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-8000:])
    if r.returncode != 0:
        raise RuntimeError("nvcc build of libwq4.so failed")
    return SO_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not SO_PATH.exists():
        raise RuntimeError(
            f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(whisper_burn_b200/csrc/build.sh). There is no CPU fallback.")
    L = C.CDLL(str(SO_PATH))
    vp, sz, i32, i64, u32 = C.c_void_p, C.c_size_t, C.c_int32, C.c_int64, C.c_uint32
    L.wq4_last_error.restype = C.c_char_p
    L.wq4_version.restype = i32
    L.wq4_device_count.restype = i32
    L.wq4_tensor_create.argtypes = [vp, sz, i64, i64, i32, vp, C.POINTER(vp)]
    L.wq4_tensor_destroy.argtypes = [vp]
    L.wq4_tensor_shape.argtypes = [vp, C.POINTER(i64), C.POINTER(i64)]
    L.wq4_tensor_num_blocks.argtypes = [vp]
    L.wq4_tensor_num_blocks.restype = i64
    L.wq4_tensor_device.argtypes = [vp]
    L.wq4_tensor_device_bytes.argtypes = [vp]
    L.wq4_tensor_device_bytes.restype = sz
    L.wq4_tensor_dequantize.argtypes = [vp, vp, vp]
    L.wq4_tensor_export_gguf.argtypes = [vp, vp, vp]
    L.wq4_linear_workspace_bytes.argtypes = [vp, i64, i64]
    L.wq4_linear_workspace_bytes.restype = sz
    L.wq4_matmul.argtypes = [vp, vp, vp, i64, i64, vp, sz, vp]
    L.wq4_linear_forward.argtypes = [vp, vp, vp, vp, i64, i64, vp, sz, vp]
    L.wq4_linear_forward_ex.argtypes = [vp, vp, vp, vp, vp, vp, vp, i64, u32, vp, sz, vp]
    L.wq4_ffn_workspace_bytes.argtypes = [vp, vp, i64, i64]
    L.wq4_ffn_workspace_bytes.restype = sz
    L.wq4_ffn_forward.argtypes = [vp, vp, vp, vp, vp, vp, i64, i64, vp, sz, vp]
    L.wq4_ffn_forward_static.argtypes = [vp, vp, vp, vp, vp, vp, i64, i64, vp]
    L.wq4_ffn_forward_static.restype = i32
    L.wq4_ffn_static_limits.argtypes = [C.POINTER(i64), C.POINTER(i64)]
    L.wq4_ffn_static_limits.restype = None
    L.wq4_conv1d_create.argtypes = [vp, vp, i64, i64, i64, i32, vp, C.POINTER(vp)]
    L.wq4_conv1d_destroy.argtypes = [vp]
    L.wq4_conv1d_workspace_bytes.argtypes = [vp, i64, i64, i64, i64]
    L.wq4_conv1d_workspace_bytes.restype = sz
    L.wq4_conv1d_forward.argtypes = [vp, vp, vp, i64, i64, i64, i64, u32, vp, sz, vp]
    L.wq4_logmel_create.argtypes = [i32, i32, vp, C.POINTER(vp)]
    L.wq4_logmel_destroy.argtypes = [vp]
    L.wq4_logmel_forward.argtypes = [vp, vp, i64, i64, vp, vp]
    L.wq4_logmel_forward_ragged.argtypes = [vp, vp, i64, i64, vp, vp, vp]
    L.wq4_linear_plan.argtypes = [vp, i64, u32, C.POINTER(i32)]
    L.wq4_launch_count.restype = i64
    L.wq4_launch_count_reset.restype = None
    L.wq4_device_error_flag.argtypes = [i32]
    L.wq4_debug_f16_overflow_count.argtypes = [i32, i32]
    L.wq4_debug_f16_overflow_count.restype = i64
    for name in dir(L):
        pass
    for fn in ("wq4_tensor_create", "wq4_tensor_destroy", "wq4_tensor_shape", "wq4_tensor_device",
               "wq4_tensor_dequantize", "wq4_tensor_export_gguf", "wq4_matmul", "wq4_linear_forward",
               "wq4_linear_forward_ex", "wq4_ffn_forward", "wq4_conv1d_create", "wq4_conv1d_destroy",
               "wq4_conv1d_forward", "wq4_linear_plan", "wq4_device_error_flag", "wq4_logmel_create",
               "wq4_logmel_destroy", "wq4_logmel_forward", "wq4_logmel_forward_ragged"):
        getattr(L, fn).restype = i32
    _lib = L
    return L


def check(status: int) -> None:
    if status != WQ4_OK:
        raise Wq4Error(status, lib().wq4_last_error().decode("utf-8", "replace"))


def launch_count() -> int:
    return int(lib().wq4_launch_count())


def reset_launch_count() -> None:
    lib().wq4_launch_count_reset()
