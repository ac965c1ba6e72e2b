"""ctypes binding to the in-tree native library (``b200dist/_native/libb200dist.so``).

The library exposes a plain C ABI (``include/b200dist.h``); every entry point returns
0 on success and a small error code otherwise (``b200_last_error()`` has the text).
On a CUDA tensor the Python ops call straight into these kernels and FAIL LOUDLY if
the library is missing — there is no silent eager fallback on the GPU.  CPU tensors
(unit tests, gloo) use the plain PyTorch expressions in ``b200dist.ops.reference``.
"""
from __future__ import annotations

import ctypes
import os
import threading
from pathlib import Path

import torch

from ..utils import debug as _debug

_LIB_PATH = Path(__file__).resolve().parent.parent / "_native" / "libb200dist.so"
_lock = threading.Lock()
_lib = None

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_ll = ctypes.c_longlong
c_float = ctypes.c_float

# number of native kernel launches issued through this binding (bench.py reports it)
launch_count = 0


class NativeError(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    return _LIB_PATH.exists()


def load():
    """Load (once) and return the ctypes handle; build in-tree if the .so is absent."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not _LIB_PATH.exists():
            if os.environ.get("B200DIST_NO_AUTOBUILD") == "1":
                raise NativeError(f"{_LIB_PATH} is missing; run `python -m b200dist._build`")
            from .. import _build

            _build.build()
        lib = ctypes.CDLL(str(_LIB_PATH))
        lib.b200_last_error.restype = ctypes.c_char_p
        _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().b200_last_error().decode(errors="replace")
        raise NativeError(f"b200dist native call {what} failed (code {rc}): {msg}")


def ptr(t) -> c_void_p:
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def stream_ptr(device=None) -> c_void_p:
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.bfloat16:
        return 0
    if dt == torch.float32:
        return 1
    raise NativeError(f"b200dist kernels support bf16 and fp32 tensors, got {dt}")


_NVTX = os.environ.get("B200DIST_NVTX", "0") not in ("", "0")

# entry points that launch more than one kernel
_KERNELS_PER_CALL = {"rmsnorm_bwd": 2, "rmsnorm_bwd_res": 2, "layernorm_bwd": 3, "flash_attn_bwd": 3}   # (memsets are not counted)


def call(name: str, *args) -> None:
    """Invoke ``b200_<name>`` and raise on a non-zero return code."""
    global launch_count
    fn = getattr(load(), "b200_" + name)
    launch_count += _KERNELS_PER_CALL.get(name, 1)
    if _NVTX:   # B200DIST_NVTX=1: one NVTX range per native call (kernels and collectives) for nsys / torch.profiler
        torch.cuda.nvtx.range_push("b200_" + name)
        try:
            check(fn(*args), name)
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        check(fn(*args), name)
    if _debug.sync_debug_enabled():
        _debug.after_native_call(name)
