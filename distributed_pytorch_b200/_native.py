"""Loader / builder for ``libdpb200.so`` (the sm_100a kernels and the C++ runtime, C ABI in ``include/dpb200.h``).

The library is plain CUDA C++ compiled by ``nvcc -gencode arch=compute_100a,code=sm_100a``; Python talks to it with
``ctypes`` and passes raw device pointers plus the current CUDA stream, so every launch lands on the stream PyTorch is
using (and is captured by ``torch.cuda.graph``).  The ``.so`` lives in-tree (``distributed_pytorch_b200/_lib``).

There is no fallback: on a CUDA device a missing or unloadable library is a hard error (``require()``).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import threading
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_longlong, c_size_t, c_void_p

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_CSRC = os.path.join(_ROOT, "csrc")
_LIBDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_lib")
LIB_PATH = os.path.join(_LIBDIR, "libdpb200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static",
]

_lock = threading.Lock()
_lib = None
_load_error: str | None = None


def _sources() -> list[str]:
    out = [os.path.join(_ROOT, "include", "dpb200.h")]
    for name in sorted(os.listdir(_CSRC)):
        if name.endswith((".cu", ".cuh", ".h")):
            out.append(os.path.join(_CSRC, name))
    return out


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > t for s in _sources())


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/dpb200.cu for sm_100a into the in-tree shared library. Returns the library path."""
    if not force and not needs_build():
        return LIB_PATH
    os.makedirs(_LIBDIR, exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc, *NVCC_FLAGS, "-o", LIB_PATH, os.path.join(_CSRC, "dpb200.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({' '.join(cmd)}):\n{proc.stdout}\n{proc.stderr}")
    if verbose:
        print(proc.stderr)
    return LIB_PATH


_P = c_void_p
_I = c_int
_LL = c_longlong
_F = c_float

_SIGNATURES = {
    "dpb_version": (c_int, []),
    "dpb_last_error": (c_char_p, []),
    "dpb_last_device_error": (c_int, [POINTER(c_int)]),
    "dpb_debug_set_conv_dump": (c_int, [_P]),
    "dpb_conv_workspace_bytes": (c_size_t, []),
    "dpb_conv_counter_bytes": (c_size_t, []),
    "dpb_conv3x3_gemm": (c_int, [_I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, POINTER(c_int), _P]),
    "dpb_conv_last_launches": (c_int, []),
    "dpb_set_tuning": (c_int, [_I, _I]),
    "dpb_first_conv_blocks": (c_int, [_I, _I, _I]),
    "dpb_first_conv_fprop": (c_int, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_first_conv_wgrad_blocks": (c_int, []),
    "dpb_first_conv_wgrad": (c_int, [_P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_bn_finalize": (c_int, [_P, _P, _I, _I, c_double, _P, _P, _P, _P, _P, _F, _F, _P, _P, _P, _P, _P]),
    "dpb_bn_eval_scale_shift": (c_int, [_I, _P, _P, _P, _P, _F, _P, _P, _P]),
    "dpb_bn_relu_pool_fwd": (c_int, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dpb_bn_bwd_blocks": (c_int, []),
    "dpb_bn_bwd_launches": (c_int, [_I, _I, _I, _I, _I]),
    "dpb_bn_relu_pool_bwd": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dpb_relu_fwd": (c_int, [_P, _P, _LL, _P]),
    "dpb_relu_bwd": (c_int, [_P, _P, _P, _LL, _P]),
    "dpb_maxpool2x2_fwd": (c_int, [_P, _P, _I, _I, _I, _I, _P]),
    "dpb_maxpool2x2_bwd": (c_int, [_P, _P, _P, _I, _I, _I, _I, _P]),
    "dpb_column_stats_blocks": (c_int, []),
    "dpb_column_stats": (c_int, [_P, _P, _P, _LL, _I, POINTER(c_int), _P]),
    "dpb_column_sum": (c_int, [_P, _P, _P, _LL, _I, _P]),
    "dpb_bn_apply": (c_int, [_P, _P, _P, _P, _LL, _I, _P]),
    "dpb_bn_plain_bwd": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _LL, _I, _P]),
    "dpb_noalloc_max_channels": (c_int, []),
    "dpb_bn_relu_pool_bwd_noalloc": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dpb_linear_fwd": (c_int, [_P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_linear_bwd": (c_int, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_cross_entropy_fwd": (c_int, [_P, _P, _P, _P, _I, _I, _P]),
    "dpb_cross_entropy_bwd": (c_int, [_P, _P, _P, _P, _I, _I, _P]),
    "dpb_sgd_step": (c_int, [_P, _P, _P, _LL, _F, _F, _F, _F, _I, _P]),
    "dpb_symm_alloc": (c_int, [c_size_t, POINTER(c_void_p)]),
    "dpb_symm_free": (c_int, [_P]),
    "dpb_symm_export": (c_int, [_P, ctypes.c_char_p]),
    "dpb_symm_import": (c_int, [ctypes.c_char_p, POINTER(c_void_p)]),
    "dpb_symm_close": (c_int, [_P]),
    "dpb_signal_pad_bytes": (c_size_t, []),
    "dpb_comm_create": (c_int, [_I, _I, POINTER(c_void_p), _I, POINTER(c_void_p)]),
    "dpb_comm_destroy": (c_int, [_P]),
    "dpb_allreduce": (c_int, [_P, _I, POINTER(c_void_p), _LL, _LL, _F, _I, _I, _P]),
    "dpb_allreduce_sgd": (c_int, [_P, _I, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p), _LL, _LL, _F, _F, _F, _F, _I, _I, _I, _P]),
    "dpb_gather_mean_scatter": (c_int, [_P, _I, POINTER(c_void_p), _LL, _LL, _F, _I, _I, _P]),
    "dpb_broadcast": (c_int, [_P, _I, POINTER(c_void_p), _LL, _LL, _I, _I, _P]),
    "dpb_barrier": (c_int, [_P, _I, _P]),
    "dpb_plan_buckets": (c_int, [_I, POINTER(c_longlong), _LL, _LL, POINTER(c_int)]),
    "dpb_reducer_create": (c_int, [_P, _I, POINTER(c_longlong), POINTER(c_longlong), POINTER(c_int), _I, POINTER(c_void_p), _F, POINTER(c_void_p)]),
    "dpb_reducer_destroy": (c_int, [_P]),
    "dpb_reducer_prepare": (c_int, [_P]),
    "dpb_reducer_mark_ready": (c_int, [_P, _I, _P]),
    "dpb_reducer_wait_stream": (c_int, [_P, _P]),
    "dpb_reducer_finalize": (c_int, [_P, _P]),
    "dpb_reducer_launched": (c_int, [_P]),
}


def _bind(lib: ctypes.CDLL) -> None:
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def load(build_if_missing: bool = True):
    """Return the ctypes handle, or ``None`` if the library cannot be loaded (reason in ``load_error()``)."""
    global _lib, _load_error
    with _lock:
        if _lib is not None:
            return _lib
        try:
            if not os.path.exists(LIB_PATH) and build_if_missing:
                build()
            lib = ctypes.CDLL(LIB_PATH)
            _bind(lib)
            for key, env in ((0, "DPB_CONV_TARGET_CTAS"), (1, "DPB_CONV_NT"), (2, "DPB_COMM_BLOCKS"), (3, "DPB_CONV_MIN_K_ITERS"), (4, "DPB_PDL"), (5, "DPB_CONV_CLUSTER"), (6, "DPB_CONV_PAIR"), (7, "DPB_FIRST_TC"), (8, "DPB_PAIR_REDUCE"), (9, "DPB_TAIL_MAX_PIXELS"), (10, "DPB_CONV_PERSIST"), (11, "DPB_COMM_WEAK")):
                if os.environ.get(env):
                    lib.dpb_set_tuning(key, int(os.environ[env]))
            _lib = lib
            _load_error = None
        except Exception as exc:  # noqa: BLE001 - reported through load_error()/require()
            _load_error = f"{type(exc).__name__}: {exc}"
            _lib = None
        return _lib


def load_error() -> str | None:
    return _load_error


def require():
    """The native library, or a loud failure. Every CUDA code path goes through this."""
    lib = load()
    if lib is None:
        raise RuntimeError(
            "distributed_pytorch_b200: native library libdpb200.so is not available "
            f"({_load_error}). Build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "There is no PyTorch fallback for CUDA tensors.")
    return lib


class NativeError(RuntimeError):
    pass


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        lib = require()
        msg = lib.dpb_last_error()
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg.decode() if msg else ''}")


# launch accounting (bench.py reports how many of our kernels ran inside the timed region)
_launches = 0


def count_launches(n: int = 1) -> None:
    global _launches
    _launches += n


def launches() -> int:
    return _launches


def ptr(t) -> int | None:
    """Device pointer of a tensor (None passes NULL)."""
    return None if t is None else t.data_ptr()


def stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
