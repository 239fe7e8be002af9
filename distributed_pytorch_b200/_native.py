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

ABI_VERSION = 202  # must equal dpb_version() of the loaded library (bumped whenever the C ABI changes)

_lock = threading.Lock()
_lib = None
_load_error: str | None = None


def _sources() -> list[str]:
    out = [os.path.join(_ROOT, "include", "dpb200.h")]
    for name in sorted(os.listdir(_CSRC)):
        if name.endswith((".cu", ".cuh", ".h")):
            out.append(os.path.join(_CSRC, name))
    return out


HASH_PATH = LIB_PATH + ".srchash"


def _source_hash() -> str:
    import hashlib
    h = hashlib.sha256()
    for path in _sources():
        h.update(os.path.basename(path).encode())
        with open(path, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def needs_build() -> bool:
    """True if the library is missing or was built from other sources.  Compared by content hash (kept next to the
    library), not by mtime: the tree may have been copied to another machine in any order."""
    if not os.path.exists(LIB_PATH) or not os.path.exists(HASH_PATH):
        return True
    want = _source_hash() + ("+experiments" if os.environ.get("DPB_EXPERIMENTS") == "1" else "")
    try:
        with open(HASH_PATH) as fh:
            return fh.read().strip() != want
    except OSError:
        return True


def build(force: bool = False, verbose: bool = False, experiments: bool = False) -> str:
    """Compile csrc/dpb200.cu for sm_100a into the in-tree shared library. Returns the library path.

    Several ranks may get here at once on a fresh checkout (torchrun starts them together): the compile is serialised
    with a file lock, goes to a temporary file and is moved into place atomically, so no process can dlopen a
    half-written library; whoever gets the lock second finds the library up to date and returns.
    ``experiments`` adds ``-DDPB_EXPERIMENTS`` (in-kernel tracing used by tools/conv_timing.py)."""
    import fcntl
    if not force and not needs_build():
        return LIB_PATH
    os.makedirs(_LIBDIR, exist_ok=True)
    with open(os.path.join(_LIBDIR, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not needs_build():
                return LIB_PATH
            nvcc = os.environ.get("NVCC", "nvcc")
            tmp = f"{LIB_PATH}.tmp.{os.getpid()}"
            cmd = [nvcc, *NVCC_FLAGS, "-o", tmp, os.path.join(_CSRC, "dpb200.cu")]
            if experiments or os.environ.get("DPB_EXPERIMENTS") == "1":
                cmd.insert(1, "-DDPB_EXPERIMENTS")
            if verbose:
                cmd.insert(1, "-Xptxas")
                cmd.insert(2, "-v")
            proc = subprocess.run(cmd, capture_output=True, text=True)
            if proc.returncode != 0:
                if os.path.exists(tmp):
                    os.unlink(tmp)
                raise RuntimeError(f"nvcc failed ({' '.join(cmd)}):\n{proc.stdout}\n{proc.stderr}")
            os.replace(tmp, LIB_PATH)
            with open(HASH_PATH + ".tmp", "w") as fh:
                fh.write(_source_hash() + ("+experiments" if "-DDPB_EXPERIMENTS" in cmd else ""))
            os.replace(HASH_PATH + ".tmp", HASH_PATH)
            if verbose:
                print(proc.stderr)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB_PATH


_P = c_void_p
_I = c_int
_LL = c_longlong
_F = c_float

_SIGNATURES = {
    "dpb_version": (c_int, []),
    "dpb_last_error": (c_char_p, []),
    "dpb_init_device": (c_int, []),
    "dpb_last_device_error": (c_int, [_I, POINTER(c_int)]),
    "dpb_debug_set_conv_dump": (c_int, [_P]),
    "dpb_debug_set_trace": (c_int, [_P]),
    "dpb_conv_workspace_bytes": (c_size_t, []),
    "dpb_conv3x3_gemm": (c_int, [_I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_first_conv_uses_tensor_cores": (c_int, [_I, _I]),
    "dpb_bn_relu_pool_fwd_train": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _F, _F, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dpb_conv_last_launches": (c_int, []),
    "dpb_set_tuning": (c_int, [_I, _I]),
    "dpb_first_conv_blocks": (c_int, [_I, _I, _I]),
    "dpb_first_conv_fprop": (c_int, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_first_conv_wgrad_blocks": (c_int, []),
    "dpb_first_conv_wgrad": (c_int, [_P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_bn_finalize": (c_int, [_P, _P, _I, _I, c_double, _P, _P, _P, _P, _P, _F, _F, _P, _P, _P, _P, _P]),
    "dpb_bn_eval_scale_shift": (c_int, [_I, _P, _P, _P, _P, _F, _P, _P, _P]),
    "dpb_bn_relu_pool_fwd": (c_int, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dpb_bn_bwd_blocks": (c_int, []),
    "dpb_bn_bwd_launches": (c_int, [_I, _I]),
    "dpb_bn_relu_pool_bwd": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
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
    "dpb_head_supported": (c_int, [_I, _I]),
    "dpb_head_fwd": (c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_head_bwd": (c_int, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dpb_augment_batch": (c_int, [_P, _P, _P, _I, _I, _I, _I, ctypes.c_uint, ctypes.c_uint, _P, _F, _F, _F, _F, _F, _F, _P, _P, _P]),
    "dpb_sgd_step": (c_int, [_P, _P, _P, _LL, _F, _F, _F, _F, _I, _P]),
    "dpb_symm_alloc": (c_int, [c_size_t, POINTER(c_void_p)]),
    "dpb_symm_free": (c_int, [_P]),
    "dpb_symm_export": (c_int, [_P, ctypes.c_char_p]),
    "dpb_symm_import": (c_int, [ctypes.c_char_p, POINTER(c_void_p)]),
    "dpb_symm_close": (c_int, [_P]),
    "dpb_signal_pad_bytes": (c_size_t, []),
    "dpb_vmm_probe": (c_int, [_I, POINTER(c_longlong)]),
    "dpb_vmm_alloc": (c_int, [c_size_t, c_size_t, POINTER(c_void_p), POINTER(ctypes.c_ulonglong), POINTER(c_int)]),
    "dpb_vmm_import": (c_int, [_I, c_size_t, c_size_t, _I, POINTER(c_void_p), POINTER(ctypes.c_ulonglong)]),
    "dpb_vmm_unmap": (c_int, [_P, c_size_t, ctypes.c_ulonglong]),
    "dpb_mc_create": (c_int, [_I, c_size_t, POINTER(ctypes.c_ulonglong), POINTER(c_int)]),
    "dpb_mc_add_device": (c_int, [ctypes.c_ulonglong]),
    "dpb_mc_bind_map": (c_int, [ctypes.c_ulonglong, ctypes.c_ulonglong, c_size_t, c_size_t, POINTER(c_void_p)]),
    "dpb_mc_unbind": (c_int, [ctypes.c_ulonglong, c_size_t]),
    "dpb_comm_debug_withhold": (c_int, [_P, _I]),
    "dpb_allreduce_nvls": (c_int, [_P, _I, _P, _P, _P, _P, _LL, _LL, _F, _F, _F, _F, _I, _I, _I, _P]),
    "dpb_nvls_enabled": (c_int, []),
    "dpb_reducer_set_sgd": (c_int, [_P, POINTER(c_void_p), POINTER(c_void_p), _F, _F, _F, _I]),
    "dpb_reducer_set_multicast": (c_int, [_P, _P, _P, _P, _P]),
    "dpb_reducer_side_stream": (c_int, [_P, POINTER(c_void_p)]),
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


# dpb_set_tuning keys and the environment variables that set them at load time (include/dpb200.h)
TUNING_ENV = ((0, "DPB_CONV_TARGET_CTAS"), (1, "DPB_CONV_NT"), (2, "DPB_COMM_BLOCKS"), (3, "DPB_CONV_MIN_K_ITERS"),
              (4, "DPB_PDL"), (5, "DPB_CONV_CLUSTER_MAX"), (7, "DPB_FIRST_TC"), (11, "DPB_COMM_WEAK"),
              (13, "DPB_DDP_BLOCKS"), (14, "DPB_NVLS"), (15, "DPB_NVLS_BLOCKS"), (16, "DPB_CONV_CLUSTER_REACH"))


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
            # a library older than its sources (a stale git-ignored .so from another checkout) is rebuilt, never loaded
            if build_if_missing and needs_build():
                build()
            lib = ctypes.CDLL(LIB_PATH)
            lib.dpb_version.restype = c_int
            if lib.dpb_version() != ABI_VERSION:
                if not build_if_missing:
                    raise RuntimeError(f"libdpb200.so has ABI version {lib.dpb_version()}, this package needs {ABI_VERSION}")
                build(force=True)
                lib = ctypes.CDLL(LIB_PATH)
                lib.dpb_version.restype = c_int
                if lib.dpb_version() != ABI_VERSION:
                    raise RuntimeError(f"rebuilt libdpb200.so still reports ABI {lib.dpb_version()} (expected {ABI_VERSION})")
            _bind(lib)
            for key, env in TUNING_ENV:
                if os.environ.get(env):
                    if lib.dpb_set_tuning(key, int(os.environ[env])) != 0 and key != 12:
                        raise RuntimeError(f"{env}={os.environ[env]} rejected: {lib.dpb_last_error().decode()}")
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


_initialised_devices: set[int] = set()


def init_device(index: int | None = None) -> None:
    """Register the device-error record and the peer time-out on a CUDA device (once per device and process)."""
    import torch
    idx = torch.cuda.current_device() if index is None else index
    if idx in _initialised_devices:
        return
    lib = require()
    with torch.cuda.device(idx):
        if os.environ.get("DPB_PEER_TIMEOUT_MS"):
            lib.dpb_set_tuning(12, int(os.environ["DPB_PEER_TIMEOUT_MS"]))
        rc = lib.dpb_init_device()
    if rc != 0:
        raise NativeError(f"dpb_init_device failed (rc={rc}): {lib.dpb_last_error().decode()}")
    _initialised_devices.add(idx)


def device_error_message(device: int = -1) -> str | None:
    """What a kernel of ours recorded before it gave up (None if nothing).  Works after the CUDA context has become
    unusable, because the record lives in host memory."""
    lib = load(build_if_missing=False)
    if lib is None:
        return None
    out = (c_int * 6)()
    lib.dpb_last_device_error(device, out)
    code, where, block, aux0, aux1, rank = list(out)
    if code == 0:
        return None
    if code == 2:
        return (f"peer-signal timeout: rank {rank} waited for rank {aux0}, channel {where}, block {block}, "
                f"epoch {aux1} (a peer never reached the collective: dead, hung or not launched)")
    if code == 1:
        return f"pipeline (mbarrier) timeout in a compute kernel: site {where}, block {block}, parity {aux0}"
    return f"device error code {code}: where={where} block={block} aux=({aux0},{aux1})"


def raise_if_device_error(exc: BaseException | None = None, device: int = -1) -> None:
    """Turn a bare CUDA failure into the diagnostic our kernel left behind, if there is one."""
    msg = device_error_message(device)
    if msg is not None:
        raise NativeError(msg) from exc


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        lib = require()
        msg = lib.dpb_last_error()
        dev = device_error_message()
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg.decode() if msg else ''}"
                          + (f" [{dev}]" if dev else ""))


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
