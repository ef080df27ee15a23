"""ctypes binding of libb200_transducer.so (the C ABI declared in include/b200_transducer.h).

There is no CPU fallback: if the shared library is missing or a call fails, the
wrappers raise.  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes
from ctypes import c_char_p, c_float, c_int, c_size_t, c_uint64, c_void_p, c_double
from pathlib import Path

import torch

import os

_PKG = Path(__file__).resolve().parent
# B200_LIB: load another build of the same library (kernel experiments, scripts/grad_variants.sh)
LIB_PATH = Path(os.environ["B200_LIB"]) if os.environ.get("B200_LIB") else _PKG / "libb200_transducer.so"

F32, F16, BF16 = 0, 1, 2
_DTYPES = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16}

_lib = None


class B200Error(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Load the library once; fail loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise B200Error(
                f"{LIB_PATH} not found: build it with `python -m wenet_celoss_b200.build` "
                "(there is no CPU or PyTorch fallback for this path)")
        _lib = ctypes.CDLL(str(LIB_PATH))
        _declare(_lib)
    return _lib


def _declare(L: ctypes.CDLL) -> None:
    L.b200_version.restype = c_int
    L.b200_last_error.restype = c_char_p
    L.b200_compiled_arch.restype = c_int
    L.b200_launch_count.restype = c_uint64
    # every compute entry point returns int status; arguments are passed as
    # c_void_p / c_int / c_float explicitly at the call sites via _call().


def dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise B200Error(f"unsupported dtype {t.dtype}") from None


def ptr(t) -> c_void_p:
    if t is None:
        return c_void_p(0)
    return c_void_p(t.data_ptr())


def stream_handle(device=None) -> c_void_p:
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


# Measurement hook (bench.py, scripts/): while TRACE is a list, every entry-point call is bracketed by
# two CUDA events on the current stream - the stream the wrappers hand to the kernels - and
# (name, start, end) is appended, so a caller can read per-kernel times INSIDE a real step.
TRACE = None


def call(fn_name: str, *args) -> None:
    """Invoke an int-returning entry point and raise on a non-zero status."""
    L = lib()
    fn = getattr(L, fn_name)
    fn.restype = c_int
    trace = TRACE
    if trace is not None:
        ev0 = torch.cuda.Event(enable_timing=True)
        ev0.record()
    rc = fn(*args)
    if trace is not None:
        ev1 = torch.cuda.Event(enable_timing=True)
        ev1.record()
        trace.append((fn_name, ev0, ev1))
    if rc != 0:
        msg = L.b200_last_error().decode("utf-8", "replace")
        raise B200Error(f"{fn_name} failed (status {rc}): {msg}")


def launch_count() -> int:
    return int(lib().b200_launch_count())


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise B200Error(f"{name} must be a CUDA tensor (no CPU path in wenet_celoss_b200)")


i32 = c_int
f32 = c_float
f64 = c_double
size_t = c_size_t


# ---- host-side validation reads -------------------------------------------------------------
# The reference kernels (torchaudio rnnt_loss, torch.nn.CTCLoss) validate the length tensors on the
# host, which costs one device-to-host read - and a stream synchronisation - per call.  The result is
# remembered per (storage, version) of the tensors it was computed from, so that a batch that is used
# again (same tensors, not modified in place since) is not read back a second time.
_HOST_READS: list = []  # [(weak references, versions, result)], most recent last


def cached_host_read(tensors, compute):
    """``compute() -> list`` (does the device-to-host read), memoised on the IDENTITY of the tensor
    objects (weak references: a new tensor that happens to reuse the storage address is a miss) and on
    their version counters (an in-place update is a miss)."""
    import weakref
    for refs, versions, result in reversed(_HOST_READS):
        if len(refs) == len(tensors) and all(r() is t and t._version == v
                                             for r, t, v in zip(refs, tensors, versions)):
            return result
    result = compute()
    if len(_HOST_READS) >= 8:
        del _HOST_READS[0]
    _HOST_READS.append((tuple(weakref.ref(t) for t in tensors), tuple(t._version for t in tensors), result))
    return result


_CONVERTED: list = []  # [(weak reference, version, device, dtype, converted tensor)]


def cached_convert(t, device, dtype):
    """``t.to(device, dtype).contiguous()`` - remembered per tensor object and version when it needs a copy
    (length tensors given as int64 are converted to the kernels' int32 once per batch, not once per call)."""
    import weakref
    if t.device == device and t.dtype == dtype and t.is_contiguous():
        return t
    for ref, ver, dev, dt, out in reversed(_CONVERTED):
        if ref() is t and t._version == ver and dev == device and dt == dtype:
            return out
    out = t.to(device=device, dtype=dtype).contiguous()
    if len(_CONVERTED) >= 8:
        del _CONVERTED[0]
    _CONVERTED.append((weakref.ref(t), t._version, device, dtype, out))
    return out
