"""oracle/c_oracle.py — TEST INFRASTRUCTURE.  ctypes front-end of oracle/lattice_ref.c.

Builds oracle/_build/liboracle_ref.so with gcc on first use.  Only tests/,
__graft_entry__.smoke() and bench.py's CPU legs import this.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
_SRC = _DIR / "lattice_ref.c"
_SO = _DIR / "_build" / "liboracle_ref.so"
_lib = None


def build(force: bool = False) -> Path:
    _SO.parent.mkdir(exist_ok=True)
    if force or not _SO.exists() or _SO.stat().st_mtime < _SRC.stat().st_mtime:
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-o", str(_SO), str(_SRC), "-lm"],
                       check=True)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(str(_SO))
    return _lib


def _np(x, dtype):
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.ascontiguousarray(x, dtype=dtype)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else ctypes.c_void_p(0)


def rnnt_loss_c(logits, targets, logit_len, target_len, blank=0, dy=None, want_grad=True,
                nthreads=0):
    """-> costs [B] f64, grad [B,T,U1,V] f32 (or None), alpha, beta [B,T,U1] f64."""
    x = _np(logits, np.float32)
    B, T, U1, V = x.shape
    tg = _np(targets, np.int32).reshape(B, U1 - 1)
    tl = _np(logit_len, np.int32)
    ul = _np(target_len, np.int32)
    dyv = None if dy is None else _np(dy, np.float64)
    costs = np.empty(B, dtype=np.float64)
    alpha = np.empty((B, T, U1), dtype=np.float64)
    beta = np.empty((B, T, U1), dtype=np.float64)
    grad = np.empty_like(x) if want_grad else None
    rc = lib().oracle_rnnt_loss(_p(x), _p(tg), _p(tl), _p(ul), B, T, U1, V, int(blank), _p(dyv),
                                _p(costs), _p(alpha), _p(beta), _p(grad), int(nthreads))
    if rc != 0:
        raise RuntimeError("oracle_rnnt_loss: bad arguments")
    return costs, grad, alpha, beta


def ctc_loss_c(logits, ys_pad, hlens, ys_lens, blank=0, scale=1.0, want_grad=True, nthreads=0):
    """-> nll [B] f64, grad [B,T,V] f32 of scale*sum(nll) w.r.t. the (pre-softmax) logits."""
    x = _np(logits, np.float32)
    B, T, V = x.shape
    ys = _np(ys_pad, np.int64)
    Lmax = ys.shape[1] if ys.ndim == 2 else 0
    ys = ys.reshape(B, Lmax)
    hl = _np(hlens, np.int32)
    yl = _np(ys_lens, np.int32)
    nll = np.empty(B, dtype=np.float64)
    grad = np.empty_like(x) if want_grad else None
    fn = lib().oracle_ctc_loss
    fn.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int] * 5 + [ctypes.c_double] + \
        [ctypes.c_void_p] * 2 + [ctypes.c_int]
    rc = fn(_p(x), _p(ys), _p(hl), _p(yl), B, T, V, Lmax, int(blank), float(scale), _p(nll), _p(grad),
            int(nthreads))
    if rc != 0:
        raise RuntimeError("oracle_ctc_loss: bad arguments")
    return nll, grad


def cpu_threads() -> int:
    return os.cpu_count() or 1
