"""Drop-in ``rnnt_loss`` with the torchaudio signature, backed by sm_100a kernels.

Mirrors ``torchaudio.functional.rnnt_loss(logits, targets, logit_lengths,
target_lengths, blank, clamp, reduction, fused_log_softmax)`` as called by the
reference at wenet/transducer/transducer.py:142-147 (``reduction="mean"``) and
:296-301 (``reduction="none"``): same argument meaning, same validation errors,
same result dtype, gradient w.r.t. ``logits`` only.  Extension: bf16 logits.

Unlike torchaudio the gradient is not materialised in forward: backward re-reads
the logits and writes the gradient once with ``dy`` folded in (3*N*s bytes of HBM
traffic per fwd+bwd instead of >= 6*N*s).
"""
from __future__ import annotations

import torch
from torch.amp import custom_bwd, custom_fwd

from . import _lib
from ._lib import c_void_p, f32, i32, ptr


def _validate(logits, targets, logit_lengths, target_lengths, blank):
    # same checks, same wording as torchaudio's compute.cpp (observed messages)
    if logits.dtype not in (torch.float32, torch.float16, torch.bfloat16):
        raise RuntimeError("logits must be float32 or float16 (half) type (bfloat16 also accepted here)")
    if targets.dtype != torch.int32:
        raise RuntimeError("targets must be int32 type")
    if logit_lengths.dtype != torch.int32:
        raise RuntimeError("logit_lengths must be int32 type")
    if target_lengths.dtype != torch.int32:
        raise RuntimeError("target_lengths must be int32 type")
    if not logits.is_contiguous():
        raise RuntimeError("logits must be contiguous")
    if not targets.is_contiguous():
        raise RuntimeError("targets must be contiguous")
    if not logit_lengths.is_contiguous():
        raise RuntimeError("logit_lengths must be contiguous")
    if not target_lengths.is_contiguous():
        raise RuntimeError("target_lengths must be contiguous")
    if logits.dim() != 4:
        raise RuntimeError("logits must be 4-D (batch, time, target, class)")
    if targets.dim() != 2:
        raise RuntimeError("targets must be 2-D (batch, max target length)")
    if logit_lengths.dim() != 1:
        raise RuntimeError("logit_lengths must be 1-D")
    if target_lengths.dim() != 1:
        raise RuntimeError("target_lengths must be 1-D")
    B = logits.size(0)
    if logit_lengths.size(0) != B:
        raise RuntimeError("batch dimension mismatch between logits and logit_lengths")
    if target_lengths.size(0) != B:
        raise RuntimeError("batch dimension mismatch between logits and target_lengths")
    if targets.size(0) != B:
        raise RuntimeError("batch dimension mismatch between logits and targets")
    if not (0 <= blank < logits.size(-1)):
        raise RuntimeError("blank must be within [0, logits.shape[-1])")
    for name, t in (("logits", logits), ("targets", targets), ("logit_lengths", logit_lengths),
                    ("target_lengths", target_lengths)):
        _lib.require_cuda(t, name)
    # one small D2H read, as torchaudio does for the same two checks
    mx = _lib.cached_host_read((logit_lengths, target_lengths),
                               lambda: torch.stack([logit_lengths.max(), target_lengths.max()]).tolist())
    if mx[0] != logits.size(1):
        raise RuntimeError("input length mismatch")
    if mx[1] != targets.size(1) or logits.size(2) != targets.size(1) + 1:
        raise RuntimeError("output length mismatch")


def rnnt_lattice(lp_blank, lp_emit, logit_lengths, target_lengths):
    """Rescaled alpha, beta [B,T,U1], costs [B] and the fp64 per-diagonal offsets
    [B,2,T+U1] (see include/b200_transducer.h) from per-cell blank / emit log-probs."""
    B, T, U1 = lp_blank.shape
    alpha = torch.empty_like(lp_blank)
    beta = torch.empty_like(lp_blank)
    costs = torch.empty(B, dtype=torch.float32, device=lp_blank.device)
    diag_off = torch.empty(B, 2, T + U1, dtype=torch.float64, device=lp_blank.device)
    _lib.call("b200_rnnt_lattice", ptr(lp_blank), ptr(lp_emit), ptr(logit_lengths),
              ptr(target_lengths), i32(B), i32(T), i32(U1), ptr(alpha), ptr(beta), ptr(costs),
              ptr(diag_off), _lib.stream_handle(lp_blank.device))
    return alpha, beta, costs, diag_off


def rnnt_grad_coef(lse, alpha, beta, diag_off, dy, targets, logit_lengths, target_lengths, blank):
    """float4 per lattice cell: {c0, cb, ce, scale} (see include/b200_transducer.h)."""
    B, T, U1 = lse.shape
    coef = torch.empty(B, T, U1, 4, dtype=torch.float32, device=lse.device)
    _lib.call("b200_rnnt_grad_coef", ptr(lse), ptr(alpha), ptr(beta), ptr(diag_off), ptr(dy),
              ptr(targets), ptr(logit_lengths), ptr(target_lengths), i32(B), i32(T), i32(U1),
              i32(blank), ptr(coef), _lib.stream_handle(lse.device))
    return coef


class _RnntLossFn(torch.autograd.Function):
    @staticmethod
    @custom_fwd(device_type="cuda")
    def forward(ctx, logits, targets, logit_lengths, target_lengths, blank, clamp):
        B, T, U1, V = logits.shape
        dev = logits.device
        with torch.cuda.device(dev):
            st = _lib.stream_handle(dev)
            lat = torch.empty(5, B, T, U1, dtype=torch.float32, device=dev)
            lse, lpb, lpe, alpha, beta = lat.unbind(0)
            costs = torch.empty(B, dtype=torch.float32, device=dev)
            diag_off = torch.empty(B, 2, T + U1, dtype=torch.float64, device=dev)
            _lib.call("b200_rnnt_loss_fwd", ptr(logits), i32(_lib.dtype_code(logits)), ptr(targets),
                      ptr(logit_lengths), ptr(target_lengths), i32(B), i32(T), i32(U1), i32(V),
                      i32(blank), ptr(costs), ptr(lse), ptr(lpb), ptr(lpe), ptr(alpha), ptr(beta),
                      ptr(diag_off), st)
        ctx.save_for_backward(logits, targets, logit_lengths, target_lengths, lse, alpha, beta,
                              diag_off)
        ctx.blank = blank
        ctx.clamp = clamp
        # torchaudio returns costs in the logits dtype (f32/f16); bf16 (extension) stays f32
        return costs.to(torch.float16) if logits.dtype == torch.float16 else costs

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, dy):
        logits, targets, logit_lengths, target_lengths, lse, alpha, beta, diag_off = ctx.saved_tensors
        B, T, U1, V = logits.shape
        dev = logits.device
        with torch.cuda.device(dev):
            dyf = dy.detach().to(torch.float32).contiguous()
            coef = rnnt_grad_coef(lse, alpha, beta, diag_off, dyf, targets, logit_lengths,
                                  target_lengths, ctx.blank)
            grad = torch.empty_like(logits)
            _lib.call("b200_rnnt_grad", ptr(logits), i32(_lib.dtype_code(logits)), ptr(targets),
                      ptr(coef), i32(B), i32(T), i32(U1), i32(V), i32(ctx.blank), f32(ctx.clamp),
                      ptr(grad), _lib.stream_handle(dev))
        return grad, None, None, None, None, None


def rnnt_loss(logits: torch.Tensor,
              targets: torch.Tensor,
              logit_lengths: torch.Tensor,
              target_lengths: torch.Tensor,
              blank: int = -1,
              clamp: float = -1,
              reduction: str = "mean",
              fused_log_softmax: bool = True) -> torch.Tensor:
    """RNN-T loss (Graves 2012) over joint logits ``[B, T, U+1, V]``.

    Same contract as ``torchaudio.functional.rnnt_loss``; see module docstring.
    """
    if reduction not in ("none", "mean", "sum"):
        raise ValueError('reduction should be one of "none", "mean", or "sum"')
    if not fused_log_softmax:
        raise NotImplementedError(
            "fused_log_softmax=False (log-prob inputs) is not on the reference's path "
            "(wenet/transducer/transducer.py:142 uses the default) and is not implemented")
    if blank < 0:
        blank = logits.shape[-1] + blank
    _validate(logits, targets, logit_lengths, target_lengths, blank)
    costs = _RnntLossFn.apply(logits, targets, logit_lengths, target_lengths, int(blank),
                              float(clamp))
    if reduction == "mean":
        return costs.mean()
    if reduction == "sum":
        return costs.sum()
    return costs
