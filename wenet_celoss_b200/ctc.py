"""CTC module and loss with the reference's interface, backed by sm_100a kernels.

``CTC`` mirrors wenet/transformer/ctc.py:21-84 (same constructor, same parameter
name ``ctc_lo`` so checkpoints load, same ``forward`` / ``log_softmax`` / ``argmax``).
``forward`` keeps the ``ctc_lo`` projection as a library GEMM (ctc.py:57) and
replaces ``log_softmax`` + ``torch.nn.CTCLoss(reduction='sum')`` (ctc.py:60-61) by
three kernels working directly on the batch-major ``[B,T,V]`` logits.
"""
from __future__ import annotations

import types

import torch
from torch.amp import custom_bwd, custom_fwd
import torch.nn.functional as F

from . import _lib
from ._lib import i32, ptr


class _CtcLossFn(torch.autograd.Function):
    @staticmethod
    @custom_fwd(device_type="cuda")
    def forward(ctx, logits, ys_pad, hlens, ys_lens, blank):
        B, T, V = logits.shape
        Lmax = ys_pad.size(1)
        Smax = 2 * Lmax + 1
        dev = logits.device
        with torch.cuda.device(dev):
            st = _lib.stream_handle(dev)
            lse = torch.empty(B, T, dtype=torch.float32, device=dev)
            lat = torch.empty(3, B, T, Smax, dtype=torch.float32, device=dev)
            lp_ext, alpha, beta = lat.unbind(0)
            nll = torch.empty(B, dtype=torch.float32, device=dev)
            leader = torch.empty(B, Smax, dtype=torch.int32, device=dev)
            frame_off = torch.zeros(B, 2, T + 1, dtype=torch.float64, device=dev)
            dt = i32(_lib.dtype_code(logits))
            _lib.call("b200_ctc_logprobs", ptr(logits), dt, ptr(ys_pad), ptr(hlens), ptr(ys_lens),
                      i32(B), i32(T), i32(V), i32(Lmax), i32(blank), ptr(lse), ptr(lp_ext), st)
            _lib.call("b200_ctc_lattice", ptr(lp_ext), ptr(ys_pad), ptr(hlens), ptr(ys_lens), i32(B),
                      i32(T), i32(Lmax), i32(blank), ptr(alpha), ptr(beta), ptr(nll), ptr(leader),
                      ptr(frame_off), st)
        ctx.save_for_backward(logits, ys_pad, hlens, ys_lens, lse, lp_ext, alpha, beta, frame_off,
                              leader)
        ctx.blank = blank
        return nll

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, dy):
        logits, ys_pad, hlens, ys_lens, lse, lp_ext, alpha, beta, frame_off, leader = ctx.saved_tensors
        B, T, V = logits.shape
        Lmax = ys_pad.size(1)
        dev = logits.device
        with torch.cuda.device(dev):
            scale = dy.detach().to(torch.float32).contiguous()
            grad = torch.empty_like(logits)
            _lib.call("b200_ctc_grad", ptr(logits), i32(_lib.dtype_code(logits)), ptr(lse),
                      ptr(lp_ext), ptr(alpha), ptr(beta), ptr(frame_off), ptr(scale), ptr(ys_pad),
                      ptr(hlens), ptr(ys_lens), ptr(leader), i32(B), i32(T), i32(V), i32(Lmax),
                      i32(ctx.blank), ptr(grad), _lib.stream_handle(dev))
        return grad, None, None, None, None


def ctc_loss(logits: torch.Tensor, ys_pad: torch.Tensor, hlens: torch.Tensor, ys_lens: torch.Tensor,
             blank: int = 0, reduction: str = "sum") -> torch.Tensor:
    """Negative log-likelihood per utterance over pre-softmax ``logits [B,T,V]``.

    ``ys_pad [B,Lmax]`` is the reference's padded target tensor (int64, pad -1);
    infeasible alignments give ``inf`` (zero_infinity=False, as ctc.py:44)."""
    if reduction not in ("none", "sum", "mean"):
        raise ValueError(f"{reduction} is not a valid value for reduction")
    _lib.require_cuda(logits, "logits")
    if logits.dim() != 3:
        raise RuntimeError("logits must be 3-D (batch, time, class)")
    if logits.dtype not in (torch.float32, torch.float16, torch.bfloat16):
        raise RuntimeError("logits must be float32, float16 or bfloat16")
    B = logits.size(0)
    if ys_pad.dim() != 2 or ys_pad.size(0) != B or hlens.numel() != B or ys_lens.numel() != B:
        raise RuntimeError("batch dimension mismatch between logits, ys_pad, hlens and ys_lens")
    dev = logits.device
    logits = logits.contiguous()
    ys = ys_pad.to(device=dev, dtype=torch.int64).contiguous()
    hl = hlens.to(device=dev, dtype=torch.int32).contiguous()
    yl = ys_lens.to(device=dev, dtype=torch.int32).contiguous()
    # lengths beyond the tensors would run the trellis kernels out of bounds; torch.nn.CTCLoss raises too
    mx = torch.stack([hl.max(), yl.max(), -hl.min(), -yl.min()]).tolist()
    if mx[0] > logits.size(1) or mx[2] > 0:
        raise RuntimeError(f"Expected input_lengths to have value at most {logits.size(1)}, but got value "
                           f"{mx[0] if mx[0] > logits.size(1) else -mx[2]} (while checking arguments for ctc_loss)")
    if mx[1] > ys.size(1) or mx[3] > 0:
        raise RuntimeError(f"Expected target_lengths to have value at most {ys.size(1)}, but got value "
                           f"{mx[1] if mx[1] > ys.size(1) else -mx[3]} (while checking arguments for ctc_loss)")
    nll = _CtcLossFn.apply(logits, ys, hl, yl, int(blank))
    if reduction == "sum":
        return nll.sum()
    if reduction == "mean":  # torch semantics: per-utterance nll / target length, then batch mean
        return (nll / yl.clamp_min(1).to(nll.dtype)).mean()
    return nll


def _ctc_forward(self, hs_pad: torch.Tensor, hlens: torch.Tensor, ys_pad: torch.Tensor,
                 ys_lens: torch.Tensor) -> torch.Tensor:
    """CTC.forward (ctc.py:46-64): ``sum_b nll_b / B`` (or ``nll / B`` when reduce=False)."""
    ys_hat = self.ctc_lo(F.dropout(hs_pad, p=self.dropout_rate))  # [B,T,V], library GEMM
    reduce = getattr(self, "reduce", None)
    if reduce is None:  # a reference CTC instance: recover the flag from its criterion
        reduce = getattr(getattr(self, "ctc_loss", None), "reduction", "sum") == "sum"
    loss = ctc_loss(ys_hat, ys_pad, hlens, ys_lens, blank=0, reduction="sum" if reduce else "none")
    return loss / ys_hat.size(0)


class CTC(torch.nn.Module):
    """CTC module (same constructor and parameter names as the reference's)."""

    def __init__(self, odim: int, encoder_output_size: int, dropout_rate: float = 0.0,
                 reduce: bool = True):
        super().__init__()
        self.dropout_rate = dropout_rate
        self.reduce = reduce
        self.ctc_lo = torch.nn.Linear(encoder_output_size, odim)

    forward = _ctc_forward

    def log_softmax(self, hs_pad: torch.Tensor) -> torch.Tensor:
        return F.log_softmax(self.ctc_lo(hs_pad), dim=2)

    def argmax(self, hs_pad: torch.Tensor) -> torch.Tensor:
        return torch.argmax(self.ctc_lo(hs_pad), dim=2)


def patch_ctc(module: torch.nn.Module) -> torch.nn.Module:
    """Swap the forward of an existing reference ``CTC`` instance (anything with
    ``ctc_lo`` and ``dropout_rate``) for the sm_100a path; parameters untouched."""
    if not hasattr(module, "ctc_lo"):
        raise TypeError("patch_ctc expects a wenet CTC module (ctc_lo)")
    if not hasattr(module, "dropout_rate"):
        module.dropout_rate = 0.0  # the reference default (ctc.py:26)
    module.forward = types.MethodType(_ctc_forward, module)
    return module
