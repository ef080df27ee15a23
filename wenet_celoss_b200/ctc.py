"""CTC module and loss with the reference's interface, backed by sm_100a kernels.

``CTC`` mirrors wenet/transformer/ctc.py:21-84 (same constructor, same parameter
name ``ctc_lo`` so checkpoints load, same ``forward`` / ``log_softmax`` / ``argmax``).
``forward`` fuses the ``ctc_lo`` projection (ctc.py:57), ``log_softmax`` and
``torch.nn.CTCLoss(reduction='sum')`` (ctc.py:60-61): the ``[B,T,V]`` logits are produced
tile by tile in tensor memory by the fp32-accurate tcgen05 GEMM of csrc/split_gemm.cu and
consumed by its softmax epilogues, the trellis reads exact fp32 dot products for the
2 L + 1 label states.  ``ctc_loss`` is the stand-alone loss over given logits (three
kernels working directly on the batch-major ``[B,T,V]`` tensor); encoder dims the fused
kernels do not tile (not 256 / 512) go through a library GEMM and ``ctc_loss``.
"""
from __future__ import annotations

import types

import torch
from torch.amp import custom_bwd, custom_fwd
import torch.nn.functional as F

from . import _lib
from ._lib import i32, ptr
from .joint import _i64, _pad_bias, split_f32


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


class _FusedCtcFn(torch.autograd.Function):
    """nll [B] from hs [B,T,E] and ctc_lo directly; differentiable w.r.t. hs, weight and bias."""

    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, hs, weight, bias, ys_pad, hlens, ys_lens, blank):
        B, T, E = hs.shape
        V = weight.shape[0]
        Lmax = ys_pad.size(1)
        Smax = 2 * Lmax + 1
        dev = hs.device
        with torch.cuda.device(dev):
            st = _lib.stream_handle(dev)
            hs32 = hs.detach().float().contiguous()
            w32 = weight.detach().float().contiguous()
            b32 = bias.detach().float().contiguous()
            hs_hi, hs_lo = split_f32(hs32.view(B * T, E))
            w_hi, w_lo = split_f32(w32)
            bias_pad = _pad_bias(b32, V)
            lse = torch.empty(B, T, dtype=torch.float32, device=dev)
            _lib.call("b200_split_gemm_lse", ptr(hs_hi), ptr(hs_lo), _i64(B * T), i32(E), i32(E), ptr(w_hi),
                      ptr(w_lo), i32(V), i32(E), ptr(bias_pad), ptr(lse), i32(0), st)
            lat = torch.empty(3, B, T, Smax, dtype=torch.float32, device=dev)
            lp_ext, alpha, beta = lat.unbind(0)
            _lib.call("b200_ctc_lp_ext", ptr(hs32), ptr(w32), ptr(b32), ptr(ys_pad), ptr(hlens), ptr(ys_lens),
                      i32(B), i32(T), i32(E), i32(V), i32(Lmax), i32(blank), ptr(lse), ptr(lp_ext), st)
            nll = torch.empty(B, dtype=torch.float32, device=dev)
            leader = torch.empty(B, Smax, dtype=torch.int32, device=dev)
            frame_off = torch.zeros(B, 2, T + 1, dtype=torch.float64, device=dev)
            _lib.call("b200_ctc_lattice", ptr(lp_ext), ptr(ys_pad), ptr(hlens), ptr(ys_lens), i32(B),
                      i32(T), i32(Lmax), i32(blank), ptr(alpha), ptr(beta), ptr(nll), ptr(leader),
                      ptr(frame_off), st)
        ctx.save_for_backward(hs_hi, hs_lo, w_hi, w_lo, bias_pad, ys_pad, hlens, ys_lens, lse, lp_ext, alpha,
                              beta, frame_off, leader)
        ctx.shape = (B, T, E, V, Lmax)
        ctx.blank = blank
        ctx.in_dtypes = (hs.dtype, weight.dtype, bias.dtype)
        return nll

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, dy):
        (hs_hi, hs_lo, w_hi, w_lo, bias_pad, ys_pad, hlens, ys_lens, lse, lp_ext, alpha, beta, frame_off,
         leader) = ctx.saved_tensors
        B, T, E, V, Lmax = ctx.shape
        dev = hs_hi.device
        rows = B * T
        with torch.cuda.device(dev):
            st = _lib.stream_handle(dev)
            scale = dy.detach().to(torch.float32).contiguous()
            # G = scale_b * softmax for t < hlens[b] (zero rows beyond) minus the state posteriors
            live = torch.arange(T, device=dev)[None, :] < hlens[:, None]
            coef = torch.zeros(B, T, 4, dtype=torch.float32, device=dev)
            coef[..., 0] = -lse
            coef[..., 3] = scale[:, None] * live
            ldg = (V + 15) // 16 * 16
            G_hi = torch.empty(rows, ldg, dtype=torch.bfloat16, device=dev)
            G_lo = torch.empty_like(G_hi)
            # the sparse part (state posteriors of blank and the labels) goes into the EXP epilogue: one pass over G
            nwords = (V + 31) // 32
            words = torch.empty(B, nwords, 2, dtype=torch.int32, device=dev)
            val = torch.empty(B, Lmax + 1, T, dtype=torch.float32, device=dev)
            _lib.call("b200_ctc_post", ptr(lp_ext), ptr(alpha), ptr(beta), ptr(frame_off), ptr(scale), ptr(ys_pad),
                      ptr(hlens), ptr(ys_lens), ptr(leader), i32(B), i32(T), i32(V), i32(Lmax), i32(ctx.blank),
                      ptr(words), ptr(val), st)
            _lib.call("b200_split_gemm_exp_sub", ptr(hs_hi), ptr(hs_lo), _i64(rows), i32(E), i32(E), ptr(w_hi),
                      ptr(w_lo), i32(V), i32(E), ptr(bias_pad), ptr(coef), ptr(words), ptr(val), i32(T),
                      i32(Lmax + 1), ptr(G_hi), ptr(G_lo), _i64(ldg), i32(0), st)
            dW = torch.zeros(V, E, dtype=torch.float32, device=dev)
            db = torch.zeros(V, dtype=torch.float32, device=dev)
            dhs = torch.empty(B, T, E, dtype=torch.float32, device=dev)
            # d ctc_lo.weight and, from the same operand tiles, d ctc_lo.bias (column sums of G)
            _lib.call("b200_joint_dw_split_db", ptr(G_hi), ptr(G_lo), i32(ldg), ptr(hs_hi), ptr(hs_lo), i32(E),
                      _i64(rows), i32(V), i32(E), ptr(dW), None, ptr(db), i32(0), st)
            _lib.call("b200_split_gemm_nn", ptr(G_hi), ptr(G_lo), i32(ldg), ptr(w_hi), ptr(w_lo), _i64(rows),
                      i32(V), i32(E), ptr(dhs), i32(0), st)
        dt_h, dt_w, dt_b = ctx.in_dtypes
        return dhs.to(dt_h), dW.to(dt_w), db.to(dt_b), None, None, None, None


def _fusable(E: int) -> bool:
    return E in (256, 512)  # tile shapes of csrc/joint_dw.cu / joint_dh.cu


def _check_lengths(hl, yl, T, Lmax, key=None):
    # lengths beyond the tensors would run the trellis kernels out of bounds; torch.nn.CTCLoss raises too.
    # `key`: the caller's own length tensors (hl / yl may be fresh int32 copies of them)
    mx = _lib.cached_host_read(key or (hl, yl),
                               lambda: torch.stack([hl.max(), yl.max(), -hl.min(), -yl.min()]).tolist())
    if mx[0] > T or mx[2] > 0:
        raise RuntimeError(f"Expected input_lengths to have value at most {T}, but got value "
                           f"{mx[0] if mx[0] > T else -mx[2]} (while checking arguments for ctc_loss)")
    if mx[1] > Lmax or mx[3] > 0:
        raise RuntimeError(f"Expected target_lengths to have value at most {Lmax}, but got value "
                           f"{mx[1] if mx[1] > Lmax else -mx[3]} (while checking arguments for ctc_loss)")


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
    ys = _lib.cached_convert(ys_pad, dev, torch.int64)
    hl = _lib.cached_convert(hlens, dev, torch.int32)
    yl = _lib.cached_convert(ys_lens, dev, torch.int32)
    _check_lengths(hl, yl, logits.size(1), ys.size(1), key=(hlens, ys_lens))
    nll = _CtcLossFn.apply(logits, ys, hl, yl, int(blank))
    if reduction == "sum":
        return nll.sum()
    if reduction == "mean":  # torch semantics: per-utterance nll / target length, then batch mean
        return (nll / yl.clamp_min(1).to(nll.dtype)).mean()
    return nll


def _ctc_forward(self, hs_pad: torch.Tensor, hlens: torch.Tensor, ys_pad: torch.Tensor,
                 ys_lens: torch.Tensor) -> torch.Tensor:
    """CTC.forward (ctc.py:46-64): ``sum_b nll_b / B`` (or ``nll / B`` when reduce=False)."""
    hs = F.dropout(hs_pad, p=self.dropout_rate)  # training=True by default, exactly as ctc.py:56
    reduce = getattr(self, "reduce", None)
    if reduce is None:  # a reference CTC instance: recover the flag from its criterion
        reduce = getattr(getattr(self, "ctc_loss", None), "reduction", "sum") == "sum"
    B, T, E = hs.shape
    if _fusable(E) and hs.is_cuda:
        if ys_pad.dim() != 2 or ys_pad.size(0) != B or hlens.numel() != B or ys_lens.numel() != B:
            raise RuntimeError("batch dimension mismatch between hs_pad, ys_pad, hlens and ys_lens")
        dev = hs.device
        ys = _lib.cached_convert(ys_pad, dev, torch.int64)
        hl = _lib.cached_convert(hlens, dev, torch.int32)
        yl = _lib.cached_convert(ys_lens, dev, torch.int32)
        _check_lengths(hl, yl, T, ys.size(1), key=(hlens, ys_lens))
        nll = _FusedCtcFn.apply(hs, self.ctc_lo.weight, self.ctc_lo.bias, ys, hl, yl, 0)
        return (nll.sum() if reduce else nll) / B
    ys_hat = self.ctc_lo(hs)  # [B,T,V], library GEMM
    loss = ctc_loss(ys_hat, ys_pad, hlens, ys_lens, blank=0, reduction="sum" if reduce else "none")
    return loss / ys_hat.size(0)


def _ctc_log_softmax(self, hs_pad: torch.Tensor) -> torch.Tensor:
    """CTC.log_softmax (ctc.py:66-74): fp32 log-probs [B,T,V], projection and softmax in one pass pair."""
    B, T, E = hs_pad.shape
    V = self.ctc_lo.out_features
    if not (hs_pad.is_cuda and E % 8 == 0 and V % 4 == 0) or torch.is_grad_enabled() and (
            hs_pad.requires_grad or self.ctc_lo.weight.requires_grad):
        return F.log_softmax(self.ctc_lo(hs_pad), dim=2)  # autograd wanted / shapes TMA cannot address
    dev = hs_pad.device
    with torch.cuda.device(dev):
        st = _lib.stream_handle(dev)
        hs_hi, hs_lo = split_f32(hs_pad.reshape(B * T, E))
        w_hi, w_lo = split_f32(self.ctc_lo.weight)
        bias_pad = _pad_bias(self.ctc_lo.bias, V)
        lse = torch.empty(B * T, dtype=torch.float32, device=dev)
        out = torch.empty(B, T, V, dtype=torch.float32, device=dev)
        args = (ptr(hs_hi), ptr(hs_lo), _i64(B * T), i32(E), i32(E), ptr(w_hi), ptr(w_lo), i32(V), i32(E),
                ptr(bias_pad))
        _lib.call("b200_split_gemm_lse", *args, ptr(lse), i32(0), st)
        _lib.call("b200_split_gemm_lsm", *args, ptr(lse), ptr(out), _i64(V), i32(0), st)
    return out


class CTC(torch.nn.Module):
    """CTC module (same constructor and parameter names as the reference's)."""

    def __init__(self, odim: int, encoder_output_size: int, dropout_rate: float = 0.0,
                 reduce: bool = True):
        super().__init__()
        self.dropout_rate = dropout_rate
        self.reduce = reduce
        self.ctc_lo = torch.nn.Linear(encoder_output_size, odim)

    forward = _ctc_forward

    log_softmax = _ctc_log_softmax

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
    module.log_softmax = types.MethodType(_ctc_log_softmax, module)
    return module
