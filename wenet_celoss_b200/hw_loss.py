"""Hot-word cross-entropy loss of the fork ("celoss") and the hook that computes ``hw_loss`` in
``Transducer.forward`` (wenet/transducer/transducer.py:66,170-262).

``hw_ce_loss(hw_output [B,C,U], labels [B,U])`` stands in for ``self.hw_criterion``
(``nn.CrossEntropyLoss()``, mean reduction, ignore_index -100) on the device kernels of
csrc/ce_loss.cu.  ``hw_loss_from_streams`` mirrors the three ``loss_mode`` branches; the hot-word
classifier heads themselves (``ContextBias.forward_hw_pred*``, context_bias.py:382-399) are model
modules and are called exactly where the reference calls them.
"""
from __future__ import annotations

import ctypes

import torch
from torch.amp import custom_bwd, custom_fwd

from . import _lib
from ._lib import i32, ptr


class _CeLossFn(torch.autograd.Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, logits, labels, ignore_index):
        B, C, U = logits.shape
        dev = logits.device
        with torch.cuda.device(dev):
            x = logits.detach().float().contiguous()
            y = labels.to(device=dev, dtype=torch.int64).contiguous()
            # nn.CrossEntropyLoss asserts on the device for labels outside [0, C); so does this path
            torch._assert_async(((y == ignore_index) | ((y >= 0) & (y < C))).all(),
                                "hw_ce_loss: label out of range [0, C) and not ignore_index")
            out = torch.empty(2, dtype=torch.float32, device=dev)
            _lib.call("b200_ce_loss_fwd", ptr(x), ptr(y), i32(B), i32(C), i32(U),
                      ctypes.c_longlong(ignore_index), ptr(out[0:1]), ptr(out[1:2]),
                      _lib.stream_handle(dev))
        ctx.save_for_backward(x, y, out)
        ctx.ignore_index = ignore_index
        ctx.in_dtype = logits.dtype
        return out[0].clone()

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, dy):
        x, y, out = ctx.saved_tensors
        B, C, U = x.shape
        dev = x.device
        with torch.cuda.device(dev):
            d = dy.detach().float().reshape(1).contiguous()
            grad = torch.empty_like(x)
            _lib.call("b200_ce_loss_bwd", ptr(x), ptr(y), i32(B), i32(C), i32(U),
                      ctypes.c_longlong(ctx.ignore_index), ptr(d), ptr(out[1:2]), ptr(grad),
                      _lib.stream_handle(dev))
        return grad.to(ctx.in_dtype), None, None


def hw_ce_loss(hw_output: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """``nn.CrossEntropyLoss()(hw_output, labels)`` for ``hw_output [B, C, U]`` (class dim 1) and
    ``labels [B, U]``: scalar mean over the positions whose label is not ``ignore_index``."""
    _lib.require_cuda(hw_output, "hw_output")
    if hw_output.dim() != 3 or labels.dim() != 2:
        raise RuntimeError("hw_ce_loss expects hw_output [B, C, U] and labels [B, U]")
    if hw_output.size(0) != labels.size(0) or hw_output.size(2) != labels.size(1):
        raise RuntimeError(f"Expected target size {[hw_output.size(0), hw_output.size(2)]}, "
                           f"got {list(labels.shape)}")
    return _CeLossFn.apply(hw_output, labels, int(ignore_index))


def end_blank(ys_pad: torch.Tensor, blank: int, ignore_id: int) -> torch.Tensor:
    """[B,L] -> [B,L+1]: blank APPENDED, padding replaced by blank (wenet/utils/common.py:89-120)."""
    tail = torch.full((ys_pad.size(0), 1), blank, dtype=torch.long, device=ys_pad.device)
    out = torch.cat([ys_pad, tail], dim=1)
    return torch.where(out == ignore_id, torch.full_like(out, blank), out)


def hw_loss_from_streams(model, bias_hidden, predictor_out, encoder_out_bias, predictor_out_bias,
                         hw_label, add_blank_fn):
    """``hw_loss`` of Transducer.forward (transducer.py:170-200) for ``loss_mode`` 'pred', 'both'
    or anything else (the separate-heads branch).  ``predictor_out`` is the BIASED predictor stream
    (the reference clones it after forward_predictor_bias, :128-129)."""
    cb = model.context_bias
    ignore_id = getattr(model, "ignore_id", -1)
    mode = getattr(model, "loss_mode", "both")
    if mode == "pred":
        hw_output = cb.forward_hw_pred(bias_hidden, predictor_out).permute(0, 2, 1)
        labels = end_blank(hw_label, model.blank, ignore_id)[..., :-1]
        return hw_ce_loss(hw_output[..., :-1], labels)
    if mode == "both":
        hw_output = cb.forward_hw_pred_both(encoder_out_bias, predictor_out_bias).permute(0, 2, 1)
        labels = end_blank(hw_label, model.blank, ignore_id)[..., :-1]
        return hw_ce_loss(hw_output[..., :-1], labels)
    _, hw_output_dec = cb.forward_hw_pred_both_sep(encoder_out_bias, predictor_out_bias)
    labels = add_blank_fn(hw_label, model.blank, ignore_id)
    return hw_ce_loss(hw_output_dec.permute(0, 2, 1), labels)
