"""CTC prefix beam search with the reference's interface, on the device for a whole batch.

Mirrors ``ASRModel._ctc_prefix_beam_search`` / ``ctc_prefix_beam_search``
(wenet/transformer/asr_model.py:326-443), which ``Transducer`` inherits and
``transducer_attention_rescoring(beam_search_type='ctc')`` calls
(wenet/transducer/transducer.py:447-456).  The reference handles one utterance with a Python
dict per frame; here one CTA per utterance runs all frames (csrc/ctc_beam.cu).
"""
from __future__ import annotations

import ctypes
from typing import List, Tuple

import torch

from . import _lib
from ._lib import i32, ptr


@torch.no_grad()
def ctc_prefix_beam_search_batch(ctc_probs: torch.Tensor, lens: torch.Tensor, beam_size: int,
                                 blank: int = 0) -> List[List[Tuple[Tuple[int, ...], float]]]:
    """``ctc_probs [B,T,V]``: log-probs (``CTC.log_softmax``, ctc.py:66-74); ``lens [B]`` frames
    per utterance.  Returns per utterance the reference's ``hyps`` list
    ``[(prefix tuple, log_add(pb, pnb)), ...]`` best first (asr_model.py:408)."""
    _lib.require_cuda(ctc_probs, "ctc_probs")
    if ctc_probs.dim() != 3:
        raise RuntimeError("ctc_probs must be 3-D (batch, time, class)")
    dev = ctc_probs.device
    B, T, V = ctc_probs.shape
    with torch.cuda.device(dev):
        lp = ctc_probs.float().contiguous()
        ln = lens.reshape(-1).to(device=dev, dtype=torch.int32).clamp(max=T).contiguous()
        if ln.numel() != B:
            raise RuntimeError("lens must have one entry per utterance")
        out_tok = torch.empty(B, beam_size, T, dtype=torch.int32, device=dev)
        out_len = torch.empty(B, beam_size, dtype=torch.int32, device=dev)
        out_score = torch.empty(B, beam_size, dtype=torch.float64, device=dev)
        out_n = torch.empty(B, dtype=torch.int32, device=dev)
        L_ = _lib.lib()
        L_.b200_ctc_prefix_beam_workspace_bytes.restype = ctypes.c_size_t
        ws_bytes = L_.b200_ctc_prefix_beam_workspace_bytes(B, T, beam_size)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        _lib.call("b200_ctc_prefix_beam", ptr(lp), ptr(ln), i32(B), i32(T), i32(V), i32(beam_size),
                  i32(blank), ptr(out_tok), ptr(out_len), ptr(out_score), ptr(out_n), ptr(ws),
                  ctypes.c_size_t(ws_bytes), _lib.stream_handle(dev))
        n = out_n.cpu().tolist()
        lens_h = out_len.cpu().tolist()
        scores = out_score.cpu().tolist()
        width = max(max(max(r) for r in lens_h), 1) if B else 1
        toks = out_tok[:, :, :width].cpu().numpy()
    # one flat .tolist() per hypothesis (a nested .tolist() of the padded [B, beam, width] block was 2/3 of
    # the call at long prefixes)
    return [[(tuple(toks[b, k, :lens_h[b][k]].tolist()), scores[b][k]) for k in range(n[b])] for b in range(B)]


def _ctc_prefix_beam_search(self, speech: torch.Tensor, speech_lengths: torch.Tensor, beam_size: int,
                            decoding_chunk_size: int = -1, num_decoding_left_chunks: int = -1,
                            simulate_streaming: bool = False):
    """Drop-in for ``ASRModel._ctc_prefix_beam_search`` (asr_model.py:326-409): same arguments,
    returns ``(hyps, encoder_out)``; batches of more than one utterance return a list of hyps
    lists (the reference asserts batch_size == 1, :358)."""
    assert speech.shape[0] == speech_lengths.shape[0]
    assert decoding_chunk_size != 0
    encoder_out, encoder_mask = self._forward_encoder(speech, speech_lengths, decoding_chunk_size,
                                                      num_decoding_left_chunks, simulate_streaming)
    B, maxlen = encoder_out.size(0), encoder_out.size(1)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        ctc_probs = self.ctc.log_softmax(encoder_out)  # library GEMM + log_softmax (:368)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    # the reference walks all maxlen frames of its single utterance (:374)
    lens = torch.full((B,), maxlen, dtype=torch.int32)
    hyps = ctc_prefix_beam_search_batch(ctc_probs, lens, beam_size, blank=0)
    return (hyps[0] if B == 1 else hyps), encoder_out


def ctc_prefix_beam_search(self, speech: torch.Tensor, speech_lengths: torch.Tensor, beam_size: int,
                           decoding_chunk_size: int = -1, num_decoding_left_chunks: int = -1,
                           simulate_streaming: bool = False):
    """Drop-in for ``ASRModel.ctc_prefix_beam_search`` (asr_model.py:411-443): ``hyps[0]``, the
    best ``(prefix, score)`` pair, exactly what the reference returns (:443)."""
    hyps, _ = _ctc_prefix_beam_search(self, speech, speech_lengths, beam_size, decoding_chunk_size,
                                      num_decoding_left_chunks, simulate_streaming)
    return hyps[0]


@torch.no_grad()
def forced_align_batch(ctc_probs: torch.Tensor, ys_pad: torch.Tensor, hlens: torch.Tensor,
                       ys_lens: torch.Tensor, blank_id: int = 0) -> List[List[int]]:
    """CTC forced alignment of a batch: ``ctc_probs [B,T,V]`` log-probs, ``ys_pad [B,Lmax]``
    labels (padding ignored), ``hlens`` / ``ys_lens [B]``.  Returns per utterance the aligned token
    of every frame (ctc_util.py:27-83 per utterance)."""
    _lib.require_cuda(ctc_probs, "ctc_probs")
    dev = ctc_probs.device
    B, T, V = ctc_probs.shape
    Lmax = ys_pad.size(1)
    with torch.cuda.device(dev):
        lp = ctc_probs.float().contiguous()
        ys = ys_pad.to(device=dev, dtype=torch.int64).contiguous()
        hl = hlens.reshape(-1).to(device=dev, dtype=torch.int32).clamp(max=T).contiguous()
        yl = ys_lens.reshape(-1).to(device=dev, dtype=torch.int32).contiguous()
        if int(yl.min()) < 1:
            raise RuntimeError("forced_align needs at least one label per utterance")
        out = torch.empty(B, T, dtype=torch.int32, device=dev)
        L_ = _lib.lib()
        L_.b200_ctc_forced_align_workspace_bytes.restype = ctypes.c_size_t
        ws_bytes = L_.b200_ctc_forced_align_workspace_bytes(B, T, Lmax)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        _lib.call("b200_ctc_forced_align", ptr(lp), ptr(ys), ptr(hl), ptr(yl), i32(B), i32(T), i32(V),
                  i32(Lmax), i32(blank_id), ptr(out), ptr(ws), ctypes.c_size_t(ws_bytes),
                  _lib.stream_handle(dev))
        rows = out.cpu().tolist()
        n = hl.cpu().tolist()
    return [rows[b][:n[b]] for b in range(B)]


def forced_align(ctc_probs: torch.Tensor, y: torch.Tensor, blank_id: int = 0) -> list:
    """Drop-in for ``forced_align(ctc_probs [T,D], y [L], blank_id)`` (ctc_util.py:27-83)."""
    T = ctc_probs.size(0)
    return forced_align_batch(ctc_probs.unsqueeze(0), y.reshape(1, -1), torch.tensor([T]),
                              torch.tensor([y.numel()]), blank_id)[0]
