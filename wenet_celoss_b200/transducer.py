"""Model-level drop-in for the reference ``Transducer`` (wenet/transducer/transducer.py).

``patch_transducer(model)`` rebinds, on an existing reference ``Transducer`` instance (or
any module with ``encoder``, ``predictor``, ``joint``, ``ctc``, ``blank``, ...), the methods
on the hot path to their sm_100a versions and leaves everything else untouched:

  forward               transducer.py:79-270   same signature, same loss dict keys
                        ``loss, loss_att, loss_ctc, loss_rnnt, hw_loss``; joint + rnnt_loss
                        (:132-147) -> fused_joint_rnnt_loss, self.ctc(...) (:159) -> CTC kernels
  greedy_search         transducer.py:515-598  batched (the reference asserts batch == 1)
  beam_search           transducer.py:332-377  returns (hyp, score) of the best hypothesis
  _cal_transducer_score transducer.py:277-302  -cost per hypothesis, forward only

Out of scope and kept as in the reference object: encoder, attention decoder
(``_calc_att_loss``), context biasing (called exactly where the reference calls it, so an
identity or a real ``ContextBias`` both work).  The hot-word CE loss (``hw_weight != 0``,
transducer.py:170-262) runs through ``hw_loss.hw_ce_loss`` (csrc/ce_loss.cu) on the outputs of the
model's own ``ContextBias.forward_hw_pred*`` heads.
"""
from __future__ import annotations

import types
from typing import Dict, Optional

import torch

from .ctc import patch_ctc
from .hw_loss import hw_loss_from_streams
from .joint import fused_joint_rnnt_loss
from .search import greedy_search_batch, prefix_beam_search_batch

IGNORE_ID = -1


def add_blank(ys_pad: torch.Tensor, blank: int, ignore_id: int) -> torch.Tensor:
    """[B,L] -> [B,L+1]: blank prepended, padding replaced by blank (wenet/utils/common.py:56-87)."""
    head = torch.full((ys_pad.size(0), 1), blank, dtype=ys_pad.dtype, device=ys_pad.device)
    out = torch.cat([head, ys_pad], dim=1)
    return torch.where(out == ignore_id, torch.full_like(out, blank), out)


def rnnt_targets(text: torch.Tensor, text_lengths: torch.Tensor, encoder_out_lens: torch.Tensor,
                 ignore_id: int = IGNORE_ID):
    """The int32 / zero-padded views torchaudio wants (transducer.py:137-141)."""
    t = text.to(torch.int64)
    t = torch.where(t == ignore_id, torch.zeros_like(t), t).to(torch.int32).contiguous()
    return t, encoder_out_lens.to(torch.int32).contiguous(), text_lengths.to(torch.int32).contiguous()


def _bias(model, name, *args):
    cb = getattr(model, "context_bias", None)
    if cb is None:
        return None
    return getattr(cb, name)(*args)


def transducer_forward(self, speech: torch.Tensor, speech_lengths: torch.Tensor, text: torch.Tensor,
                       text_lengths: torch.Tensor, context_list=None, context_lengths=None,
                       hw_label=None) -> Dict[str, Optional[torch.Tensor]]:
    assert text_lengths.dim() == 1, text_lengths.shape
    assert (speech.shape[0] == speech_lengths.shape[0] == text.shape[0] == text_lengths.shape[0]), (
        speech.shape, speech_lengths.shape, text.shape, text_lengths.shape)
    ignore_id = getattr(self, "ignore_id", IGNORE_ID)
    bias_hidden = _bias(self, "forward_bias_hidden", context_list, context_lengths)
    encoder_out, encoder_mask = self.encoder(speech, speech_lengths)
    encoder_out_lens = encoder_mask.squeeze(1).sum(1)
    encoder_out_bias = predictor_out_bias = None
    if getattr(self, "context_bias", None) is not None:
        encoder_out, encoder_out_bias = self.context_bias.forward_encoder_bias(bias_hidden, encoder_out)
    ys_in_pad = add_blank(text, self.blank, ignore_id)
    predictor_out = self.predictor(ys_in_pad)
    if getattr(self, "context_bias", None) is not None:
        predictor_out, predictor_out_bias = self.context_bias.forward_predictor_bias(bias_hidden,
                                                                                  predictor_out)
    rnnt_text, enc_lens32, text_lens32 = rnnt_targets(text, text_lengths, encoder_out_lens, ignore_id)
    loss = fused_joint_rnnt_loss(self.joint, encoder_out, predictor_out, rnnt_text, enc_lens32,
                                 text_lens32, blank=self.blank, reduction="mean")
    loss_rnnt = loss
    loss = self.transducer_weight * loss
    loss_att: Optional[torch.Tensor] = None
    if getattr(self, "attention_decoder_weight", 0.0) != 0.0 and getattr(self, "decoder", None) is not None:
        loss_att, _ = self._calc_att_loss(encoder_out, encoder_mask, text, text_lengths)
    loss_ctc: Optional[torch.Tensor] = None
    if self.ctc_weight != 0.0 and self.ctc is not None:
        loss_ctc = self.ctc(encoder_out, enc_lens32, text, text_lengths)
    if loss_ctc is not None:
        loss = loss + self.ctc_weight * loss_ctc.sum()
    if loss_att is not None:
        loss = loss + self.attention_decoder_weight * loss_att.sum()
    hw_loss: Optional[torch.Tensor] = None
    if getattr(self, "hw_weight", 0.0) != 0.0:  # transducer.py:170-262
        if getattr(self, "context_bias", None) is None or hw_label is None:
            raise RuntimeError("hw_weight != 0 needs model.context_bias and hw_label")
        hw_loss = hw_loss_from_streams(self, bias_hidden, predictor_out, encoder_out_bias,
                                       predictor_out_bias, hw_label, add_blank)
        loss = loss + self.hw_weight * hw_loss
    return {"loss": loss, "loss_att": loss_att, "loss_ctc": loss_ctc, "loss_rnnt": loss_rnnt,
            "hw_loss": hw_loss}


@torch.no_grad()
def transducer_score(self, encoder_out: torch.Tensor, encoder_mask: torch.Tensor,
                     hyps_lens: torch.Tensor, hyps_pad: torch.Tensor) -> torch.Tensor:
    ignore_id = getattr(self, "ignore_id", IGNORE_ID)
    hyps_pad_blank = add_blank(hyps_pad, self.blank, ignore_id)
    xs_in_lens = encoder_mask.squeeze(1).sum(1).int()
    predictor_out = self.predictor(hyps_pad_blank)
    rnnt_text, xs32, hl32 = rnnt_targets(hyps_pad, hyps_lens, xs_in_lens, ignore_id)
    cost = fused_joint_rnnt_loss(self.joint, encoder_out, predictor_out, rnnt_text, xs32, hl32,
                                 blank=self.blank, reduction="none")
    return cost * -1


@torch.no_grad()
def transducer_greedy_search(self, speech: torch.Tensor, speech_lengths: torch.Tensor,
                             decoding_chunk_size: int = -1, num_decoding_left_chunks: int = -1,
                             simulate_streaming: bool = False, n_steps: int = 64,
                             context_list: Optional[torch.Tensor] = None,
                             context_lengths: Optional[torch.Tensor] = None,
                             context_filter_state: str = "on",
                             context_decoder_labels_padded: Optional[torch.Tensor] = None):
    """``Transducer.greedy_search`` (transducer.py:515-598): same positional order and defaults -
    ``context_filter_state='on'`` (the fork's gated decode with go-back), context tensors defaulting to
    ``IntTensor([0])`` - and the same return shape ``(hyps, dist)``: ``dist`` = edit distance between
    ``context_decoder_labels_padded`` (default ``[0]``) and the gate decisions (greedy_search.py:174),
    0 for loss modes that return no distance (transducer.py:589-598).  Batched: ``hyps`` holds one
    token list per utterance (the reference asserts batch == 1) and ``dist`` is a list when B > 1.

    A model WITHOUT a ``ContextBias`` cannot run the gate; it decodes with the upstream algorithm
    ("greedy_search copy.py":6-63) and returns ``dist = 0``."""
    assert speech.shape[0] == speech_lengths.shape[0]
    assert decoding_chunk_size != 0
    if context_filter_state not in ("on", "off"):
        raise ValueError("context_filter_state must be 'on' or 'off'")
    _ = simulate_streaming
    encoder_out, encoder_mask = self.encoder(speech, speech_lengths, decoding_chunk_size,
                                             num_decoding_left_chunks)
    encoder_out_lens = encoder_mask.squeeze(1).sum(1)
    from .search import _has_hotword_bias, edit_distance
    if not _has_hotword_bias(self):
        return greedy_search_batch(self, encoder_out, encoder_out_lens, n_steps=n_steps), 0
    if context_list is None:
        context_list = torch.IntTensor([0])
    if context_lengths is None:
        context_lengths = torch.IntTensor([0])
    if context_decoder_labels_padded is None:
        context_decoder_labels_padded = torch.IntTensor([0])
    hyps, decisions = greedy_search_batch(self, encoder_out, encoder_out_lens, n_steps=n_steps,
                                          context_list=context_list, context_lengths=context_lengths,
                                          context_filter_state=context_filter_state,
                                          return_decisions=True)
    if getattr(self, "loss_mode", "pred") not in ("pred", "both"):
        return hyps, 0  # transducer.py:589-598: this branch returns the hypotheses only, dist stays 0
    lab_t = torch.as_tensor(context_decoder_labels_padded)
    per_utt = lab_t.dim() > 1 and lab_t.size(0) == len(hyps)
    lab = lab_t.reshape(len(hyps), -1).tolist() if per_utt else [lab_t.reshape(-1).tolist()] * len(hyps)
    dist = [edit_distance(lab[b], decisions[b]) for b in range(len(hyps))]
    return hyps, (dist[0] if len(dist) == 1 else dist)


@torch.no_grad()
def transducer_beam_search(self, speech: torch.Tensor, speech_lengths: torch.Tensor,
                           decoding_chunk_size: int = -1, beam_size: int = 5,
                           num_decoding_left_chunks: int = -1, simulate_streaming: bool = False,
                           ctc_weight: float = 0.3, transducer_weight: float = 0.7):
    """(best hyp without the seed blank, score), as transducer.py:377; a batch returns a list."""
    encoder_out, encoder_mask = self.encoder(speech, speech_lengths, decoding_chunk_size,
                                             num_decoding_left_chunks)
    lens = encoder_mask.squeeze(1).sum(1)
    beams = prefix_beam_search_batch(self, encoder_out, lens, beam_size, ctc_weight, transducer_weight)
    out = [(b[0].hyp[1:], b[0].score) for b in beams]
    return out[0] if len(out) == 1 else out


def _add_sos_eos_in(hyps, sos: int, eos: int, device) -> torch.Tensor:
    """``ys_in`` of add_sos_eos (wenet/utils/common.py:122-163): <sos> + tokens, padded with <eos>."""
    L = max(len(h) for h in hyps) + 1
    out = torch.full((len(hyps), L), eos, dtype=torch.long)
    for i, h in enumerate(hyps):
        out[i, 0] = sos
        if h:
            out[i, 1:1 + len(h)] = torch.tensor(h, dtype=torch.long)
    return out.to(device)


@torch.no_grad()
def transducer_attention_rescoring(self, speech: torch.Tensor, speech_lengths: torch.Tensor,
                                   beam_size: int, decoding_chunk_size: int = -1,
                                   num_decoding_left_chunks: int = -1, simulate_streaming: bool = False,
                                   reverse_weight: float = 0.0, ctc_weight: float = 0.0,
                                   attn_weight: float = 0.0, transducer_weight: float = 0.0,
                                   search_ctc_weight: float = 1.0, search_transducer_weight: float = 0.0,
                                   beam_search_type: str = "transducer"):
    """``Transducer.transducer_attention_rescoring`` (transducer.py:378-513), same arguments and return
    ``(best hypothesis, best score)``.

    n-best search on the device (RNN-T or CTC prefix beam search), then one rescoring step:
    transducer score of every hypothesis through the fused joint + RNN-T loss kernels with ONE encoder
    projection shared by the beam (the reference ``repeat``s encoder_out beam_size times, :467, and
    materialises [beam, T, U+1, V] logits), the attention decoder (a library module of the model, called
    as the reference calls it), and the score combine + argmax in ``b200_attn_rescore`` instead of a
    ``.cpu().numpy()`` round trip and Python loops (:487-511)."""
    import ctypes

    from . import _lib
    from ._lib import i32, ptr
    from .joint import fused_joint_rnnt_loss as _fused

    assert speech.shape[0] == speech_lengths.shape[0]
    assert decoding_chunk_size != 0
    if reverse_weight > 0.0:
        assert hasattr(self.decoder, "right_decoder")
    assert speech.shape[0] == 1  # as the reference: attention rescoring is per utterance
    dev = speech.device
    _lib.require_cuda(speech, "speech")
    if beam_search_type == "transducer":
        encoder_out, encoder_mask = self.encoder(speech, speech_lengths, decoding_chunk_size,
                                                 num_decoding_left_chunks)
        lens = encoder_mask.squeeze(1).sum(1)
        beam = prefix_beam_search_batch(self, encoder_out, lens, beam_size, search_ctc_weight,
                                        search_transducer_weight)[0]
        beam_score = [s.score for s in beam]
        hyps = [list(s.hyp[1:]) for s in beam]
    elif beam_search_type == "ctc":
        nbest, encoder_out = self._ctc_prefix_beam_search(
            speech, speech_lengths, beam_size=beam_size, decoding_chunk_size=decoding_chunk_size,
            num_decoding_left_chunks=num_decoding_left_chunks, simulate_streaming=simulate_streaming)
        beam_score = [h[1] for h in nbest]
        hyps = [list(h[0]) for h in nbest]
    else:
        raise ValueError("beam_search_type must be 'transducer' or 'ctc'")
    assert len(hyps) == beam_size
    N, T = beam_size, encoder_out.size(1)
    ignore_id = getattr(self, "ignore_id", IGNORE_ID)
    Lmax = max(len(h) for h in hyps)
    hyps_pad = torch.full((N, max(Lmax, 1)), ignore_id, dtype=torch.long)
    for i, h in enumerate(hyps):
        if h:
            hyps_pad[i, :len(h)] = torch.tensor(h, dtype=torch.long)
    hyps_pad = hyps_pad[:, :Lmax].contiguous().to(dev)
    hyps_lens = torch.tensor([len(h) for h in hyps], dtype=torch.long, device=dev)

    # 2.1 transducer score: predictor over the n-best, ONE encoder projection for the whole beam
    td_score = transducer_score_shared(self, encoder_out, hyps_lens, hyps_pad)

    # 2.2 attention decoder (model module; encoder_out broadcast as a stride-0 view, not a copy)
    enc_rep = encoder_out.expand(N, -1, -1)
    enc_mask = torch.ones(N, 1, T, dtype=torch.bool, device=dev)
    ys_in = _add_sos_eos_in(hyps, self.sos, self.eos, dev)
    r_hyps = [list(reversed(h)) for h in hyps]
    r_ys_in = _add_sos_eos_in(r_hyps, self.sos, self.eos, dev)
    decoder_out, r_decoder_out, _ = self.decoder(enc_rep, enc_mask, ys_in, hyps_lens + 1, r_ys_in,
                                                 reverse_weight)
    dec = decoder_out.float().contiguous()
    use_r = reverse_weight > 0.0 and torch.is_tensor(r_decoder_out) and r_decoder_out.dim() == 3
    rdec = r_decoder_out.float().contiguous() if use_r else None
    L1, V = dec.size(1), dec.size(2)
    with torch.cuda.device(dev):
        scores = torch.empty(N, dtype=torch.float64, device=dev)
        best = torch.empty(1, dtype=torch.int32, device=dev)
        bs = torch.tensor(beam_score, dtype=torch.float64, device=dev)
        _lib.call("b200_attn_rescore", ptr(dec), ptr(rdec), ptr(hyps_pad), ptr(hyps_lens.int().contiguous()),
                  i32(N), i32(L1), i32(Lmax), i32(V), i32(self.eos), ctypes.c_double(reverse_weight),
                  ctypes.c_double(attn_weight), ptr(bs), ctypes.c_double(ctc_weight),
                  ptr(td_score.float().contiguous()), ctypes.c_double(transducer_weight), ptr(scores),
                  ptr(None), ptr(best), _lib.stream_handle(dev))
        bi = int(best.item())
    return hyps[bi], float(scores[bi].item())


@torch.no_grad()
def transducer_score_shared(self, encoder_out: torch.Tensor, hyps_lens: torch.Tensor,
                            hyps_pad: torch.Tensor) -> torch.Tensor:
    """``_cal_transducer_score`` (transducer.py:277-302) for N hypotheses of ONE utterance:
    ``encoder_out`` is ``[1, T, E]`` - the encoder projection runs once and is shared by the beam."""
    ignore_id = getattr(self, "ignore_id", IGNORE_ID)
    N, T = hyps_pad.size(0), encoder_out.size(1)
    hyps_pad_blank = add_blank(hyps_pad, self.blank, ignore_id)
    predictor_out = self.predictor(hyps_pad_blank)
    xs_in_lens = torch.full((N,), T, dtype=torch.int32, device=encoder_out.device)
    rnnt_text, xs32, hl32 = rnnt_targets(hyps_pad, hyps_lens, xs_in_lens, ignore_id)
    cost = fused_joint_rnnt_loss(_SharedEncJoint(self.joint, N), encoder_out, predictor_out, rnnt_text, xs32,
                                 hl32, blank=self.blank, reduction="none")
    return cost * -1


class _SharedEncJoint:
    """The model's joint with ``enc_ffn`` evaluated on the single utterance and broadcast over the
    n-best: [1,T,E] -> one GEMM -> [N,T,J] view (what fused_joint_rnnt_loss consumes)."""

    def __init__(self, joint, n):
        self._joint, self._n = joint, n
        self.pred_ffn, self.ffn_out = joint.pred_ffn, joint.ffn_out
        for name in ("postjoin_linear", "post_ffn", "activatoin"):
            if hasattr(joint, name):
                setattr(self, name, getattr(joint, name))

    def enc_ffn(self, x):
        return self._joint.enc_ffn(x).expand(self._n, -1, -1)


def patch_transducer(model: torch.nn.Module) -> torch.nn.Module:
    """Rebind the hot-path methods of a reference ``Transducer`` instance in place."""
    for attr in ("encoder", "predictor", "joint", "blank"):
        if not hasattr(model, attr):
            raise TypeError(f"patch_transducer expects a Transducer-like module (missing .{attr})")
    if getattr(model, "ctc", None) is not None:
        patch_ctc(model.ctc)
        if hasattr(model, "_forward_encoder"):  # ASRModel's CTC prefix beam search (asr_model.py:326-443)
            from . import ctc_search
            model._ctc_prefix_beam_search = types.MethodType(ctc_search._ctc_prefix_beam_search, model)
            model.ctc_prefix_beam_search = types.MethodType(ctc_search.ctc_prefix_beam_search, model)
    model.forward = types.MethodType(transducer_forward, model)
    model._cal_transducer_score = types.MethodType(transducer_score, model)
    model.greedy_search = types.MethodType(transducer_greedy_search, model)
    model.beam_search = types.MethodType(transducer_beam_search, model)
    if getattr(model, "decoder", None) is not None and hasattr(model, "sos"):
        model.transducer_attention_rescoring = types.MethodType(transducer_attention_rescoring, model)
    return model
