"""Batched RNN-T greedy search (and, below, prefix beam search) on the device.

Mirrors the reference entry points
  wenet/transducer/search/greedy_search.py:34   basic_greedy_search(model, encoder_out, encoder_out_lens, ..., n_steps=64)
  wenet/transducer/search/greedy_search copy.py:6 (the same algorithm without hot-word gating)
and adds the batched form the reference lacks (it asserts batch == 1,
wenet/transducer/transducer.py:545): ``greedy_search_batch(model, encoder_out [B,T,E],
encoder_out_lens [B])`` returns B token lists.  ``model`` is anything exposing what the
reference functions read: ``model.predictor`` (``embed``, ``rnn`` = nn.LSTM, ``projection``),
``model.joint`` (``enc_ffn``, ``pred_ffn``, ``ffn_out``) and ``model.blank``.

With a real ``ContextBias`` on the model the fork's hot-word streams are applied as
``context_filter_state='off'`` does (greedy_search.py:62-70,140-144: encoder stream biased once,
every predictor output biased with the empty context inside the device loop), or as the fork's
DEFAULT ``context_filter_state='on'`` does (greedy_search.py:100-150; transducer.py:525): per
predictor call a hot-word gate (``ContextBias.forward_hw_pred``), decisions 0 / 1 that swap which
stream is biased with the phrase list, and the "go-back" re-decode - all inside the device loop
(``b200_rnnt_greedy_gated``).
"""
from __future__ import annotations

import ctypes
from typing import List

import torch

from . import _lib
from ._lib import i32, ptr


def _model_tensors(model):
    pred, joint = model.predictor, model.joint
    rnn = getattr(pred, "rnn", None)
    if not isinstance(rnn, torch.nn.LSTM) or rnn.bidirectional or not rnn.bias or rnn.proj_size != 0:
        raise NotImplementedError("device search implements the LSTM RNNPredictor only "
                                  "(wenet/transducer/predictor.py:58-200, rnn_type='lstm')")
    for name in ("enc_ffn", "pred_ffn", "ffn_out"):
        if getattr(joint, name, None) is None:
            raise NotImplementedError(f"device search needs joint.{name} (prejoin_linear=True)")
    f32 = lambda t: t.detach().to(torch.float32).contiguous()
    L, H = rnn.num_layers, rnn.hidden_size
    lstm = []
    for l in range(L):
        lstm += [f32(getattr(rnn, f"weight_ih_l{l}")), f32(getattr(rnn, f"weight_hh_l{l}")),
                 f32(getattr(rnn, f"bias_ih_l{l}")), f32(getattr(rnn, f"bias_hh_l{l}"))]
    # projection followed by pred_ffn has no non-linearity in between: compose once in fp64
    wp, bp = pred.projection.weight.detach().double(), pred.projection.bias.detach().double()
    wf, bf = joint.pred_ffn.weight.detach().double(), joint.pred_ffn.bias.detach().double()
    w_c = (wf @ wp).to(torch.float32).contiguous()
    b_c = (wf @ bp + bf).to(torch.float32).contiguous()
    return dict(embed=f32(pred.embed.weight), lstm=lstm, w_c=w_c, b_c=b_c,
                w_out=f32(joint.ffn_out.weight), b_out=f32(joint.ffn_out.bias), L=L, H=H,
                Emb=pred.embed.weight.shape[1], J=joint.ffn_out.in_features,
                V=joint.ffn_out.out_features)


def _enc_proj(model, encoder_out: torch.Tensor) -> torch.Tensor:
    """enc_ffn(encoder_out) in full fp32 (one library GEMM, joint.py:55)."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        w, b = model.joint.enc_ffn.weight, model.joint.enc_ffn.bias
        return torch.nn.functional.linear(encoder_out.float(), w.float(), b.float()).contiguous()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


class _PredBias(ctypes.Structure):
    """b200_pred_bias_t (include/b200_transducer.h)."""
    _fields_ = ([(n, ctypes.c_void_p) for n in
                 ("proj_w", "proj_b", "wq", "bq", "k", "v", "wo", "bo", "ln1_g", "ln1_b", "wc", "bc",
                  "ln2_g", "ln2_b", "wf", "bf")] +
                [("P", ctypes.c_int), ("Nc", ctypes.c_int), ("heads", ctypes.c_int),
                 ("ln_eps", ctypes.c_float)])


class _Gate(ctypes.Structure):
    """b200_gate_t (include/b200_transducer.h)."""
    _fields_ = ([(n, ctypes.c_void_p) for n in
                 ("enc_proj_empty", "k0", "v0", "hq_w", "hq_b", "hk", "hv", "ho_w", "ho_b",
                  "hln_g", "hln_b", "hout_w", "hout_b", "gate_frames")] +
                [("Nc0", ctypes.c_int), ("Nc_hw", ctypes.c_int), ("hw_heads", ctypes.c_int),
                 ("n_labels", ctypes.c_int)])


def edit_distance(a, b) -> float:
    """Levenshtein distance, the ``dist`` the fork's greedy search returns
    (wenet/transducer/search/greedy_search.py:6-32,174): host bookkeeping on two short lists."""
    a, b = [int(x) for x in a], [int(x) for x in b]
    prev = list(range(len(b) + 1))
    for i, x in enumerate(a, 1):
        cur = [i] + [0] * len(b)
        for j, y in enumerate(b, 1):
            cur[j] = min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y))
        prev = cur
    return float(prev[-1])


def _has_hotword_bias(model) -> bool:
    cb = getattr(model, "context_bias", None)
    return cb is not None and hasattr(cb, "predictor_bias") and hasattr(cb, "predictor_bias_combine")


def _hotword_setup(model, encoder_out, context_list, context_lengths, gated: bool = False):
    """What greedy_search.py:62-70 does before its loop: phrase vectors of the given and of the
    EMPTY context list, encoder stream biased with the given one; plus K / V of the context the
    per-step predictor bias attends to (fixed for the whole call): the EMPTY list in
    ``basic_greedy_search`` (loss_mode 'pred', :140-144), the GIVEN list in
    ``basic_greedy_search_both`` (every other loss_mode, :357, transducer.py:559-597)."""
    cb = model.context_bias
    dev = encoder_out.device
    if context_list is None:
        context_list = torch.IntTensor([0])       # the reference's defaults (greedy_search.py:38-39)
    if context_lengths is None:
        context_lengths = torch.IntTensor([0])
    context_list, context_lengths = context_list.to(dev), context_lengths.to(dev)
    bias_hidden = cb.forward_bias_hidden(context_list, context_lengths)
    empty = torch.zeros((1, 1), dtype=torch.int, device=dev)
    bias_hidden_empty = cb.forward_bias_hidden(empty, context_lengths[0].unsqueeze(0))
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        encoder_plain = encoder_out
        encoder_out, encoder_out_bias = cb.forward_encoder_bias(bias_hidden, encoder_out)
        mha = cb.predictor_bias
        given = getattr(model, "loss_mode", "pred") != "pred"
        ctx = (bias_hidden if given else bias_hidden_empty)[0].float()
        f32 = lambda t: t.detach().to(torch.float32).contiguous()
        keep = dict(k=f32(mha.linear_k(ctx)), v=f32(mha.linear_v(ctx)))
        gate_keep = None
        if gated:   # greedy_search.py:70 / :338 (encoder stream over the EMPTY list)
            enc_empty, _ = cb.forward_encoder_bias(bias_hidden_empty, encoder_plain)
            # decision 0 takes the OTHER phrase list on the predictor stream (:140-144 / :399-400)
            c0 = (bias_hidden_empty if given else bias_hidden)[0].float()
            gate_keep = dict(enc_proj_empty=_enc_proj(model, enc_empty),
                             k0=f32(mha.linear_k(c0)), v0=f32(mha.linear_v(c0)))
            if given:
                # forward_hw_pred_both (:360, context_bias.py:388-394) attends over ONE key, the frame's
                # encoder bias: its softmax is 1 and the gate is a function of the frame alone, so the
                # module runs once over all frames (library call) and the loop reads a table
                Bq, Tq, Eq = encoder_out_bias.shape
                frames = encoder_out_bias.reshape(Bq * Tq, 1, Eq)
                hw_out = cb.forward_hw_pred_both(frames, torch.zeros_like(frames))
                gate_keep["gate_frames"] = hw_out.reshape(Bq, Tq, -1).argmax(-1).to(torch.int32).contiguous()
            else:   # :99-110: gate head over the predictor output, inside the device loop
                hw = cb.hw_bias
                cg = bias_hidden[0].float()
                gate_keep.update(hk=f32(hw.linear_k(cg)), hv=f32(hw.linear_v(cg)))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    pred, joint = model.predictor, model.joint
    named = dict(proj_w=pred.projection.weight, proj_b=pred.projection.bias,
                 wq=mha.linear_q.weight, bq=mha.linear_q.bias, wo=mha.linear_out.weight,
                 bo=mha.linear_out.bias, ln1_g=cb.predictor_bias_bias_norm.weight,
                 ln1_b=cb.predictor_bias_bias_norm.bias, wc=cb.predictor_bias_combine.weight,
                 bc=cb.predictor_bias_combine.bias, ln2_g=cb.predictor_bias_out_norm.weight,
                 ln2_b=cb.predictor_bias_out_norm.bias, wf=joint.pred_ffn.weight, bf=joint.pred_ffn.bias)
    keep.update({n: f32(t) for n, t in named.items()})
    P = pred.projection.out_features
    if keep["wc"].shape != (P, 2 * P) or keep["wq"].shape != (P, P):
        raise NotImplementedError("predictor bias expects embedding_size == predictor output size")
    desc = _PredBias(**{n: keep[n].data_ptr() for n in keep}, P=P, Nc=keep["k"].shape[0],
                     heads=int(mha.h), ln_eps=float(cb.predictor_bias_bias_norm.eps))
    if not gated:
        return encoder_out, desc, keep
    hw = cb.hw_bias
    if "gate_frames" not in gate_keep:
        if hw.linear_q.weight.shape != (P, P) or cb.hw_output_layer.in_features != P:
            raise NotImplementedError("forward_hw_pred needs unified_hw_odim == predictor output size "
                                      "(conf/encoder_bias_conformer_rnnt_4_head_bi_1_layer_pred.yaml:40,51)")
        gnamed = dict(hq_w=hw.linear_q.weight, hq_b=hw.linear_q.bias, ho_w=hw.linear_out.weight,
                      ho_b=hw.linear_out.bias, hln_g=cb.hw_bias_norm.weight, hln_b=cb.hw_bias_norm.bias,
                      hout_w=cb.hw_output_layer.weight, hout_b=cb.hw_output_layer.bias)
        gate_keep.update({n: f32(t) for n, t in gnamed.items()})
    gdesc = _Gate(**{n: gate_keep[n].data_ptr() for n in gate_keep}, Nc0=gate_keep["k0"].shape[0],
                  Nc_hw=gate_keep["hk"].shape[0] if "hk" in gate_keep else 0,
                  hw_heads=int(hw.h), n_labels=int(cb.hw_output_layer.out_features))
    return encoder_out, desc, (keep, gate_keep, gdesc)


@torch.no_grad()
def greedy_search_batch(model, encoder_out: torch.Tensor, encoder_out_lens: torch.Tensor,
                        n_steps: int = 64, return_rounds: bool = False, context_list=None,
                        context_lengths=None, hotword: bool = None, context_filter_state: str = "off",
                        return_decisions: bool = False):
    """Greedy RNN-T decode of a batch: encoder_out [B,T,E] -> B lists of token ids.

    When the model carries a real ``ContextBias`` (``model.context_bias.predictor_bias``) the
    fork's hot-word streams are applied as greedy_search.py:62-144 does with
    ``context_filter_state='off'``: the encoder stream is biased with ``context_list`` once, every
    predictor output with the empty context inside the device loop (``hotword=False`` turns it off).
    ``context_filter_state='on'`` runs the fork's default gated decode with go-back
    (greedy_search.py:100-150); ``return_decisions`` adds the per-utterance gate decisions (the
    reference's ``result`` list)."""
    _lib.require_cuda(encoder_out, "encoder_out")
    if encoder_out.dim() != 3:
        raise RuntimeError("encoder_out must be 3-D (batch, time, dim)")
    dev = encoder_out.device
    B, T, _ = encoder_out.shape
    m = _model_tensors(model)
    if hotword is None:
        hotword = _has_hotword_bias(model)
    gated = context_filter_state == "on"
    if gated and not hotword:
        raise RuntimeError("context_filter_state='on' needs a model with a ContextBias (model.context_bias)")
    bias_desc = gate_desc = None
    if hotword:
        with torch.cuda.device(dev):
            encoder_out, bias_desc, _keep_alive = _hotword_setup(model, encoder_out, context_list,
                                                                 context_lengths, gated=gated)
            if gated:
                gate_desc = _keep_alive[2]
    with torch.cuda.device(dev):
        lens = encoder_out_lens.reshape(-1).to(device=dev, dtype=torch.int32).clamp(max=T).contiguous()
        if lens.numel() != B:
            raise RuntimeError("encoder_out_lens must have one entry per utterance")
        enc_proj = _enc_proj(model, encoder_out)
        # a go-back decrements the per-frame emission count (greedy_search.py:124): one extra symbol per frame
        max_tokens = int(min(T * (n_steps + (1 if gated else 0)) + (8 if gated else 0), 1 << 20))
        tokens = torch.empty(B, max_tokens, dtype=torch.int32, device=dev)
        n_tok = torch.zeros(B, dtype=torch.int32, device=dev)
        L_ = _lib.lib()
        L_.b200_greedy_workspace_bytes.restype = ctypes.c_size_t
        L_.b200_greedy_gated_workspace_bytes.restype = ctypes.c_size_t
        ws_fn = L_.b200_greedy_gated_workspace_bytes if gated else L_.b200_greedy_workspace_bytes
        ws_bytes = ws_fn(B, m["V"], m["H"], m["L"], m["J"])
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        ptrs = (ctypes.c_void_p * (4 * m["L"]))(*[t.data_ptr() for t in m["lstm"]])
        rounds = ctypes.c_int(0)
        if bias_desc is None:
            _lib.call("b200_rnnt_greedy", ptr(enc_proj), ptr(lens), i32(B), i32(T), ptr(m["embed"]), ptrs,
                      ptr(m["w_c"]), ptr(m["b_c"]), ptr(m["w_out"]), ptr(m["b_out"]), i32(m["V"]),
                      i32(m["Emb"]), i32(m["H"]), i32(m["L"]), i32(m["J"]), i32(int(model.blank)),
                      i32(n_steps), ptr(tokens), ptr(n_tok), i32(max_tokens), ptr(ws),
                      ctypes.c_size_t(ws_bytes), ctypes.byref(rounds), _lib.stream_handle(dev))
        elif gated:
            max_dec = max_tokens + 1
            decisions = torch.empty(B, max_dec, dtype=torch.int32, device=dev)
            n_dec = torch.zeros(B, dtype=torch.int32, device=dev)
            _lib.call("b200_rnnt_greedy_gated", ptr(enc_proj), ptr(lens), i32(B), i32(T), ptr(m["embed"]),
                      ptrs, ptr(m["w_out"]), ptr(m["b_out"]), i32(m["V"]), i32(m["Emb"]), i32(m["H"]),
                      i32(m["L"]), i32(m["J"]), i32(int(model.blank)), i32(n_steps),
                      ctypes.byref(bias_desc), ctypes.byref(gate_desc), ptr(tokens), ptr(n_tok),
                      i32(max_tokens), ptr(decisions), ptr(n_dec), i32(max_dec), ptr(ws),
                      ctypes.c_size_t(ws_bytes), ctypes.byref(rounds), _lib.stream_handle(dev))
        else:
            _lib.call("b200_rnnt_greedy_bias", ptr(enc_proj), ptr(lens), i32(B), i32(T), ptr(m["embed"]),
                      ptrs, ptr(m["w_out"]), ptr(m["b_out"]), i32(m["V"]), i32(m["Emb"]), i32(m["H"]),
                      i32(m["L"]), i32(m["J"]), i32(int(model.blank)), i32(n_steps),
                      ctypes.byref(bias_desc), ptr(tokens), ptr(n_tok), i32(max_tokens), ptr(ws),
                      ctypes.c_size_t(ws_bytes), ctypes.byref(rounds), _lib.stream_handle(dev))
        n_host = n_tok.cpu().tolist()  # one D2H copy of the results at the end
        width = max(n_host) if n_host else 0
        tok_host = tokens[:, :max(width, 1)].cpu().tolist()
        dec_host = None
        if gated and return_decisions:
            nd = n_dec.cpu().tolist()
            dh = decisions[:, :max(max(nd), 1)].cpu().tolist()
            dec_host = [dh[b][:nd[b]] for b in range(B)]
    hyps = [tok_host[b][:min(n_host[b], max_tokens)] for b in range(B)]
    out = (hyps,)
    if return_decisions:
        out += (dec_host if dec_host is not None else [[1] * (len(h) + 1) for h in hyps],)
    if return_rounds:
        out += (rounds.value,)
    return out if len(out) > 1 else hyps


def basic_greedy_search(model, encoder_out: torch.Tensor, encoder_out_lens, context_list=None,
                        context_lengths=None, n_steps: int = 64, context_filter_state: str = "off",
                        context_decoder_labels_padded=None) -> List[List[int]]:
    """Single-utterance entry point with the reference's argument order.  Returns ``[hyps]`` (the
    upstream shape, "greedy_search copy.py":63); with ``context_decoder_labels_padded`` given it
    returns the fork's triple ``([hyps], dist, result)`` (greedy_search.py:174-176): ``result`` =
    the gate decisions, ``dist`` = their edit distance to the labels."""
    if context_filter_state not in ("on", "off"):
        raise ValueError("context_filter_state must be 'on' or 'off'")
    if encoder_out.size(0) != 1:
        raise RuntimeError("basic_greedy_search takes one utterance; use greedy_search_batch")
    lens = torch.as_tensor(encoder_out_lens).reshape(-1)[:1]
    if context_filter_state == "off" and context_decoder_labels_padded is None:
        return [greedy_search_batch(model, encoder_out, lens, n_steps=n_steps, context_list=context_list,
                                    context_lengths=context_lengths)[0]]
    hyps, decisions = greedy_search_batch(model, encoder_out, lens, n_steps=n_steps,
                                          context_list=context_list, context_lengths=context_lengths,
                                          context_filter_state=context_filter_state, return_decisions=True)
    if context_decoder_labels_padded is None:
        return [hyps[0]]
    labels = torch.as_tensor(context_decoder_labels_padded).reshape(-1).tolist()
    return [hyps[0]], edit_distance(labels, decisions[0]), decisions[0]


# --------------------------------------------------------------------------- prefix beam search
class Sequence:
    """Hypothesis record with the reference's fields (prefix_beam_search.py:7-19).  ``hyp``
    starts with the seed blank; ``cache`` is not populated (callers only read ``hyp`` and
    ``score``: transducer.py:377,444-445)."""
    __slots__ = ("hyp", "score", "cache")

    def __init__(self, hyp, score, cache=None):
        self.hyp, self.score, self.cache = hyp, score, cache


@torch.no_grad()
def prefix_beam_search_batch(model, encoder_out: torch.Tensor, encoder_out_lens: torch.Tensor,
                             beam_size: int = 10, ctc_weight: float = 0.3,
                             transducer_weight: float = 0.7) -> List[List[Sequence]]:
    """RNN-T prefix beam search with CTC shallow fusion for a batch: encoder_out [B,T,E] ->
    per utterance the final beam (best first), exactly one expansion per frame."""
    _lib.require_cuda(encoder_out, "encoder_out")
    dev = encoder_out.device
    B, T, _ = encoder_out.shape
    m = _model_tensors(model)
    with torch.cuda.device(dev):
        lens = encoder_out_lens.reshape(-1).to(device=dev, dtype=torch.int32).clamp(max=T).contiguous()
        if lens.numel() != B:
            raise RuntimeError("encoder_out_lens must have one entry per utterance")
        enc_proj = _enc_proj(model, encoder_out)
        prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = False
        try:  # CTC.log_softmax (ctc.py:66-74): library GEMM + log_softmax, once per utterance
            ctc_lp = model.ctc.log_softmax(encoder_out.float()).float().contiguous()
        finally:
            torch.backends.cuda.matmul.allow_tf32 = prev
        ML = T + 2
        out_hyp = torch.empty(B, beam_size, ML, dtype=torch.int32, device=dev)
        out_len = torch.empty(B, beam_size, dtype=torch.int32, device=dev)
        out_score = torch.empty(B, beam_size, dtype=torch.float64, device=dev)
        out_n = torch.empty(B, dtype=torch.int32, device=dev)
        L_ = _lib.lib()
        L_.b200_beam_workspace_bytes.restype = ctypes.c_size_t
        ws_bytes = L_.b200_beam_workspace_bytes(B, T, m["V"], m["H"], m["L"], m["J"], beam_size)
        if ws_bytes == 0:
            raise _lib.B200Error("b200_beam_workspace_bytes: bad shape")
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        ptrs = (ctypes.c_void_p * (4 * m["L"]))(*[t.data_ptr() for t in m["lstm"]])
        _lib.call("b200_rnnt_beam", ptr(enc_proj), ptr(ctc_lp), ptr(lens), i32(B), i32(T),
                  ptr(m["embed"]), ptrs, ptr(m["w_c"]), ptr(m["b_c"]), ptr(m["w_out"]), ptr(m["b_out"]),
                  i32(m["V"]), i32(m["Emb"]), i32(m["H"]), i32(m["L"]), i32(m["J"]),
                  i32(int(model.blank)), i32(beam_size), ctypes.c_float(ctc_weight),
                  ctypes.c_float(transducer_weight), ptr(out_hyp), ptr(out_len), ptr(out_score),
                  ptr(out_n), ptr(ws), ctypes.c_size_t(ws_bytes), _lib.stream_handle(dev))
        n = out_n.cpu().tolist()
        lens_h = out_len.cpu().tolist()
        scores = out_score.cpu().tolist()
        width = max(max(r) for r in lens_h) if B else 0
        hyps = out_hyp[:, :, :max(width, 1)].cpu().tolist()
    return [[Sequence(hyps[b][k][:lens_h[b][k]], scores[b][k]) for k in range(n[b])] for b in range(B)]


class PrefixBeamSearch:
    """Same constructor and method as the reference class (prefix_beam_search.py:22-148)."""

    def __init__(self, encoder, predictor, joint, ctc, blank):
        self.encoder, self.predictor, self.joint, self.ctc, self.blank = encoder, predictor, joint, ctc, blank

    def prefix_beam_search(self, speech, speech_lengths, decoding_chunk_size: int = -1,
                           beam_size: int = 5, num_decoding_left_chunks: int = -1,
                           simulate_streaming: bool = False, ctc_weight: float = 0.3,
                           transducer_weight: float = 0.7):
        assert speech.shape[0] == speech_lengths.shape[0]
        assert decoding_chunk_size != 0
        encoder_out, _ = self.encoder(speech, speech_lengths, decoding_chunk_size,
                                      num_decoding_left_chunks)
        # the reference runs all maxlen frames of its single utterance (:75)
        lens = torch.full((encoder_out.size(0),), encoder_out.size(1), dtype=torch.int32)
        beams = prefix_beam_search_batch(self, encoder_out, lens, beam_size, ctc_weight,
                                         transducer_weight)
        return (beams[0] if encoder_out.size(0) == 1 else beams), encoder_out
