"""oracle/reference.py — TEST INFRASTRUCTURE, NOT THE PRODUCT.

CPU restatement (PyTorch fp32 on CPU + the same third-party kernels the reference
calls) of the transducer hot path of Vindicator645/wenet-celoss.  Only tests/,
``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` / ``--impl reference``
legs may import this module; nothing under wenet_celoss_b200/ does.

What is restated, and from where (paths under /root/reference):
  joint_logits            wenet/transducer/joint.py:45-70
  OracleJoint             wenet/transducer/joint.py:11-43 (parameter names kept)
  OraclePredictor         wenet/transducer/predictor.py:58-200 (RNNPredictor, lstm)
  OracleCTC               wenet/transformer/ctc.py:21-84
  add_blank               wenet/utils/common.py:56-87
  log_add                 wenet/utils/common.py:268-276
  transducer_loss         wenet/transducer/transducer.py:124-147
  transducer_score        wenet/transducer/transducer.py:277-302
  greedy_search           wenet/transducer/search/greedy_search copy.py:6-63 and
                          wenet/transducer/search/greedy_search.py:34-176 with
                          context_filter_state='off' and an identity context bias
  prefix_beam_search      wenet/transducer/search/prefix_beam_search.py:22-148

Third-party arithmetic is NOT re-derived here: like the reference, this module
calls torchaudio.functional.rnnt_loss (reference pin torchaudio==0.10.0,
README.md:57) and torch's CTC loss / LSTM (pin torch==1.10.0).  The fp64 C
restatement of those two recurrences lives in oracle/lattice_ref.c.

Pinning: the reference's own tests hold no vectors for this path.
oracle/gen_golden.py runs the real reference modules from /root/reference on
seeded inputs, checks every function here against them and writes the fixtures
in tests/golden/ that the CPU test-suite replays.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F
from torch import nn


# --------------------------------------------------------------------------- helpers
def add_blank(ys_pad: torch.Tensor, blank: int, ignore_id: int) -> torch.Tensor:
    """[B,L] -> [B,L+1]: blank in column 0, padding (ignore_id) replaced by blank."""
    head = ys_pad.new_full((ys_pad.size(0), 1), blank)
    ys = torch.cat([head, ys_pad], dim=1)
    return ys.masked_fill(ys == ignore_id, blank)


def log_add(values: Sequence[float]) -> float:
    """log(sum(exp(v))) in Python double; all -inf stays -inf."""
    if all(v == -math.inf for v in values):
        return -math.inf
    top = max(values)
    return top + math.log(sum(math.exp(v - top) for v in values))


# --------------------------------------------------------------------------- modules
class OracleJoint(nn.Module):
    """Add-mode joint with tanh: parameter names match TransducerJoint."""

    def __init__(self, vocab: int, enc_dim: int, pred_dim: int, join_dim: int):
        super().__init__()
        self.enc_ffn = nn.Linear(enc_dim, join_dim)
        self.pred_ffn = nn.Linear(pred_dim, join_dim)
        self.ffn_out = nn.Linear(join_dim, vocab)

    def forward(self, enc_out: torch.Tensor, pred_out: torch.Tensor) -> torch.Tensor:
        return joint_logits(enc_out, pred_out, self)


def joint_logits(enc_out: torch.Tensor, pred_out: torch.Tensor, joint) -> torch.Tensor:
    """[B,T,E], [B,U1,P] -> [B,T,U1,V] = ffn_out(tanh(enc_ffn(e)[:, :, None] + pred_ffn(p)[:, None]))."""
    e = F.linear(enc_out, joint.enc_ffn.weight, joint.enc_ffn.bias)
    p = F.linear(pred_out, joint.pred_ffn.weight, joint.pred_ffn.bias)
    h = torch.tanh(e[:, :, None, :] + p[:, None, :, :])
    return F.linear(h, joint.ffn_out.weight, joint.ffn_out.bias)


class OraclePredictor(nn.Module):
    """Embedding -> dropout -> LSTM stack -> Linear; names match RNNPredictor."""

    def __init__(self, vocab: int, embed: int, out_dim: int, hidden: int, layers: int,
                 embed_dropout: float = 0.1, dropout: float = 0.1):
        super().__init__()
        self.n_layers = layers
        self.hidden_size = hidden
        self.embed = nn.Embedding(vocab, embed)
        self.dropout = nn.Dropout(embed_dropout)
        self.rnn = nn.LSTM(input_size=embed, hidden_size=hidden, num_layers=layers, bias=True,
                           batch_first=True, dropout=dropout)
        self.projection = nn.Linear(hidden, out_dim)

    def init_state(self, batch_size: int, device, method: str = "zero") -> List[torch.Tensor]:
        z = torch.zeros(self.n_layers, batch_size, self.hidden_size, device=device)
        return [z, z.clone()]

    def forward(self, tokens: torch.Tensor, cache: Optional[List[torch.Tensor]] = None):
        state = cache if cache is not None else self.init_state(tokens.size(0), tokens.device)
        out, _ = self.rnn(self.dropout(self.embed(tokens)), (state[0], state[1]))
        return self.projection(out)

    def forward_step(self, token: torch.Tensor, padding: torch.Tensor, cache: List[torch.Tensor]):
        h0, c0 = cache
        out, (h1, c1) = self.rnn(self.dropout(self.embed(token)), (h0, c0))
        out = self.projection(out)
        # padding==1 keeps the old state (always 0 on the search paths)
        h1 = padding * h0 + h1 * (1 - padding)
        c1 = padding * c0 + c1 * (1 - padding)
        return out, [h1, c1]

    def cache_to_batch(self, caches):
        return [torch.cat([c[0] for c in caches], dim=1), torch.cat([c[1] for c in caches], dim=1)]

    def batch_to_cache(self, cache):
        hs, cs = cache
        return [[hs[:, i:i + 1], cs[:, i:i + 1]] for i in range(hs.size(1))]


class OracleCTC(nn.Module):
    def __init__(self, vocab: int, enc_dim: int):
        super().__init__()
        self.ctc_lo = nn.Linear(enc_dim, vocab)

    def forward(self, hs_pad, hlens, ys_pad, ys_lens):
        return ctc_forward(hs_pad, hlens, ys_pad, ys_lens, self)

    def log_softmax(self, hs_pad):
        return F.log_softmax(self.ctc_lo(hs_pad), dim=2)

    def argmax(self, hs_pad):
        return torch.argmax(self.ctc_lo(hs_pad), dim=2)


def ctc_forward(hs_pad, hlens, ys_pad, ys_lens, ctc) -> torch.Tensor:
    """sum_b nll_b / B with blank=0, zero_infinity=False (ctc.py:57-63)."""
    logp = F.log_softmax(ctc.ctc_lo(hs_pad).transpose(0, 1), dim=2)  # [T,B,V]
    total = F.ctc_loss(logp, ys_pad, hlens, ys_lens, blank=0, reduction="sum", zero_infinity=False)
    return total / logp.size(1)


class OracleModel(nn.Module):
    """What the search functions read from the model: predictor, joint, ctc, blank."""

    def __init__(self, vocab: int, enc_dim: int = 256, pred_dim: int = 256, join_dim: int = 512,
                 embed: int = 256, hidden: int = 256, layers: int = 2, blank: int = 0):
        super().__init__()
        self.blank = blank
        self.vocab_size = vocab
        self.predictor = OraclePredictor(vocab, embed, pred_dim, hidden, layers)
        self.joint = OracleJoint(vocab, enc_dim, pred_dim, join_dim)
        self.ctc = OracleCTC(vocab, enc_dim)


# --------------------------------------------------------------------------- losses
def rnnt_targets(text: torch.Tensor, text_lengths: torch.Tensor, enc_lens: torch.Tensor,
                 ignore_id: int = -1):
    """int32 zero-padded targets and int32 lengths (transducer.py:137-141)."""
    t = text.to(torch.int64)
    t = torch.where(t == ignore_id, torch.zeros_like(t), t).to(torch.int32)
    return t, enc_lens.to(torch.int32), text_lengths.to(torch.int32)


def transducer_loss(model, enc_out, enc_lens, text, text_lengths, reduction="mean",
                    ignore_id: int = -1) -> torch.Tensor:
    """predictor -> joint -> torchaudio rnnt_loss, exactly the chain at transducer.py:124-147."""
    import torchaudio
    ys_in = add_blank(text, model.blank, ignore_id)
    pred_out = model.predictor(ys_in)
    logits = joint_logits(enc_out, pred_out, model.joint)
    tgt, tl, ul = rnnt_targets(text, text_lengths, enc_lens, ignore_id)
    return torchaudio.functional.rnnt_loss(logits, tgt, tl, ul, blank=model.blank, reduction=reduction)


def transducer_score(model, enc_out, enc_lens, hyps_pad, hyps_lens, ignore_id: int = -1):
    """-cost per hypothesis (transducer.py:277-302)."""
    return -transducer_loss(model, enc_out, enc_lens, hyps_pad, hyps_lens, reduction="none",
                            ignore_id=ignore_id)


# --------------------------------------------------------------------------- greedy
@torch.no_grad()
def greedy_search(model, enc_out: torch.Tensor, enc_len: int, n_steps: int = 64,
                  record_margin: bool = False):
    """One utterance ([1,T,E]).  Returns the emitted token list (and, optionally, the
    smallest top-1/top-2 log-prob margin seen, to attribute mismatches to ties)."""
    dev = enc_out.device
    blank = model.blank
    pad = torch.zeros(1, 1, device=dev)
    token = torch.tensor([[blank]], device=dev)
    state = model.predictor.init_state(1, device=dev)
    pending_state = state
    pred_step = None
    need_predictor = True
    emitted_here = 0
    t = 0
    hyp: List[int] = []
    min_margin = math.inf
    while t < int(enc_len):
        if need_predictor:
            pred_step, pending_state = model.predictor.forward_step(token, pad, state)
        logp = joint_logits(enc_out[:, t:t + 1, :], pred_step, model.joint).log_softmax(dim=-1)
        flat = logp.reshape(-1)
        k = int(flat.argmax())
        if record_margin:
            top2 = flat.topk(2).values
            min_margin = min(min_margin, float(top2[0] - top2[1]))
        if k != blank:
            hyp.append(k)
            need_predictor = True
            emitted_here += 1
            token = torch.tensor([[k]], device=dev)
            state = pending_state
        if k == blank or emitted_here >= n_steps:
            if k == blank:
                need_predictor = False
            t += 1
            emitted_here = 0
    return (hyp, min_margin) if record_margin else hyp


# --------------------------------------------------------------------------- beam
class Hyp:
    __slots__ = ("hyp", "score", "cache")

    def __init__(self, hyp, score, cache):
        self.hyp, self.score, self.cache = hyp, score, cache


@torch.no_grad()
def prefix_beam_search(model, enc_out: torch.Tensor, beam_size: int = 10, ctc_weight: float = 0.3,
                       transducer_weight: float = 0.7) -> List[Hyp]:
    """One utterance ([1,T,E]); exactly one expansion per frame; returns the beam
    (each ``hyp`` starts with the blank the search was seeded with)."""
    dev = enc_out.device
    blank = model.blank
    T = enc_out.size(1)
    ctc_lp = model.ctc.log_softmax(enc_out).squeeze(0)  # [T,V]
    beam = [Hyp([blank], 0.0, model.predictor.init_state(1, device=dev))]
    for i in range(T):
        last = torch.tensor([h.hyp[-1] for h in beam], dtype=torch.int, device=dev)
        state = model.predictor.cache_to_batch([h.cache for h in beam])
        base = torch.tensor([h.score for h in beam]).to(dev)  # python float -> fp32
        pad = torch.zeros(last.size(0), 1, device=dev)
        pred, new_state = model.predictor.forward_step(last.unsqueeze(-1), pad, state)
        logp = joint_logits(enc_out[:, i:i + 1, :], pred, model.joint).log_softmax(dim=-1)
        logp = logp.squeeze(1).squeeze(1)  # [N,V]
        per_hyp_state = model.predictor.batch_to_cache(new_state)
        fused = torch.log(transducer_weight * torch.exp(logp) + ctc_weight * torch.exp(ctc_lp[i].unsqueeze(0)))
        top_lp, top_ix = fused.topk(beam_size)
        cand = base.unsqueeze(1) + top_lp  # fp32
        expanded: List[Hyp] = []
        for j, h in enumerate(beam):
            for k in range(beam_size):
                sym = int(top_ix[j, k])
                sc = cand[j, k].item()
                if sym == blank:
                    expanded.append(Hyp(list(h.hyp), sc, h.cache))
                else:
                    expanded.append(Hyp(h.hyp + [sym], sc, per_hyp_state[j]))
        merged: List[Hyp] = [expanded[0]]
        for cnd in expanded[1:]:
            for kept in merged:
                if kept.hyp == cnd.hyp:
                    kept.score = log_add([kept.score, cnd.score])
                    break
            else:
                merged.append(cnd)
        merged.sort(key=lambda h: h.score, reverse=True)  # stable
        beam = merged[:beam_size]
    return beam


def beam_search_best(model, enc_out, **kw) -> Tuple[List[int], float]:
    """(hyp without the seed blank, score) as Transducer.beam_search returns (transducer.py:377)."""
    beam = prefix_beam_search(model, enc_out, **kw)
    return beam[0].hyp[1:], beam[0].score


def ctc_prefix_beam_search(ctc_probs: torch.Tensor, beam_size: int, blank: int = 0):
    """CTC prefix beam search over the log-probs of one utterance ``[T,V]`` (restates
    wenet/transformer/asr_model.py:370-408; blank is hard-wired to 0 there).  Returns
    ``[(prefix tuple, log_add(pb, pnb)), ...]`` best first."""
    from collections import defaultdict
    ninf = -float("inf")
    cur = [(tuple(), (0.0, ninf))]
    for t in range(ctc_probs.size(0)):
        logp = ctc_probs[t]
        nxt = defaultdict(lambda: (ninf, ninf))
        _, top_ix = logp.topk(beam_size)
        for s in top_ix.tolist():
            ps = logp[s].item()
            for prefix, (pb, pnb) in cur:
                last = prefix[-1] if prefix else None
                if s == blank:
                    n_pb, n_pnb = nxt[prefix]
                    nxt[prefix] = (log_add([n_pb, pb + ps, pnb + ps]), n_pnb)
                elif s == last:
                    n_pb, n_pnb = nxt[prefix]
                    nxt[prefix] = (n_pb, log_add([n_pnb, pnb + ps]))
                    ext = prefix + (s,)
                    n_pb, n_pnb = nxt[ext]
                    nxt[ext] = (n_pb, log_add([n_pnb, pb + ps]))
                else:
                    ext = prefix + (s,)
                    n_pb, n_pnb = nxt[ext]
                    nxt[ext] = (n_pb, log_add([n_pnb, pb + ps, pnb + ps]))
        ranked = sorted(nxt.items(), key=lambda kv: log_add(list(kv[1])), reverse=True)  # stable
        cur = ranked[:beam_size]
    return [(prefix, log_add([pb, pnb])) for prefix, (pb, pnb) in cur]


def forced_align(ctc_probs: torch.Tensor, y: torch.Tensor, blank_id: int = 0) -> List[int]:
    """CTC forced alignment of one utterance (restates wenet/utils/ctc_util.py:27-83 with numpy
    rows instead of per-cell tensors; fp32 max/add; Python's negative-index wrap of state 0 kept)."""
    import numpy as np
    lp = ctc_probs.detach().cpu().numpy().astype(np.float32)
    T = lp.shape[0]
    lab = [blank_id]
    for tok in y.tolist():
        lab += [int(tok), blank_id]
    S = len(lab)
    alpha = np.full((T, S), -np.inf, dtype=np.float32)
    path = np.full((T, S), -1, dtype=np.int64)
    alpha[0, 0] = lp[0, lab[0]]
    alpha[0, 1] = lp[0, lab[1]]
    for t in range(1, T):
        for s in range(S):
            prev = [s, s - 1]
            if not (lab[s] == blank_id or s < 2 or lab[s] == lab[s - 2]):
                prev.append(s - 2)
            cand = np.array([alpha[t - 1, q] for q in prev], dtype=np.float32)  # q = -1 wraps
            k = int(np.argmax(cand))  # first maximum
            alpha[t, s] = cand[k] + lp[t, lab[s]]
            path[t, s] = prev[k]
    st = [S - 1, S - 2][int(np.argmax(np.array([alpha[-1, S - 1], alpha[-1, S - 2]])))]
    seq = [0] * T
    seq[-1] = st
    for t in range(T - 2, -1, -1):
        seq[t] = int(path[t + 1, seq[t + 1]])
    return [lab[q] for q in seq]


def predictor_bias(cb, ctx: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """ContextBias.forward_predictor_bias (wenet/transformer/context_bias.py:375-381) restated
    with functional ops: ``x [1,1,P]`` predictor output, ``ctx [1,Nc,P]`` phrase vectors; ``cb``
    only supplies the parameters (predictor_bias.{linear_q,linear_k,linear_v,linear_out},
    predictor_bias_bias_norm, predictor_bias_combine, predictor_bias_out_norm)."""
    import torch.nn.functional as F
    mha = cb.predictor_bias
    nh = int(mha.h)
    P = x.size(-1)
    dk = P // nh
    q = F.linear(x, mha.linear_q.weight, mha.linear_q.bias).view(1, -1, nh, dk).transpose(1, 2)
    k = F.linear(ctx, mha.linear_k.weight, mha.linear_k.bias).view(1, -1, nh, dk).transpose(1, 2)
    v = F.linear(ctx, mha.linear_v.weight, mha.linear_v.bias).view(1, -1, nh, dk).transpose(1, 2)
    att = torch.softmax(q @ k.transpose(-2, -1) / math.sqrt(dk), dim=-1)   # attention.py:48-120
    a = (att @ v).transpose(1, 2).reshape(1, -1, P)
    a = F.linear(a, mha.linear_out.weight, mha.linear_out.bias)
    n1, n2 = cb.predictor_bias_bias_norm, cb.predictor_bias_out_norm
    a = F.layer_norm(a, (P,), n1.weight, n1.bias, n1.eps)
    y = F.linear(torch.cat([x, a], dim=-1), cb.predictor_bias_combine.weight, cb.predictor_bias_combine.bias)
    return F.layer_norm(y, (P,), n2.weight, n2.bias, n2.eps)


@torch.no_grad()
def greedy_search_hotword(model, enc_biased: torch.Tensor, enc_len: int, bias_hidden_empty: torch.Tensor,
                          n_steps: int = 64) -> List[int]:
    """The fork's greedy search with ``context_filter_state='off'``
    (wenet/transducer/search/greedy_search.py:62-165): ``enc_biased`` is the encoder stream after
    forward_encoder_bias with the given phrase list (:69), every predictor output passes through
    forward_predictor_bias with the EMPTY phrase list (:140-144) before the joint."""
    dev = enc_biased.device
    blank = model.blank
    pad = torch.zeros(1, 1, device=dev)
    token = torch.tensor([[blank]], device=dev)
    state = model.predictor.init_state(1, device=dev)
    pending_state = state
    pred_step = None
    need_predictor = True
    emitted_here = 0
    t = 0
    hyp: List[int] = []
    while t < int(enc_len):
        if need_predictor:
            raw, pending_state = model.predictor.forward_step(token, pad, state)
            pred_step = predictor_bias(model.context_bias, bias_hidden_empty, raw)
        logp = joint_logits(enc_biased[:, t:t + 1, :], pred_step, model.joint).log_softmax(dim=-1)
        k = int(logp.reshape(-1).argmax())
        if k != blank:
            hyp.append(k)
            need_predictor = True
            emitted_here += 1
            token = torch.tensor([[k]], device=dev)
            state = pending_state
        if k == blank or emitted_here >= n_steps:
            if k == blank:
                need_predictor = False
            t += 1
            emitted_here = 0
    return hyp


def hw_pred_gate(cb, ctx: torch.Tensor, y: torch.Tensor) -> int:
    """ContextBias.forward_hw_pred (wenet/transformer/context_bias.py:382-387) + ``topk(1)`` index
    (greedy_search.py:103-110): multi-head attention ``hw_bias`` of the BIASED predictor output
    ``y [1,1,D]`` over the phrase vectors ``ctx [1,Nc,D]``, ``hw_bias_norm``, ``hw_output_layer``;
    returns the arg-max label (the gate).  Needs unified_hw_odim == embedding size, as in
    conf/encoder_bias_conformer_rnnt_4_head_bi_1_layer_pred.yaml:40,51."""
    import torch.nn.functional as F
    mha = cb.hw_bias
    nh = int(mha.h)
    D = y.size(-1)
    dk = D // nh
    q = F.linear(y, mha.linear_q.weight, mha.linear_q.bias).view(1, -1, nh, dk).transpose(1, 2)
    k = F.linear(ctx, mha.linear_k.weight, mha.linear_k.bias).view(1, -1, nh, dk).transpose(1, 2)
    v = F.linear(ctx, mha.linear_v.weight, mha.linear_v.bias).view(1, -1, nh, dk).transpose(1, 2)
    att = torch.softmax(q @ k.transpose(-2, -1) / math.sqrt(dk), dim=-1)
    a = (att @ v).transpose(1, 2).reshape(1, -1, D)
    a = F.linear(a, mha.linear_out.weight, mha.linear_out.bias)
    a = F.layer_norm(a, (D,), cb.hw_bias_norm.weight, cb.hw_bias_norm.bias, cb.hw_bias_norm.eps)
    out = F.linear(a, cb.hw_output_layer.weight, cb.hw_output_layer.bias).reshape(-1)
    return int(out.topk(1)[1].item())


@torch.no_grad()
def hw_pred_both_gate(cb, enc_bias_frame: torch.Tensor) -> int:
    """ContextBias.forward_hw_pred_both (wenet/transformer/context_bias.py:388-394) + ``topk(1)``
    index as basic_greedy_search_both uses it (greedy_search.py:360-366): the attention
    ``hw_bias(h_dec, h_enc, h_enc)`` has ONE key / value (``h_enc = hw_output_layer_enc(encoder bias
    of the frame)``, [1,1,D]); a softmax over one score is exactly 1, so the attention output is
    ``linear_out(linear_v(h_enc))`` whatever the predictor-side query is."""
    import torch.nn.functional as F
    mha = cb.hw_bias
    h_enc = F.linear(enc_bias_frame, cb.hw_output_layer_enc.weight, cb.hw_output_layer_enc.bias)
    a = F.linear(F.linear(h_enc, mha.linear_v.weight, mha.linear_v.bias), mha.linear_out.weight, mha.linear_out.bias)
    D = a.size(-1)
    a = F.layer_norm(a, (D,), cb.hw_bias_norm.weight, cb.hw_bias_norm.bias, cb.hw_bias_norm.eps)
    out = F.linear(a, cb.hw_output_layer.weight, cb.hw_output_layer.bias).reshape(-1)
    return int(out.topk(1)[1].item())


@torch.no_grad()
def greedy_search_hotword_gated(model, enc_biased: torch.Tensor, enc_empty: torch.Tensor, enc_len: int,
                                bias_hidden: torch.Tensor, bias_hidden_empty: torch.Tensor,
                                n_steps: int = 64, enc_bias: torch.Tensor = None):
    """The fork's DEFAULT greedy search, ``context_filter_state='on'``
    (wenet/transducer/search/greedy_search.py:88-165; default at transducer.py:525 and
    wenet/bin/recognize.py:139-141), restated as a state machine without the reference's
    ``step_list / cache_list / input_list`` stacks.  Per predictor call (:92-99):
      gate = argmax forward_hw_pred(bias_hidden, forward_predictor_bias(bias_hidden, raw))  (:100-110)
      not going back, gate 0           -> decision 0, remember this frame as ``last_t``     (:111-114)
      not going back, gate 1 after a 0 -> GO BACK (:116-131): drop the last decision and the last
          token, rewind ``t`` to ``last_t``, restore the predictor input and LSTM state of the
          PREVIOUS call (the stacks' [-1] after the pops), decrement the per-frame emission count,
          and re-run that call with decisions forced to 1 until a call at ``t >= go_back_end``
      otherwise                        -> decision 1                                         (:132-137)
    decision 1: predictor stream biased with the EMPTY list, encoder stream biased with the
    given list; decision 0: predictor stream biased with the given list, encoder stream biased
    with the empty list (:140-150).  Only the previous call is ever restored (two go-backs need
    a decision-0 call between them), so three LSTM states suffice: before the previous call,
    before this call, after this call.  Returns (hyp, decisions).

    With ``enc_bias`` (the second output of forward_encoder_bias, [1,T,E]) the loop is
    ``basic_greedy_search_both`` (loss_mode 'both', greedy_search.py:339-407) instead: the gate is
    forward_hw_pred_both of the frame's encoder bias (:360), and decision 1 keeps the GIVEN list on
    the predictor stream while decision 0 takes the EMPTY one (:358,399-400)."""
    both = enc_bias is not None
    dev = enc_biased.device
    blank = model.blank
    cb = model.context_bias
    pad = torch.zeros(1, 1, device=dev)
    token = torch.tensor([[blank]], device=dev)
    state = model.predictor.init_state(1, device=dev)
    pending_state = state
    prev_call = None                 # (input token, LSTM state before the call) of the previous call
    pred_step = None
    need_predictor = True
    emitted_here = 0
    t = last_t = 0
    going_back, go_back_end = False, -1
    hyp: List[int] = []
    result: List[int] = []
    while t < int(enc_len):
        if need_predictor:
            raw, pending_state = model.predictor.forward_step(token, pad, state)
            y_given = predictor_bias(cb, bias_hidden, raw)
            gate = (hw_pred_both_gate(cb, enc_bias[:, t:t + 1, :]) if both
                    else hw_pred_gate(cb, bias_hidden, y_given))
            if not going_back:
                if gate == 0:
                    result.append(0)
                    last_t = t
                elif result and result[-1] == 0:
                    go_back_end, t, going_back = t, last_t, True
                    result.pop()
                    hyp.pop()
                    emitted_here -= 1
                    token, state = prev_call
                    continue         # prev_call stays: the re-run call has the same input and state
                else:
                    result.append(1)
            else:
                result.append(1)
                if t >= go_back_end:
                    going_back = False
            prev_call = (token, state)
            if both:
                pred_step = y_given if result[-1] == 1 else predictor_bias(cb, bias_hidden_empty, raw)
            else:
                pred_step = predictor_bias(cb, bias_hidden_empty, raw) if result[-1] == 1 else y_given
        enc = enc_biased if result[-1] == 1 else enc_empty
        logp = joint_logits(enc[:, t:t + 1, :], pred_step, model.joint).log_softmax(dim=-1)
        k = int(logp.reshape(-1).argmax())
        if k != blank:
            hyp.append(k)
            need_predictor = True
            emitted_here += 1
            token = torch.tensor([[k]], device=dev)
            state = pending_state
        if k == blank or emitted_here >= n_steps:
            if k == blank:
                need_predictor = False
            t += 1
            emitted_here = 0
    return hyp, result


def edit_distance(a, b) -> float:
    """Levenshtein distance of two label sequences, as the helper at
    wenet/transducer/search/greedy_search.py:6-32 (returned as ``dist`` by the fork's greedy)."""
    import numpy as np
    a, b = list(a), list(b)
    prev = np.arange(len(b) + 1, dtype=np.float64)
    for i in range(1, len(a) + 1):
        cur = np.empty_like(prev)
        cur[0] = i
        for j in range(1, len(b) + 1):
            cur[j] = min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (0 if a[i - 1] == b[j - 1] else 1))
        prev = cur
    return float(prev[len(b)])


def hw_ce_loss(hw_output, labels, ignore_index: int = -100):
    """nn.CrossEntropyLoss() as Transducer.forward applies it to the hot-word head
    (wenet/transducer/transducer.py:66,181,189,199): ``hw_output [B,C,U]`` (class dim 1), ``labels
    [B,U]``; mean over labels != ignore_index.  numpy fp64; returns (loss, d loss / d hw_output)."""
    import numpy as np
    x = np.asarray(hw_output, dtype=np.float64)
    y = np.asarray(labels)
    m = x.max(axis=1, keepdims=True)
    lse = m + np.log(np.exp(x - m).sum(axis=1, keepdims=True))
    logp = x - lse                                   # [B,C,U]
    valid = y != ignore_index
    n = int(valid.sum())
    ysafe = np.where(valid, y, 0)
    picked = np.take_along_axis(logp, ysafe[:, None, :], axis=1)[:, 0, :]
    loss = float(-(picked * valid).sum() / n) if n else float("nan")
    onehot = np.zeros_like(x)
    np.put_along_axis(onehot, ysafe[:, None, :], 1.0, axis=1)
    grad = (np.exp(logp) - onehot) * valid[:, None, :] / max(n, 1)
    return loss, grad.astype(np.float32)


# --------------------------------------------------------------------------- n-best rescoring
def attention_rescoring(hyps, beam_score, td_score, decoder_logp, r_decoder_logp, eos: int,
                        reverse_weight: float, attn_weight: float, ctc_weight: float,
                        transducer_weight: float):
    """The score-combine loop of Transducer.transducer_attention_rescoring (transducer.py:487-511):
    ``decoder_logp`` / ``r_decoder_logp`` [N, L+1, V] log-softmax outputs of the attention decoder (numpy),
    ``td_score`` = _cal_transducer_score.  Returns (best index, best score, attention part per hypothesis)."""
    best_score, best_index, attn = -float("inf"), 0, []
    for i, hyp in enumerate(hyps):
        score = 0.0
        for j, w in enumerate(hyp):
            score += float(decoder_logp[i][j][w])
        score += float(decoder_logp[i][len(hyp)][eos])
        if reverse_weight > 0:
            r_score = 0.0
            for j, w in enumerate(hyp):
                r_score += float(r_decoder_logp[i][len(hyp) - j - 1][w])
            r_score += float(r_decoder_logp[i][len(hyp)][eos])
            score = score * (1 - reverse_weight) + r_score * reverse_weight
        attn.append(score)
        score = score * attn_weight + beam_score[i] * ctc_weight + float(td_score[i]) * transducer_weight
        if score > best_score:
            best_score, best_index = score, i
    return best_index, best_score, attn
