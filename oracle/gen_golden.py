"""oracle/gen_golden.py — TEST INFRASTRUCTURE.  Generates tests/golden/*.npz.

Runs the REAL reference modules imported from /root/reference (read-only; only
available in the build container, never on the GPU box) on seeded synthetic
inputs, checks oracle/reference.py and oracle/lattice_ref.c against them, and
writes small fixtures (inputs + weights + reference outputs) that the test-suite
replays without the reference tree.

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py

Reference entry points exercised (paths under /root/reference):
  wenet/transducer/joint.py:45            TransducerJoint.forward
  wenet/transducer/predictor.py:88,180    RNNPredictor.forward / forward_step
  wenet/transducer/transducer.py:79       Transducer.forward (loss dict)
  wenet/transducer/transducer.py:277      Transducer._cal_transducer_score
  wenet/transformer/ctc.py:46             CTC.forward
  wenet/transducer/search/greedy_search.py:34          basic_greedy_search (fork)
  wenet/transducer/search/greedy_search copy.py:6      basic_greedy_search (upstream)
  wenet/transducer/search/prefix_beam_search.py:42     PrefixBeamSearch.prefix_beam_search
"""
from __future__ import annotations

import contextlib
import importlib.util
import io
import os
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")
GOLD = ROOT / "tests" / "golden"
sys.path.insert(0, str(ROOT))
sys.dont_write_bytecode = True


def import_reference():
    """context_bias.py:5 does `from turtle import forward` (needs tkinter): stub it."""
    stub = types.ModuleType("turtle")
    stub.forward = None
    sys.modules.setdefault("turtle", stub)
    sys.path.insert(0, str(REF))
    import wenet.transducer.transducer as tr  # noqa: F401
    return tr


class PassThroughEncoder(torch.nn.Module):
    """encoder stand-in: features are already encoder outputs (encoder is out of scope)."""

    def forward(self, xs, xs_lens, decoding_chunk_size=0, num_decoding_left_chunks=-1):
        T = xs.size(1)
        mask = (torch.arange(T)[None, :] < xs_lens[:, None]).unsqueeze(1)
        return xs, mask

    def output_size(self):
        return 0


def make_identity_bias():
    """context-bias stand-in (hot-word path is out of scope): features pass unchanged.
    Subclasses the reference ContextBias only to satisfy Transducer.__init__'s typeguard."""
    from wenet.transformer.context_bias import ContextBias

    class IdentityBias(ContextBias):
        def __init__(self):
            torch.nn.Module.__init__(self)

        def forward_bias_hidden(self, context_list, context_lengths):
            return None

        def forward_encoder_bias(self, bias_hidden, x):
            return x, torch.zeros_like(x)

        def forward_predictor_bias(self, bias_hidden, x):
            return x, torch.zeros_like(x)

    return IdentityBias()


def sd_np(module, prefix=""):
    return {prefix + k: v.detach().cpu().numpy() for k, v in module.state_dict().items()}


def make_batch(g, B, T, U, V, E, ragged=True):
    enc = torch.randn(B, T, E, generator=g)
    if ragged:
        tl = torch.randint(max(1, T // 2), T + 1, (B,), generator=g)
        ul = torch.randint(max(0, U // 2), U + 1, (B,), generator=g)
        tl[0] = T
        ul[-1] = U
    else:
        tl = torch.full((B,), T)
        ul = torch.full((B,), U)
    text = torch.randint(1, V, (B, U), generator=g)
    for b in range(B):
        text[b, ul[b]:] = -1
    return enc, tl.to(torch.int64), text.to(torch.int64), ul.to(torch.int64)


def main():
    tr = import_reference()
    from wenet.transducer.joint import TransducerJoint
    from wenet.transducer.predictor import RNNPredictor
    from wenet.transducer.search.greedy_search import basic_greedy_search as fork_greedy
    from wenet.transducer.search.prefix_beam_search import PrefixBeamSearch
    from wenet.transformer.ctc import CTC
    import torchaudio

    from oracle import reference as orc
    from oracle.c_oracle import rnnt_loss_c, ctc_loss_c

    spec = importlib.util.spec_from_file_location(
        "greedy_upstream", REF / "wenet/transducer/search/greedy_search copy.py")
    greedy_up = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(greedy_up)

    GOLD.mkdir(parents=True, exist_ok=True)
    V, E, P, J, H, L = 37, 16, 16, 24, 16, 2
    torch.manual_seed(20261104)
    g = torch.Generator().manual_seed(7)

    ref_joint = TransducerJoint(V, E, P, J)
    ref_pred = RNNPredictor(V, E, P, 0.1, H, L)
    ref_ctc = CTC(V, E)
    ref_model = tr.Transducer(V, 0, PassThroughEncoder(), ref_pred, ref_joint, None, ref_ctc,
                              make_identity_bias(), ctc_weight=0.3, transducer_weight=0.7,
                              attention_weight=0.0, hw_weight=0.0, loss_mode='pred')
    ref_model.eval()  # dropout off: deterministic

    om = orc.OracleModel(V, E, P, J, E, H, L, blank=0)
    om.predictor.load_state_dict(ref_pred.state_dict())
    om.joint.load_state_dict(ref_joint.state_dict())
    om.ctc.load_state_dict(ref_ctc.state_dict())
    om.eval()
    weights = {}
    weights.update(sd_np(ref_pred, "predictor."))
    weights.update(sd_np(ref_joint, "joint."))
    weights.update(sd_np(ref_ctc, "ctc."))
    dims = np.array([V, E, P, J, H, L], dtype=np.int64)

    # ------------------------------------------------------------------ (A) joint + rnnt loss
    B, T, U = 3, 11, 5
    enc, tl, text, ul = make_batch(g, B, T, U, V, E)
    enc_r = enc.clone().requires_grad_(True)
    ys_in = orc.add_blank(text, 0, -1)
    from wenet.utils.common import add_blank as ref_add_blank
    assert torch.equal(ys_in, ref_add_blank(text, 0, -1))
    pred_out = ref_pred(ys_in).detach().requires_grad_(True)
    logits = ref_joint(enc_r, pred_out)
    logits.retain_grad()
    tgt, tl32, ul32 = orc.rnnt_targets(text, ul, tl)
    costs = torchaudio.functional.rnnt_loss(logits, tgt, tl32, ul32, blank=0, reduction="none")
    loss = costs.mean()
    ref_joint.zero_grad()
    loss.backward()
    # restatement checks
    lo = orc.joint_logits(enc, pred_out.detach(), om.joint)
    assert torch.allclose(lo, logits.detach(), atol=1e-6), "joint restatement differs"
    c_costs, c_grad, _, _ = rnnt_loss_c(logits.detach(), tgt, tl32, ul32, blank=0,
                                         dy=np.full(B, 1.0 / B))
    assert np.allclose(c_costs, costs.detach().numpy(), rtol=1e-5), (c_costs, costs)
    assert np.allclose(c_grad, logits.grad.numpy(), atol=2e-6), np.abs(c_grad - logits.grad.numpy()).max()
    np.savez_compressed(
        GOLD / "rnnt_joint_small.npz", dims=dims, enc=enc.numpy(), pred_out=pred_out.detach().numpy(),
        text=text.numpy(), enc_lens=tl.numpy(), text_lens=ul.numpy(), logits=logits.detach().numpy(),
        costs=costs.detach().numpy(), loss=loss.detach().numpy(), grad_logits=logits.grad.numpy(),
        grad_enc=enc_r.grad.numpy(), grad_pred=pred_out.grad.numpy(),
        **{"w." + k: v for k, v in weights.items()},
        **{"g.joint." + n: p.grad.numpy() for n, p in ref_joint.named_parameters()})

    # fp16 logits through torchaudio (cost returned in fp16)
    lh = logits.detach().half()
    ch = torchaudio.functional.rnnt_loss(lh, tgt, tl32, ul32, blank=0, reduction="none")
    # edge cases on materialised logits: U=0, T=1, blank=-1, target == blank, clamp
    edge = {}
    x = torch.randn(2, 4, 1, 9, generator=g, requires_grad=True)  # U = 0
    c = torchaudio.functional.rnnt_loss(x, torch.zeros(2, 0, dtype=torch.int32),
                                        torch.tensor([4, 2], dtype=torch.int32),
                                        torch.tensor([0, 0], dtype=torch.int32), blank=0, reduction="none")
    c.sum().backward()
    edge.update(u0_logits=x.detach().numpy(), u0_costs=c.detach().numpy(), u0_grad=x.grad.numpy())
    x = torch.randn(2, 1, 3, 9, generator=g, requires_grad=True)  # T = 1
    t2 = torch.tensor([[3, 5], [2, 0]], dtype=torch.int32)
    c = torchaudio.functional.rnnt_loss(x, t2, torch.tensor([1, 1], dtype=torch.int32),
                                        torch.tensor([2, 1], dtype=torch.int32), blank=-1, reduction="none")
    c.sum().backward()
    edge.update(t1_logits=x.detach().numpy(), t1_targets=t2.numpy(), t1_costs=c.detach().numpy(),
                t1_grad=x.grad.numpy())
    x = torch.randn(2, 5, 4, 7, generator=g, requires_grad=True)  # a target equal to blank + clamp
    t3 = torch.tensor([[2, 0, 4], [0, 6, 0]], dtype=torch.int32)
    l3 = (torch.tensor([5, 4], dtype=torch.int32), torch.tensor([3, 2], dtype=torch.int32))
    c = torchaudio.functional.rnnt_loss(x, t3, l3[0], l3[1], blank=0, clamp=0.05, reduction="sum")
    c.backward()
    edge.update(tb_logits=x.detach().numpy(), tb_targets=t3.numpy(), tb_tlen=l3[0].numpy(),
                tb_ulen=l3[1].numpy(), tb_sum=c.detach().numpy(), tb_grad=x.grad.numpy())
    np.savez_compressed(GOLD / "rnnt_edge.npz", fp16_costs=ch.numpy(), **edge)

    # ------------------------------------------------------------------ (B) Transducer.forward dict
    enc2, tl2, text2, ul2 = make_batch(g, 4, 9, 4, V, E)
    ref_joint.zero_grad(); ref_pred.zero_grad(); ref_ctc.zero_grad()
    enc2r = enc2.clone().requires_grad_(True)
    out = ref_model(enc2r, tl2, text2, ul2)
    out["loss"].backward()
    mine_rnnt = orc.transducer_loss(om, enc2, tl2, text2, ul2)
    mine_ctc = orc.ctc_forward(enc2, tl2.to(torch.int32), text2, ul2.to(torch.int32), om.ctc)
    assert torch.allclose(mine_rnnt, out["loss_rnnt"], rtol=1e-6), (mine_rnnt, out["loss_rnnt"])
    assert torch.allclose(mine_ctc, out["loss_ctc"], rtol=1e-6), (mine_ctc, out["loss_ctc"])
    np.savez_compressed(
        GOLD / "transducer_forward.npz", dims=dims, enc=enc2.numpy(), enc_lens=tl2.numpy(),
        text=text2.numpy(), text_lens=ul2.numpy(), ctc_weight=np.float64(0.3),
        transducer_weight=np.float64(0.7), loss=out["loss"].detach().numpy(),
        loss_rnnt=out["loss_rnnt"].detach().numpy(), loss_ctc=out["loss_ctc"].detach().numpy(),
        grad_enc=enc2r.grad.numpy(),
        **{"w." + k: v for k, v in weights.items()},
        **{"g.joint." + n: p.grad.numpy() for n, p in ref_joint.named_parameters()},
        **{"g.ctc." + n: p.grad.numpy() for n, p in ref_ctc.named_parameters()},
        **{"g.predictor." + n: p.grad.numpy() for n, p in ref_pred.named_parameters()})

    # _cal_transducer_score: one utterance repeated over N hypotheses
    N = 3
    enc_rep = enc2[:1, :int(tl2[0])].repeat(N, 1, 1)
    mask = torch.ones(N, 1, enc_rep.size(1), dtype=torch.bool)
    hyps = torch.tensor([[5, 9, 3, -1], [7, 7, -1, -1], [1, 2, 3, 4]])
    hl = torch.tensor([3, 2, 4])
    with torch.no_grad():
        sc = ref_model._cal_transducer_score(enc_rep, mask, hl, hyps)
        sc_mine = orc.transducer_score(om, enc_rep, torch.full((N,), enc_rep.size(1)), hyps, hl)
    assert torch.allclose(sc, sc_mine, rtol=1e-6)
    np.savez_compressed(GOLD / "transducer_score.npz", dims=dims, enc=enc_rep.numpy(), hyps=hyps.numpy(),
                        hyps_lens=hl.numpy(), score=sc.numpy(),
                        **{"w." + k: v for k, v in weights.items()})

    # ------------------------------------------------------------------ (C) CTC
    Bc, Tc, Lc = 4, 14, 5
    hs = torch.randn(Bc, Tc, E, generator=g)
    hl_c = torch.tensor([14, 9, 12, 11], dtype=torch.int32)
    yl_c = torch.tensor([5, 2, 0, 4], dtype=torch.int32)
    ys = torch.randint(1, V, (Bc, Lc), generator=g)
    ys[0, 1] = ys[0, 0]          # repeated label: exercises l'_s == l'_{s-2}
    ys[3, 3] = ys[3, 2]
    for b in range(Bc):
        ys[b, yl_c[b]:] = -1
    ref_ctc.zero_grad()
    hsr = hs.clone().requires_grad_(True)
    lo_ctc = ref_ctc.ctc_lo(hsr)
    lo_ctc.retain_grad()
    # CTC.forward on the logits it would produce (same arithmetic as ctc.py:57-63)
    loss_ctc = ref_ctc(hsr, hl_c, ys, yl_c)
    loss_ctc.backward()
    lo2 = ref_ctc.ctc_lo(hs).detach().requires_grad_(True)
    lp2 = lo2.transpose(0, 1).log_softmax(2)
    l2 = torch.nn.functional.ctc_loss(lp2, ys, hl_c, yl_c, reduction="sum") / Bc
    l2.backward()
    nll_c, g_c = ctc_loss_c(lo2.detach(), ys, hl_c, yl_c, blank=0, scale=1.0 / Bc)
    assert np.allclose(nll_c.sum() / Bc, loss_ctc.item(), rtol=1e-6), (nll_c, loss_ctc)
    assert np.allclose(g_c, lo2.grad.numpy(), atol=2e-7), np.abs(g_c - lo2.grad.numpy()).max()
    np.savez_compressed(
        GOLD / "ctc_small.npz", dims=dims, hs=hs.numpy(), hlens=hl_c.numpy(), ys=ys.numpy(),
        ylens=yl_c.numpy(), loss=loss_ctc.detach().numpy(), logits=lo2.detach().numpy(),
        grad_logits=lo2.grad.numpy(), grad_hs=hsr.grad.numpy(), nll=nll_c,
        **{"w.ctc." + k: v for k, v in sd_np(ref_ctc).items()},
        **{"g.ctc." + n: p.grad.numpy() for n, p in ref_ctc.named_parameters()})

    # ------------------------------------------------------------------ (D) greedy, (E) beam
    # Random-init weights make the search degenerate (every frame emits n_steps copies of
    # one symbol).  "mixed" rescales a few weights so that blank and
    # non-blank steps interleave; "degenerate" keeps the raw weights and hits the cap.
    Bs, Ts = 4, 30
    encs = torch.randn(Bs, Ts, E, generator=g) * 2.0
    lens = torch.tensor([30, 22, 27, 30])
    base_sd = {"p": ref_pred.state_dict(), "j": ref_joint.state_dict()}
    base_sd = {k: {n: t.clone() for n, t in v.items()} for k, v in base_sd.items()}
    for tag, n_steps in (("mixed", 64), ("degenerate", 3)):
        ref_pred.load_state_dict(base_sd["p"])
        ref_joint.load_state_dict(base_sd["j"])
        if tag == "mixed":
            with torch.no_grad():
                ref_joint.ffn_out.weight *= 3.0
                ref_joint.ffn_out.bias[0] += 3.0
                ref_pred.projection.weight *= 16.0
                ref_pred.embed.weight *= 4.0
        om.predictor.load_state_dict(ref_pred.state_dict())
        om.joint.load_state_dict(ref_joint.state_dict())
        sw = {}
        sw.update(sd_np(ref_pred, "predictor."))
        sw.update(sd_np(ref_joint, "joint."))
        sw.update(sd_np(ref_ctc, "ctc."))
        results = {}
        hyps_all, margins = [], []
        upm = types.SimpleNamespace(blank=0, predictor=ref_pred, joint=ref_joint,
                                    context_bias=types.SimpleNamespace(
                                        forward_bias_hidden=lambda a, b_: None,
                                        forward_encoder_bias=lambda h, x: x,
                                        forward_predictor_bias=lambda h, x: x))
        for b in range(Bs):
            e1 = encs[b:b + 1, :int(lens[b])]
            with contextlib.redirect_stdout(io.StringIO()):
                fork = fork_greedy(ref_model, e1, lens[b], torch.IntTensor([[0]]), torch.IntTensor([1]),
                                   n_steps=n_steps, context_filter_state='off',
                                   context_decoder_labels_padded=torch.IntTensor([[0]]))[0][0]
            up = greedy_up.basic_greedy_search(upm, e1, lens[b], n_steps=n_steps)[0]
            mine, margin = orc.greedy_search(om, e1, int(lens[b]), n_steps=n_steps, record_margin=True)
            assert fork == up == mine, (tag, b, fork, up, mine)
            hyps_all.append(mine)
            margins.append(margin)
        width = max(1, max(len(h) for h in hyps_all))
        arr = np.full((Bs, width), -1, dtype=np.int64)
        for b, h in enumerate(hyps_all):
            arr[b, :len(h)] = h
        results.update(greedy_hyps=arr, greedy_margin=np.array(margins), n_steps=np.int64(n_steps))
        beam_size = 4
        bs = PrefixBeamSearch(PassThroughEncoder(), ref_pred, ref_joint, ref_ctc, 0)
        nb_h = np.full((Bs, beam_size, Ts + 1), -1, dtype=np.int64)
        nb_s = np.full((Bs, beam_size), -np.inf)
        nb_n = np.zeros(Bs, dtype=np.int64)
        for b in range(Bs):
            e1 = encs[b:b + 1, :int(lens[b])]
            with torch.no_grad():
                beam, _ = bs.prefix_beam_search(e1, lens[b:b + 1], beam_size=beam_size,
                                                ctc_weight=0.3, transducer_weight=0.7)
                mine_b = orc.prefix_beam_search(om, e1, beam_size=beam_size, ctc_weight=0.3,
                                                transducer_weight=0.7)
            assert [s.hyp for s in beam] == [s.hyp for s in mine_b]
            assert np.allclose([s.score for s in beam], [s.score for s in mine_b], rtol=1e-9)
            nb_n[b] = len(beam)
            for k, s in enumerate(beam):
                nb_h[b, k, :len(s.hyp) - 1] = s.hyp[1:]
                nb_s[b, k] = s.score
        results.update(beam_hyps=nb_h, beam_scores=nb_s, beam_n=nb_n, beam_size=np.int64(beam_size),
                       ctc_weight=np.float64(0.3), transducer_weight=np.float64(0.7))
        np.savez_compressed(GOLD / f"search_{tag}.npz", dims=dims, enc=encs.numpy(), lens=lens.numpy(),
                            **results, **{"w." + k: v for k, v in sw.items()})
    print("golden fixtures written to", GOLD)
    for f in sorted(GOLD.glob("*.npz")):
        print(f"  {f.name}: {f.stat().st_size} bytes")


if __name__ == "__main__":
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    main()
