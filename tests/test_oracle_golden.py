"""CPU: the oracle (oracle/reference.py, oracle/lattice_ref.c) against the golden
fixtures that oracle/gen_golden.py produced by running the real reference modules."""
import numpy as np
import pytest
import torch
import torchaudio

from conftest import load_oracle_model
from oracle import reference as orc
from oracle.c_oracle import ctc_loss_c, rnnt_loss_c


def test_joint_and_rnnt_loss_match_reference(golden):
    d = golden("rnnt_joint_small")
    m = load_oracle_model(d)
    enc = torch.from_numpy(d["enc"]).requires_grad_(True)
    pred = torch.from_numpy(d["pred_out"]).requires_grad_(True)
    logits = orc.joint_logits(enc, pred, m.joint)
    np.testing.assert_allclose(logits.detach().numpy(), d["logits"], atol=1e-5)
    tgt, tl, ul = orc.rnnt_targets(torch.from_numpy(d["text"]), torch.from_numpy(d["text_lens"]),
                                   torch.from_numpy(d["enc_lens"]))
    costs = torchaudio.functional.rnnt_loss(logits, tgt, tl, ul, blank=0, reduction="none")
    np.testing.assert_allclose(costs.detach().numpy(), d["costs"], rtol=1e-5)
    costs.mean().backward()
    np.testing.assert_allclose(enc.grad.numpy(), d["grad_enc"], atol=2e-6)
    np.testing.assert_allclose(pred.grad.numpy(), d["grad_pred"], atol=2e-6)
    for n, p in m.joint.named_parameters():
        np.testing.assert_allclose(p.grad.numpy(), d["g.joint." + n], atol=5e-6, err_msg=n)


def test_c_lattice_matches_reference_rnnt(golden):
    d = golden("rnnt_joint_small")
    B = d["logits"].shape[0]
    tgt, tl, ul = orc.rnnt_targets(torch.from_numpy(d["text"]), torch.from_numpy(d["text_lens"]),
                                   torch.from_numpy(d["enc_lens"]))
    costs, grad, alpha, beta = rnnt_loss_c(d["logits"], tgt, tl, ul, blank=0, dy=np.full(B, 1.0 / B))
    np.testing.assert_allclose(costs, d["costs"], rtol=1e-5)
    np.testing.assert_allclose(grad, d["grad_logits"], atol=2e-6)
    # properties: padded cells exactly zero, each cell's gradient sums to ~0
    for b in range(B):
        assert np.all(grad[b, tl[b]:] == 0) and np.all(grad[b, :, ul[b] + 1:] == 0)
    assert np.abs(grad.sum(-1)).max() < 1e-6


def test_c_lattice_edge_cases(golden):
    d = golden("rnnt_edge")
    c, g, _, _ = rnnt_loss_c(d["u0_logits"], np.zeros((2, 0), np.int32), [4, 2], [0, 0], blank=0)
    np.testing.assert_allclose(c, d["u0_costs"], rtol=1e-5)
    np.testing.assert_allclose(g, d["u0_grad"], atol=2e-6)
    # U = 0: cost is minus the sum of blank log-probs
    x = torch.from_numpy(d["u0_logits"])
    lp = x.log_softmax(-1)[:, :, 0, 0]
    np.testing.assert_allclose(c, [-lp[0, :4].sum(), -lp[1, :2].sum()], rtol=1e-5)
    V = d["t1_logits"].shape[-1]
    c, g, _, _ = rnnt_loss_c(d["t1_logits"], d["t1_targets"], [1, 1], [2, 1], blank=V - 1)
    np.testing.assert_allclose(c, d["t1_costs"], rtol=1e-5)
    np.testing.assert_allclose(g, d["t1_grad"], atol=2e-6)
    c, g, _, _ = rnnt_loss_c(d["tb_logits"], d["tb_targets"], d["tb_tlen"], d["tb_ulen"], blank=0)
    np.testing.assert_allclose(c.sum(), d["tb_sum"], rtol=1e-5)
    np.testing.assert_allclose(np.clip(g, -0.05, 0.05), d["tb_grad"], atol=2e-6)


def test_transducer_forward_dict(golden):
    d = golden("transducer_forward")
    m = load_oracle_model(d)
    enc = torch.from_numpy(d["enc"]).requires_grad_(True)
    tl, text, ul = (torch.from_numpy(d[k]) for k in ("enc_lens", "text", "text_lens"))
    l_rnnt = orc.transducer_loss(m, enc, tl, text, ul)
    l_ctc = orc.ctc_forward(enc, tl.to(torch.int32), text, ul.to(torch.int32), m.ctc)
    loss = float(d["transducer_weight"]) * l_rnnt + float(d["ctc_weight"]) * l_ctc
    np.testing.assert_allclose(l_rnnt.item(), d["loss_rnnt"], rtol=1e-6)
    np.testing.assert_allclose(l_ctc.item(), d["loss_ctc"], rtol=1e-6)
    np.testing.assert_allclose(loss.item(), d["loss"], rtol=1e-6)
    loss.backward()
    np.testing.assert_allclose(enc.grad.numpy(), d["grad_enc"], atol=2e-6)
    for n, p in m.predictor.named_parameters():
        np.testing.assert_allclose(p.grad.numpy(), d["g.predictor." + n], atol=5e-6, err_msg=n)


def test_transducer_score(golden):
    d = golden("transducer_score")
    m = load_oracle_model(d)
    enc = torch.from_numpy(d["enc"])
    with torch.no_grad():
        sc = orc.transducer_score(m, enc, torch.full((enc.size(0),), enc.size(1)),
                                  torch.from_numpy(d["hyps"]), torch.from_numpy(d["hyps_lens"]))
    np.testing.assert_allclose(sc.numpy(), d["score"], rtol=1e-6)


def test_ctc_matches_reference(golden):
    d = golden("ctc_small")
    B = d["hs"].shape[0]
    nll, grad = ctc_loss_c(d["logits"], d["ys"], d["hlens"], d["ylens"], blank=0, scale=1.0 / B)
    np.testing.assert_allclose(nll.sum() / B, d["loss"], rtol=1e-6)
    np.testing.assert_allclose(grad, d["grad_logits"], rtol=1e-5, atol=2e-6)
    # L = 0 utterance: nll = -sum_t lp[t, blank]
    b = int(np.where(d["ylens"] == 0)[0][0])
    lp = torch.from_numpy(d["logits"][b]).log_softmax(-1)
    np.testing.assert_allclose(nll[b], -lp[:d["hlens"][b], 0].sum().item(), rtol=1e-6)
    assert np.abs(grad.sum(-1)).max() < 1e-6


def _unpad(row):
    return [int(x) for x in row if x >= 0]


def test_greedy_matches_reference(golden):
    for tag in ("mixed", "degenerate"):
        d = golden(f"search_{tag}")
        m = load_oracle_model(d)
        enc = torch.from_numpy(d["enc"])
        for b, n in enumerate(d["lens"]):
            hyp = orc.greedy_search(m, enc[b:b + 1, :n], int(n), n_steps=int(d["n_steps"]))
            assert hyp == _unpad(d["greedy_hyps"][b]), (tag, b)


def test_beam_matches_reference(golden):
    for tag in ("mixed", "degenerate"):
        d = golden(f"search_{tag}")
        m = load_oracle_model(d)
        enc = torch.from_numpy(d["enc"])
        for b, n in enumerate(d["lens"]):
            beam = orc.prefix_beam_search(m, enc[b:b + 1, :n], beam_size=int(d["beam_size"]),
                                          ctc_weight=float(d["ctc_weight"]),
                                          transducer_weight=float(d["transducer_weight"]))
            assert len(beam) == d["beam_n"][b]
            for k, h in enumerate(beam):
                assert h.hyp[1:] == _unpad(d["beam_hyps"][b, k]), (tag, b, k)
                np.testing.assert_allclose(h.score, d["beam_scores"][b, k], rtol=1e-6)


def test_log_add_and_add_blank():
    import math
    assert orc.log_add([-math.inf, -math.inf]) == -math.inf
    assert abs(orc.log_add([0.0, 0.0]) - math.log(2)) < 1e-15
    ys = torch.tensor([[1, 2, 3], [4, -1, -1]])
    assert orc.add_blank(ys, 0, -1).tolist() == [[0, 1, 2, 3], [0, 4, 0, 0]]


# ---- CTC prefix beam search ---------------------------------------------------------------
# Known-answer test of the reference itself: runtime/core/test/ctc_prefix_beam_search_test.cc:30-61
# (3 frames x 3 classes, beam 3; prefix scores 0.2185 / 0.1550 / 0.1525).
KAT_PROBS = [[0.25, 0.40, 0.35], [0.40, 0.35, 0.25], [0.10, 0.50, 0.40]]
KAT_HYPS = [(2, 1), (1, 2), (1,)]
KAT_SCORES = [0.2185, 0.1550, 0.1525]


def test_ctc_prefix_beam_search_reference_known_answer():
    hyps = orc.ctc_prefix_beam_search(torch.log(torch.tensor(KAT_PROBS)), 3)
    assert [h[0] for h in hyps] == KAT_HYPS
    np.testing.assert_allclose(np.exp([h[1] for h in hyps]), KAT_SCORES, rtol=1e-6)


def test_ctc_prefix_beam_search_matches_reference(golden):
    d = golden("ctc_search")
    for name in d["names"]:
        logits = torch.from_numpy(d[f"{name}.logits"])
        beam = int(d[f"{name}.beam"])
        hyps = orc.ctc_prefix_beam_search(torch.log_softmax(logits.unsqueeze(0), dim=2)[0], beam)
        assert len(hyps) == int(d[f"{name}.n"])
        for k, (prefix, score) in enumerate(hyps):
            assert list(prefix) == d[f"{name}.tokens"][k][:d[f"{name}.lens"][k]].tolist(), (name, k)
            np.testing.assert_allclose(score, d[f"{name}.scores"][k], rtol=1e-12)


def test_forced_align_matches_reference(golden):
    d = golden("ctc_search")
    for name in d["align_names"]:
        got = orc.forced_align(torch.from_numpy(d[f"align.{name}.logp"]),
                               torch.from_numpy(d[f"align.{name}.y"]), 0)
        assert got == d[f"align.{name}.alignment"].tolist(), name


def test_hotword_greedy_matches_reference(golden):
    """fork greedy search with a real ContextBias, context_filter_state='off'
    (greedy_search.py:34-176): the oracle's restatement of forward_predictor_bias inside the loop."""
    from conftest import load_hotword_model
    d = golden("hotword_greedy")
    m = load_hotword_model(d)
    cb = m.context_bias
    for b in range(d["enc"].shape[0]):
        n = int(d["lens"][b])
        got = orc.greedy_search_hotword(m, cb.enc_biased[b:b + 1, :n], n, cb.bias_hidden_empty,
                                        n_steps=int(d["n_steps"]))
        assert got == [int(x) for x in d["hyps"][b] if x >= 0], b
        # basic_greedy_search_both (loss_mode 'both', greedy_search.py:297-430): given phrase list
        got = orc.greedy_search_hotword(m, cb.enc_biased[b:b + 1, :n], n, cb.bias_hidden,
                                        n_steps=int(d["n_steps"]))
        assert got == [int(x) for x in d["hyps_both"][b] if x >= 0], ("both", b)


@pytest.mark.parametrize("mode", ["pred", "both"])
def test_hotword_gated_greedy_matches_reference(golden, mode):
    """the fork's DEFAULT greedy (context_filter_state='on', greedy_search.py:88-165 for loss_mode
    'pred', :339-407 for 'both'): gate decisions, go-back, token lists and edit distance recorded
    from the real basic_greedy_search / basic_greedy_search_both."""
    from conftest import load_hotword_model
    d = golden("hotword_greedy_gated" if mode == "pred" else "hotword_greedy_gated_both")
    m = load_hotword_model(d)
    cb = m.context_bias
    assert int(d["go_backs"]) >= 3
    labels = d["labels"].reshape(-1).tolist()
    for b in range(d["enc"].shape[0]):
        n = int(d["lens"][b])
        hyp, res = orc.greedy_search_hotword_gated(m, cb.enc_biased[b:b + 1, :n], cb.enc_empty[b:b + 1, :n], n,
                                                   cb.bias_hidden, cb.bias_hidden_empty,
                                                   n_steps=int(d["n_steps"]),
                                                   enc_bias=None if mode == "pred" else cb.enc_bias[b:b + 1, :n])
        assert hyp == [int(x) for x in d["hyps"][b] if x >= 0], b
        assert res == [int(x) for x in d["decisions"][b] if x >= 0], b
        assert orc.edit_distance(labels, res) == float(d["dist"][b])


def test_edit_distance_host():
    """product-side edit distance (host bookkeeping) against the oracle's restatement of
    greedy_search.py:6-32."""
    import random
    from wenet_celoss_b200.search import edit_distance
    rnd = random.Random(5)
    for _ in range(50):
        a = [rnd.randint(0, 1) for _ in range(rnd.randint(0, 12))]
        b = [rnd.randint(0, 1) for _ in range(rnd.randint(0, 12))]
        assert edit_distance(a, b) == orc.edit_distance(a, b)


def test_hw_ce_loss_matches_reference(golden):
    """what the real Transducer.forward handed to self.hw_criterion (transducer.py:170-200), three
    loss modes: the numpy fp64 restatement reproduces loss and gradient."""
    d = golden("hw_loss")
    for mode in d["modes"]:
        loss, grad = orc.hw_ce_loss(d[f"{mode}.hw_output"], d[f"{mode}.labels"])
        np.testing.assert_allclose(loss, float(d[f"{mode}.hw_loss"]), rtol=1e-6)
        np.testing.assert_allclose(grad, d[f"{mode}.grad"], atol=1e-7)


def test_attention_rescoring_oracle_vs_reference(golden):
    """oracle.reference.attention_rescoring reproduces what the real transducer_attention_rescoring
    (transducer.py:487-511) chose and scored (oracle/gen_golden_rescore.py)."""
    d = golden("rescore")
    eos = int(d["sos_eos"][1])
    for ci in range(int(d["ncases"])):
        pre = f"c{ci}."
        beam, rw, cw, aw, tw, _, _ = (float(x) for x in d[pre + "params"])
        hyps = [row[:n] for row, n in zip(d[pre + "hyps_pad"].tolist(), d[pre + "hyps_lens"].tolist())]
        dlp = torch.log_softmax(torch.from_numpy(d[pre + "dec"]), -1).numpy()
        rlp = torch.log_softmax(torch.from_numpy(d[pre + "rdec"]), -1).numpy() if pre + "rdec" in d.files else None
        bi, bs, attn = orc.attention_rescoring(hyps, d[pre + "beam_score"].tolist(), d[pre + "td_score"].tolist(),
                                               dlp, rlp, eos, rw, aw, cw, tw)
        assert bi == int(d[pre + "best_index"]) and hyps[bi] == d[pre + "best_hyp"].tolist()
        np.testing.assert_allclose(bs, float(d[pre + "best_score"]), rtol=1e-5)
        np.testing.assert_allclose(attn, d[pre + "attn"], rtol=1e-12)
        om = load_oracle_model(d)
        T = d[pre + "speech"].shape[1]
        N = len(hyps)
        td = orc.transducer_score(om, torch.from_numpy(d[pre + "speech"]).repeat(N, 1, 1), torch.full((N,), T),
                                  torch.from_numpy(d[pre + "hyps_pad"]), torch.from_numpy(d[pre + "hyps_lens"]))
        np.testing.assert_allclose(td.detach().numpy(), d[pre + "td_score"], rtol=1e-5, atol=1e-5)
