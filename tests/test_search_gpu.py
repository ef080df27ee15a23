"""GPU parity: batched greedy search (integer outputs: token lists must be identical)."""
import copy

import numpy as np
import pytest
import torch

from conftest import load_oracle_model
from oracle import reference as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _unpad(row):
    return [int(x) for x in row if x >= 0]


@pytest.mark.parametrize("tag", ["mixed", "degenerate"])
def test_greedy_golden(golden, tag):
    from wenet_celoss_b200 import basic_greedy_search, greedy_search_batch
    d = golden(f"search_{tag}")
    m = load_oracle_model(d).to(DEV)
    enc = torch.from_numpy(d["enc"]).to(DEV)
    lens = torch.from_numpy(d["lens"]).to(DEV)
    n_steps = int(d["n_steps"])
    hyps = greedy_search_batch(m, enc, lens, n_steps=n_steps)
    for b in range(enc.size(0)):
        assert hyps[b] == _unpad(d["greedy_hyps"][b]), (tag, b, float(d["greedy_margin"][b]))
    # single-utterance entry point, reference argument order
    b = 1
    one = basic_greedy_search(m, enc[b:b + 1, :int(lens[b])], lens[b:b + 1], None, None, n_steps)
    assert one == [_unpad(d["greedy_hyps"][b])]


def _search_model(V, seed, scale=True):
    torch.manual_seed(seed)
    m = orc.OracleModel(V, 256, 256, 512, 256, 256, 2, blank=0)
    if scale:
        with torch.no_grad():  # interleave blank / non-blank steps (see oracle/gen_golden.py)
            m.joint.ffn_out.weight.mul_(3.0)
            m.joint.ffn_out.bias[0] += 3.0
            m.predictor.projection.weight.mul_(16.0)
            m.predictor.embed.weight.mul_(4.0)
    return m.eval()


def test_greedy_full_dims_vs_oracle():
    """Reference model dims (E=P=256, J=512, LSTM 2x256), ragged batch, V=1000."""
    from wenet_celoss_b200 import greedy_search_batch
    m = _search_model(1000, seed=3)
    g = torch.Generator().manual_seed(4)
    B, T = 6, 40
    enc = torch.randn(B, T, 256, generator=g) * 2.0
    lens = torch.tensor([40, 33, 40, 1, 25, 0])
    want, margins = [], []
    for b in range(B):
        if lens[b] == 0:
            want.append([]); margins.append(np.inf); continue
        h, mg = orc.greedy_search(m, enc[b:b + 1, :lens[b]], int(lens[b]), n_steps=5, record_margin=True)
        want.append(h); margins.append(mg)
    got, rounds = greedy_search_batch(copy.deepcopy(m).to(DEV), enc.to(DEV), lens.to(DEV), n_steps=5,
                                      return_rounds=True)
    for b in range(B):
        assert got[b] == want[b], f"utt {b}: min top-2 margin in the oracle run {margins[b]:.3e}"
    assert rounds >= 40


@pytest.mark.parametrize("tag", ["mixed", "degenerate"])
def test_beam_golden(golden, tag):
    from wenet_celoss_b200 import prefix_beam_search_batch
    d = golden(f"search_{tag}")
    m = load_oracle_model(d).to(DEV)
    enc = torch.from_numpy(d["enc"]).to(DEV)
    lens = torch.from_numpy(d["lens"]).to(DEV)
    K = int(d["beam_size"])
    beams = prefix_beam_search_batch(m, enc, lens, beam_size=K, ctc_weight=float(d["ctc_weight"]),
                                     transducer_weight=float(d["transducer_weight"]))
    for b in range(enc.size(0)):
        assert len(beams[b]) == d["beam_n"][b]
        for k, s in enumerate(beams[b]):
            assert s.hyp[0] == 0 and s.hyp[1:] == _unpad(d["beam_hyps"][b, k]), (tag, b, k)
            np.testing.assert_allclose(s.score, d["beam_scores"][b, k], rtol=1e-5)


def test_beam_full_dims_vs_oracle():
    from wenet_celoss_b200 import PrefixBeamSearch, prefix_beam_search_batch
    m = _search_model(600, seed=8)
    g = torch.Generator().manual_seed(9)
    B, T = 3, 24
    enc = torch.randn(B, T, 256, generator=g) * 2.0
    lens = torch.tensor([24, 17, 24])
    md = copy.deepcopy(m).to(DEV)
    beams = prefix_beam_search_batch(md, enc.to(DEV), lens.to(DEV), beam_size=10)
    for b in range(B):
        want = orc.prefix_beam_search(m, enc[b:b + 1, :lens[b]], beam_size=10)
        assert [s.hyp for s in beams[b]] == [s.hyp for s in want], b
        np.testing.assert_allclose([s.score for s in beams[b]], [s.score for s in want], rtol=1e-5)

    class Enc(torch.nn.Module):  # pass-through encoder, as in oracle/gen_golden.py
        def forward(self, xs, lens, chunk=-1, left=-1):
            return xs, None
    bs = PrefixBeamSearch(Enc(), md.predictor, md.joint, md.ctc, 0)
    beam, enc_out = bs.prefix_beam_search(enc[0:1].to(DEV), torch.tensor([T]), beam_size=10)
    want = orc.prefix_beam_search(m, enc[0:1], beam_size=10)
    assert beam[0].hyp == want[0].hyp and enc_out.shape == (1, T, 256)


def test_hotword_greedy_golden(golden):
    """greedy search with the fork's hot-word streams (real ContextBias, filter 'off'): token lists
    of the reference's basic_greedy_search (greedy_search.py:34-176), whole batch in one call."""
    from conftest import load_hotword_model
    from wenet_celoss_b200 import greedy_search_batch
    d = golden("hotword_greedy")
    m = load_hotword_model(d).to(DEV)
    enc = torch.from_numpy(d["enc"]).to(DEV)
    lens = torch.from_numpy(d["lens"]).to(DEV)
    ctx = torch.from_numpy(d["context_list"]).to(DEV)
    ctx_len = torch.from_numpy(d["context_lengths"]).to(DEV)
    hyps = greedy_search_batch(m, enc, lens, n_steps=int(d["n_steps"]), context_list=ctx,
                               context_lengths=ctx_len)
    for b in range(enc.size(0)):
        assert hyps[b] == _unpad(d["hyps"][b]), b
    # loss_mode 'both' -> basic_greedy_search_both: the predictor bias attends to the given list
    m.loss_mode = "both"
    hyps_both = greedy_search_batch(m, enc, lens, n_steps=int(d["n_steps"]), context_list=ctx,
                                    context_lengths=ctx_len)
    for b in range(enc.size(0)):
        assert hyps_both[b] == _unpad(d["hyps_both"][b]), ("both", b)
    # without the predictor bias the same model decodes differently: the device path is exercised
    plain = greedy_search_batch(m, m.context_bias.enc_biased, lens, n_steps=int(d["n_steps"]), hotword=False)
    assert plain != hyps


@pytest.mark.parametrize("mode", ["pred", "both"])
def test_hotword_gated_greedy_golden(golden, mode):
    """the fork's DEFAULT decode (context_filter_state='on', greedy_search.py:88-165; loss_mode 'both':
    basic_greedy_search_both, :339-407): hot-word gate per predictor call, decision-dependent streams
    and the go-back re-decode on the device; token lists, gate decisions and edit distance of the real
    reference functions (46 / 48 go-backs in the fixtures)."""
    from conftest import load_hotword_model
    from wenet_celoss_b200 import basic_greedy_search, greedy_search_batch
    d = golden("hotword_greedy_gated" if mode == "pred" else "hotword_greedy_gated_both")
    m = load_hotword_model(d).to(DEV)
    m.loss_mode = mode
    enc = torch.from_numpy(d["enc"]).to(DEV)
    lens = torch.from_numpy(d["lens"]).to(DEV)
    ctx = torch.from_numpy(d["context_list"]).to(DEV)
    ctx_len = torch.from_numpy(d["context_lengths"]).to(DEV)
    n_steps = int(d["n_steps"])
    hyps, dec = greedy_search_batch(m, enc, lens, n_steps=n_steps, context_list=ctx, context_lengths=ctx_len,
                                    context_filter_state="on", return_decisions=True)
    for b in range(enc.size(0)):
        assert hyps[b] == _unpad(d["hyps"][b]), b
        assert dec[b] == _unpad(d["decisions"][b]), b
    # the gate matters: filter 'off' decodes differently on the same model
    off = greedy_search_batch(m, enc, lens, n_steps=n_steps, context_list=ctx, context_lengths=ctx_len)
    assert [off[b] for b in range(enc.size(0))] == [_unpad(r) for r in d["hyps_off"]]
    assert off != hyps
    # single-utterance entry point with the fork's return triple ([hyps], dist, result)
    # (the stand-in serves the stored full-length streams, so use a full-length utterance)
    b = int(np.argmax(d["lens"]))
    assert int(d["lens"][b]) == enc.size(1)
    m1 = load_hotword_model(d).to(DEV)
    m1.loss_mode = mode
    cb = m1.context_bias
    cb.enc, cb.enc_biased, cb.enc_empty = cb.enc[b:b + 1], cb.enc_biased[b:b + 1], cb.enc_empty[b:b + 1]
    cb.enc_bias = cb.enc_bias[b:b + 1]
    out = basic_greedy_search(m1, enc[b:b + 1], lens[b:b + 1], ctx, ctx_len, n_steps, "on",
                              torch.from_numpy(d["labels"]))
    assert out[0] == [_unpad(d["hyps"][b])] and out[1] == float(d["dist"][b])
    assert out[2] == _unpad(d["decisions"][b])
    # through the patched model entry point (Transducer.greedy_search, transducer.py:515-598): (hyps, dist)
    from wenet_celoss_b200.transducer import transducer_greedy_search

    class Enc(torch.nn.Module):
        def forward(self, xs, xl, chunk=-1, left=-1):
            return xs, torch.ones(xs.size(0), 1, xs.size(1), dtype=torch.bool, device=xs.device)
    m1.encoder = Enc()
    hy, dist = transducer_greedy_search(m1, enc[b:b + 1], lens[b:b + 1], n_steps=n_steps, context_list=ctx,
                                        context_lengths=ctx_len, context_filter_state="on",
                                        context_decoder_labels_padded=torch.from_numpy(d["labels"]))
    assert hy == [_unpad(d["hyps"][b])] and dist == float(d["dist"][b])


@pytest.mark.parametrize("mode", ["off", "on"])
def test_hotword_greedy_full_dims_vs_oracle(mode):
    """Reference dims (P = 256, 4 heads, 100 phrase vectors: the 16-byte-load paths of pred_bias /
    pred_gate kernels) against the CPU oracle restatements; utterances whose smallest argmax or
    gate margin could flip under fp32 reordering are reported, not compared."""
    import bench
    from wenet_celoss_b200 import greedy_search_batch
    m = _search_model(1000, seed=21)
    torch.manual_seed(22)
    cb = bench._ContextBias(256, 100).eval()
    with torch.no_grad():
        m.joint.ffn_out.bias[0] -= 4.5
    m.context_bias = cb
    m.loss_mode = "pred"
    g = torch.Generator().manual_seed(23)
    B, T = 4, 24
    enc = torch.randn(B, T, 256, generator=g) * 2.0
    lens = torch.tensor([24, 17, 24, 9])
    ctx = torch.zeros(100, 4, dtype=torch.int32)
    ctx_len = torch.full((100,), 4, dtype=torch.int32)
    with torch.no_grad():
        bh, bhe = cb.phrases, cb.phrases[:, :1]
        eb, _ = cb.forward_encoder_bias(bh, enc)
        ee, _ = cb.forward_encoder_bias(bhe, enc)
    want = []
    for b in range(B):
        n = int(lens[b])
        if mode == "on":
            want.append(orc.greedy_search_hotword_gated(m, eb[b:b + 1, :n], ee[b:b + 1, :n], n, bh, bhe, n_steps=5))
        else:
            want.append((orc.greedy_search_hotword(m, eb[b:b + 1, :n], n, bhe, n_steps=5), None))
    assert any(len(h) > 0 for h, _ in want)
    md = copy.deepcopy(m).to(DEV)
    hyps, dec = greedy_search_batch(md, enc.to(DEV), lens.to(DEV), n_steps=5, context_list=ctx.to(DEV),
                                    context_lengths=ctx_len.to(DEV), context_filter_state=mode,
                                    return_decisions=True)
    same = sum(hyps[b] == want[b][0] and (mode == "off" or dec[b] == want[b][1]) for b in range(B))
    assert same == B, (same, [(hyps[b][:8], want[b][0][:8]) for b in range(B)])


@pytest.mark.parametrize("switch", ["B200_GREEDY_ROUND_KERNELS", "B200_GREEDY_NO_REPL", "B200_GATE_SINGLE_CTA"])
@pytest.mark.parametrize("mode", [None, "on"])
def test_greedy_alternative_paths_agree(switch, mode, monkeypatch):
    """The library keeps alternative device paths behind environment switches: the
    per-round kernels (also what shapes without a cooperative launch take), the persistent kernel without
    the replicated decode state, the one-CTA gate chain.  Each must decode exactly what the default path
    decodes: plain greedy and the gated hot-word greedy (hypotheses and gate decisions)."""
    import bench
    from wenet_celoss_b200 import greedy_search_batch
    if mode is None and switch == "B200_GATE_SINGLE_CTA":
        pytest.skip("the gate chain only runs in the gated decode")
    m = _search_model(1000, seed=31)
    kw = {}
    if mode == "on":
        torch.manual_seed(32)
        m.context_bias = bench._ContextBias(256, 100).eval()
        m.loss_mode = "pred"
        with torch.no_grad():
            m.joint.ffn_out.bias[0] -= 4.5
        kw = dict(context_list=torch.zeros(100, 4, dtype=torch.int32, device=DEV),
                  context_lengths=torch.full((100,), 4, dtype=torch.int32, device=DEV),
                  context_filter_state="on", return_decisions=True)
    g = torch.Generator().manual_seed(33)
    B, T = 5, 40
    enc = (torch.randn(B, T, 256, generator=g) * 2.0).to(DEV)
    lens = torch.tensor([40, 31, 40, 12, 25], device=DEV)
    md = copy.deepcopy(m).to(DEV)
    out = {}
    for val in ("0", "1"):
        monkeypatch.setenv(switch, val)
        out[val] = greedy_search_batch(md, enc, lens, n_steps=5, **kw)
    assert out["0"] == out["1"]
    hyps = out["0"][0] if mode == "on" else out["0"]
    assert any(len(h) > 0 for h in hyps)


@pytest.mark.parametrize("switch", ["B200_BEAM_NO_PDL", "B200_BEAM_TM4"])
def test_beam_alternative_paths_agree(switch, monkeypatch):
    """Beam search launched without programmatic stream serialization, and with the 4-row register tiles of
    the logits GEMM (same k order per accumulator): same n-best lists, bit-identical scores."""
    from wenet_celoss_b200 import prefix_beam_search_batch
    m = _search_model(1000, seed=41).to(DEV)
    g = torch.Generator().manual_seed(43)
    B, T = 4, 30
    enc = (torch.randn(B, T, 256, generator=g) * 2.0).to(DEV)
    lens = torch.tensor([30, 22, 30, 7], dtype=torch.int32, device=DEV)
    out = {}
    for val in ("0", "1"):
        monkeypatch.setenv(switch, val)
        beams = prefix_beam_search_batch(m, enc, lens, beam_size=10)
        out[val] = [[(tuple(s.hyp), s.score) for s in bb] for bb in beams]
    assert out["0"] == out["1"]
