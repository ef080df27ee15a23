"""GPU: model-level drop-in (patch_transducer) against the oracle chain."""
import copy
import types

import numpy as np
import pytest
import torch

from oracle import reference as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


class PassThroughEncoder(torch.nn.Module):
    def forward(self, xs, xs_lens, decoding_chunk_size=0, num_decoding_left_chunks=-1):
        T = xs.size(1)
        mask = (torch.arange(T, device=xs.device)[None, :] < xs_lens[:, None]).unsqueeze(1)
        return xs, mask


class TinyTransducer(torch.nn.Module):
    """The attributes Transducer.forward reads (transducer.py:20-77), without the encoder."""

    def __init__(self, om, ctc_weight=0.3, transducer_weight=0.7):
        super().__init__()
        self.encoder = PassThroughEncoder()
        self.predictor, self.joint, self.ctc = om.predictor, om.joint, om.ctc
        self.ctc.dropout_rate = 0.0
        self.blank, self.ignore_id = 0, -1
        self.ctc_weight, self.transducer_weight = ctc_weight, transducer_weight
        self.attention_decoder_weight, self.hw_weight = 0.0, 0.0
        self.decoder, self.context_bias = None, None


def _model(V=200, seed=0):
    torch.manual_seed(seed)
    om = orc.OracleModel(V, 32, 32, 64, 32, 32, 2, blank=0)
    om.predictor.dropout.p = 0.0   # training mode (cuDNN LSTM backward needs it) without noise
    om.predictor.rnn.dropout = 0.0
    with torch.no_grad():
        om.joint.ffn_out.weight.mul_(3.0)
    return om.train()


def test_forward_loss_dict_and_grads():
    from wenet_celoss_b200 import patch_transducer
    om = _model()
    g = torch.Generator().manual_seed(1)
    B, T, U = 4, 21, 6
    enc = torch.randn(B, T, 32, generator=g)
    tl = torch.tensor([21, 15, 18, 21])
    ul = torch.tensor([6, 3, 0, 5])
    text = torch.randint(1, 200, (B, U), generator=g)
    for b in range(B):
        text[b, ul[b]:] = -1
    # oracle chain (transducer.py:124-167 restated)
    ref = copy.deepcopy(om)
    er = enc.clone().requires_grad_(True)
    l_rnnt = orc.transducer_loss(ref, er, tl, text, ul)
    l_ctc = orc.ctc_forward(er, tl.to(torch.int32), text, ul.to(torch.int32), ref.ctc)
    (0.7 * l_rnnt + 0.3 * l_ctc).backward()
    # patched model on the GPU
    tm = patch_transducer(TinyTransducer(copy.deepcopy(om)).to(DEV))
    ed = enc.to(DEV).requires_grad_(True)
    out = tm(ed, tl.to(DEV), text.to(DEV), ul.to(DEV))
    assert set(out) == {"loss", "loss_att", "loss_ctc", "loss_rnnt", "hw_loss"}
    assert out["loss_att"] is None and out["hw_loss"] is None
    np.testing.assert_allclose(out["loss_rnnt"].item(), l_rnnt.item(), rtol=1e-2)
    np.testing.assert_allclose(out["loss_ctc"].item(), l_ctc.item(), rtol=1e-4)
    np.testing.assert_allclose(out["loss"].item(), (0.7 * l_rnnt + 0.3 * l_ctc).item(), rtol=1e-2)
    out["loss"].backward()

    def close(name, a, b):
        a, b = a.detach().cpu().double(), b.double()
        rel = (a - b).norm() / b.norm().clamp_min(1e-30)
        assert rel < 1e-2, f"{name}: rel_l2 {rel:.3e}"
    close("enc", ed.grad, er.grad)
    refp = dict(ref.named_parameters())
    for n, p in list(tm.predictor.named_parameters()) + list(tm.joint.named_parameters()) + \
            list(tm.ctc.named_parameters()):
        pre = "predictor." if any(p is q for q in tm.predictor.parameters()) else \
            ("joint." if any(p is q for q in tm.joint.parameters()) else "ctc.")
        close(pre + n, p.grad, refp[pre + n].grad)


def test_transducer_score_and_decoders():
    from wenet_celoss_b200 import patch_transducer
    om = _model(seed=3)
    with torch.no_grad():
        om.joint.ffn_out.bias[0] += 3.0
        om.predictor.projection.weight.mul_(16.0)
        om.predictor.embed.weight.mul_(4.0)
    tm = patch_transducer(TinyTransducer(copy.deepcopy(om)).to(DEV))
    g = torch.Generator().manual_seed(5)
    T = 18
    enc1 = torch.randn(1, T, 32, generator=g) * 2
    hyps = torch.tensor([[5, 9, 3, -1], [7, 7, -1, -1], [1, 2, 3, 4]])
    hl = torch.tensor([3, 2, 4])
    enc_rep = enc1.repeat(3, 1, 1)
    want = orc.transducer_score(om, enc_rep, torch.full((3,), T), hyps, hl)
    mask = torch.ones(3, 1, T, dtype=torch.bool, device=DEV)
    got = tm._cal_transducer_score(enc_rep.to(DEV), mask, hl.to(DEV), hyps.to(DEV))
    np.testing.assert_allclose(got.cpu().numpy(), want.detach().numpy(), rtol=1e-2)
    # decoders: batch of 2 through the model-level entry points
    enc = torch.randn(2, T, 32, generator=g) * 2
    lens = torch.tensor([T, 11])
    # (hyps, dist) as transducer.py:598; no ContextBias on this model -> upstream decode, dist 0
    hyps_g, dist = tm.greedy_search(enc.to(DEV), lens.to(DEV), -1, -1, False, 4)
    assert dist == 0
    for b in range(2):
        assert hyps_g[b] == orc.greedy_search(om, enc[b:b + 1, :lens[b]], int(lens[b]), n_steps=4)
    best = tm.beam_search(enc[:1].to(DEV), lens[:1].to(DEV), beam_size=5)
    wh, ws = orc.beam_search_best(om, enc[:1], beam_size=5)
    assert best[0] == wh
    np.testing.assert_allclose(best[1], ws, rtol=1e-5)
