"""GPU: n-best rescoring against fixtures recorded from the REAL
``Transducer.transducer_attention_rescoring`` (wenet/transducer/transducer.py:378-513;
oracle/gen_golden_rescore.py).  The attention decoder is a model-side library module: a stand-in replays
the logits the real (Bi)TransformerDecoder produced for the fixture's n-best; everything else - the n-best
search, the transducer score with one shared encoder projection, the score combine - runs on the device
kernels through the C ABI."""
import copy
import ctypes

import numpy as np
import pytest
import torch

from conftest import load_oracle_model
from oracle import reference as orc
from test_transducer_gpu import PassThroughEncoder, TinyTransducer

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


class ReplayDecoder(torch.nn.Module):
    """Returns the recorded decoder logits; checks that it is asked about the recorded n-best."""

    def __init__(self, d, pre):
        super().__init__()
        self.dec = torch.from_numpy(d[pre + "dec"])
        self.rdec = torch.from_numpy(d[pre + "rdec"]) if pre + "rdec" in d.files else None
        self.ys_in = torch.from_numpy(d[pre + "ys_in"])
        self.r_ys_in = torch.from_numpy(d[pre + "r_ys_in"])
        self.right_decoder = True

    def forward(self, memory, memory_mask, ys_in_pad, ys_in_lens, r_ys_in_pad, reverse_weight):
        assert memory.shape[0] == self.dec.shape[0] and memory_mask.all()
        assert torch.equal(ys_in_pad.cpu(), self.ys_in), "n-best differs from the reference's"
        if reverse_weight > 0:
            assert torch.equal(r_ys_in_pad.cpu(), self.r_ys_in)
        r = self.rdec.to(memory.device) if self.rdec is not None else torch.tensor(0.0)
        return self.dec.to(memory.device), r, None


@pytest.mark.parametrize("ci", [0, 1, 2, 3])
def test_attention_rescoring_vs_reference(golden, ci):
    from wenet_celoss_b200 import patch_transducer
    d = golden("rescore")
    pre = f"c{ci}."
    om = load_oracle_model(d)
    tm = TinyTransducer(copy.deepcopy(om))
    tm.decoder = ReplayDecoder(d, pre)
    tm.sos, tm.eos = (int(x) for x in d["sos_eos"])
    # ASRModel._forward_encoder (asr_model.py:275-300): makes patch_transducer bind the CTC prefix beam search
    tm._forward_encoder = lambda sp, ln, *a, **k: tm.encoder(sp, ln)
    tm = patch_transducer(tm.to(DEV).eval())
    beam, rw, cw, aw, tw, scw, stw = (float(x) for x in d[pre + "params"])
    speech = torch.from_numpy(d[pre + "speech"]).to(DEV)
    lens = torch.tensor([speech.size(1)], device=DEV)
    hyp, score = tm.transducer_attention_rescoring(
        speech, lens, int(beam), reverse_weight=rw, ctc_weight=cw, attn_weight=aw, transducer_weight=tw,
        search_ctc_weight=scw, search_transducer_weight=stw, beam_search_type=str(d[pre + "kind"]))
    assert hyp == d[pre + "best_hyp"].tolist()
    np.testing.assert_allclose(score, float(d[pre + "best_score"]), rtol=1e-4)


@pytest.mark.parametrize("ci", [0, 1, 2, 3])
def test_transducer_score_shared_encoder(golden, ci):
    """_cal_transducer_score over the n-best with ONE encoder projection == the reference's value on the
    repeated encoder output (fp32 path, rtol 1e-4)."""
    from wenet_celoss_b200 import patch_transducer
    from wenet_celoss_b200.transducer import transducer_score_shared
    d = golden("rescore")
    pre = f"c{ci}."
    tm = patch_transducer(TinyTransducer(load_oracle_model(d)).to(DEV).eval())
    speech = torch.from_numpy(d[pre + "speech"]).to(DEV)
    hp = torch.from_numpy(d[pre + "hyps_pad"]).to(DEV)
    hl = torch.from_numpy(d[pre + "hyps_lens"]).to(DEV)
    got = transducer_score_shared(tm, speech, hl, hp)
    np.testing.assert_allclose(got.cpu().numpy(), d[pre + "td_score"], rtol=1e-4)
    # and the drop-in signature on the repeated tensor (what the reference call site passes)
    N, T = hp.size(0), speech.size(1)
    rep = tm._cal_transducer_score(speech.repeat(N, 1, 1), torch.ones(N, 1, T, dtype=torch.bool, device=DEV), hl, hp)
    np.testing.assert_allclose(rep.cpu().numpy(), d[pre + "td_score"], rtol=1e-4)


@pytest.mark.parametrize("ci", [0, 1, 2, 3])
def test_attn_rescore_kernel(golden, ci):
    """b200_attn_rescore alone on the recorded decoder logits / scores: every hypothesis' attention part
    and the argmax (first maximum), fp64 sums."""
    from wenet_celoss_b200 import _lib
    from wenet_celoss_b200._lib import i32, ptr
    d = golden("rescore")
    pre = f"c{ci}."
    beam, rw, cw, aw, tw, _, _ = (float(x) for x in d[pre + "params"])
    dec = torch.from_numpy(d[pre + "dec"]).to(DEV).contiguous()
    rdec = torch.from_numpy(d[pre + "rdec"]).to(DEV).contiguous() if pre + "rdec" in d.files else None
    hp = torch.from_numpy(d[pre + "hyps_pad"]).to(DEV).contiguous()
    hl = torch.from_numpy(d[pre + "hyps_lens"]).to(DEV).int().contiguous()
    N, L1, V = dec.shape
    bs = torch.from_numpy(d[pre + "beam_score"]).to(DEV).double().contiguous()
    td = torch.from_numpy(d[pre + "td_score"]).to(DEV).float().contiguous()
    scores = torch.empty(N, dtype=torch.float64, device=DEV)
    attn = torch.empty(N, dtype=torch.float64, device=DEV)
    best = torch.empty(1, dtype=torch.int32, device=DEV)
    eos = int(d["sos_eos"][1])
    _lib.call("b200_attn_rescore", ptr(dec), ptr(rdec if rw > 0 else None), ptr(hp), ptr(hl), i32(N), i32(L1),
              i32(hp.size(1)), i32(V), i32(eos), ctypes.c_double(rw), ctypes.c_double(aw), ptr(bs),
              ctypes.c_double(cw), ptr(td), ctypes.c_double(tw), ptr(scores), ptr(attn), ptr(best),
              _lib.stream_handle(torch.device(DEV)))
    np.testing.assert_allclose(attn.cpu().numpy(), d[pre + "attn"], rtol=1e-5)
    assert int(best.item()) == int(d[pre + "best_index"])
    np.testing.assert_allclose(float(scores[int(best.item())]), float(d[pre + "best_score"]), rtol=1e-5)
    # oracle restatement on the same inputs
    hyps = [row[:n] for row, n in zip(d[pre + "hyps_pad"].tolist(), d[pre + "hyps_lens"].tolist())]
    dlp = torch.log_softmax(dec.cpu(), -1).numpy()
    rlp = torch.log_softmax(rdec.cpu(), -1).numpy() if rdec is not None else None
    bi, bsc, at = orc.attention_rescoring(hyps, d[pre + "beam_score"].tolist(), d[pre + "td_score"].tolist(), dlp,
                                          rlp, eos, rw, aw, cw, tw)
    assert bi == int(best.item())
    np.testing.assert_allclose(scores.cpu().numpy()[bi], bsc, rtol=1e-5)
