"""GPU parity: batched CTC prefix beam search (integer outputs: n-best token lists identical to
the reference's, scores rtol 1e-9 — both sides accumulate in fp64)."""
import numpy as np
import pytest
import torch

from oracle import reference as orc
from test_oracle_golden import KAT_HYPS, KAT_PROBS, KAT_SCORES

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_reference_known_answer():
    """runtime/core/test/ctc_prefix_beam_search_test.cc:30-61 through the CUDA path."""
    from wenet_celoss_b200.ctc_search import ctc_prefix_beam_search_batch
    lp = torch.log(torch.tensor(KAT_PROBS)).unsqueeze(0).to(DEV)
    hyps = ctc_prefix_beam_search_batch(lp, torch.tensor([3]), 3)[0]
    assert [h[0] for h in hyps] == KAT_HYPS
    np.testing.assert_allclose(np.exp([h[1] for h in hyps]), KAT_SCORES, rtol=1e-6)


def test_golden_batched(golden):
    """every fixture case of the real reference, one launch per case (B = 1 as the reference)."""
    from wenet_celoss_b200.ctc_search import ctc_prefix_beam_search_batch
    d = golden("ctc_search")
    for name in d["names"]:
        logits = torch.from_numpy(d[f"{name}.logits"])
        beam = int(d[f"{name}.beam"])
        lp = torch.log_softmax(logits.unsqueeze(0), dim=2).to(DEV)  # same CPU log_softmax as the fixture
        hyps = ctc_prefix_beam_search_batch(lp, torch.tensor([logits.size(0)]), beam)[0]
        assert len(hyps) == int(d[f"{name}.n"]), name
        for k, (prefix, score) in enumerate(hyps):
            assert list(prefix) == d[f"{name}.tokens"][k][:d[f"{name}.lens"][k]].tolist(), (name, k)
            np.testing.assert_allclose(score, d[f"{name}.scores"][k], rtol=1e-9)


@pytest.mark.parametrize("B,T,V,beam", [(5, 37, 64, 4), (3, 80, 500, 10), (2, 30, 5000, 10)])
def test_ragged_batch_against_oracle(B, T, V, beam):
    from wenet_celoss_b200.ctc_search import ctc_prefix_beam_search_batch
    g = torch.Generator().manual_seed(B * 1000 + T)
    logits = torch.randn(B, T, V, generator=g) * 2.0
    logits[:, :, 0] += 2.0
    lens = torch.randint(T // 2, T + 1, (B,), generator=g)
    lens[0] = T
    lp = torch.log_softmax(logits, dim=2)
    got = ctc_prefix_beam_search_batch(lp.to(DEV), lens, beam)
    for b in range(B):
        want = orc.ctc_prefix_beam_search(lp[b, :int(lens[b])], beam)
        assert [h[0] for h in got[b]] == [h[0] for h in want], b
        np.testing.assert_allclose([h[1] for h in got[b]], [h[1] for h in want], rtol=1e-9)


def test_model_level_entry_points():
    """_ctc_prefix_beam_search / ctc_prefix_beam_search keep the reference's signatures
    (asr_model.py:326-443)."""
    from wenet_celoss_b200 import ctc_search

    class Model:
        class ctc:  # noqa: N801
            @staticmethod
            def log_softmax(x):
                return torch.log_softmax(x, dim=2)

        @staticmethod
        def _forward_encoder(speech, lengths, *a):
            return speech, None

    g = torch.Generator().manual_seed(5)
    x = torch.randn(1, 25, 40, generator=g)
    hyps, enc = ctc_search._ctc_prefix_beam_search(Model(), x.to(DEV), torch.tensor([25]), 5)
    want = orc.ctc_prefix_beam_search(torch.log_softmax(x, dim=2)[0], 5)
    assert [h[0] for h in hyps] == [h[0] for h in want]
    best = ctc_search.ctc_prefix_beam_search(Model(), x.to(DEV), torch.tensor([25]), 5)
    assert best[0] == want[0][0]


def test_forced_align_golden(golden):
    """alignments of the real forced_align (ctc_util.py:27-83), bit-identical (fp32 max/add)."""
    from wenet_celoss_b200.ctc_search import forced_align
    d = golden("ctc_search")
    for name in d["align_names"]:
        got = forced_align(torch.from_numpy(d[f"align.{name}.logp"]).to(DEV),
                           torch.from_numpy(d[f"align.{name}.y"]), 0)
        assert got == d[f"align.{name}.alignment"].tolist(), name


def test_forced_align_ragged_batch():
    from wenet_celoss_b200.ctc_search import forced_align_batch
    g = torch.Generator().manual_seed(77)
    B, T, V, Lmax = 6, 90, 300, 20
    lp = torch.log_softmax(torch.randn(B, T, V, generator=g) * 2.0, dim=2)
    hl = torch.randint(45, T + 1, (B,), generator=g)
    yl = torch.randint(1, Lmax + 1, (B,), generator=g)
    ys = torch.randint(1, V, (B, Lmax), generator=g)
    ys[1, 1] = ys[1, 0]  # a repeated label
    yl[1] = max(int(yl[1]), 2)
    got = forced_align_batch(lp.to(DEV), ys, hl, yl, 0)
    for b in range(B):
        want = orc.forced_align(lp[b, :int(hl[b])], ys[b, :int(yl[b])], 0)
        assert got[b] == want, b
