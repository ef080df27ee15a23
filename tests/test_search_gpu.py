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
