"""pytest configuration: the `gpu` marker and shared helpers.

`-m "not gpu"` runs in the build container (no GPU): oracle vs golden fixtures, host
logic, C-ABI symbol checks.  `-m gpu` runs on a B200: CUDA path vs oracle, through
the C ABI.  Nothing here reads /root/reference.
"""
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(GOLDEN / f"{name}.npz")
    return load


def load_oracle_model(d, prefix="w."):
    """OracleModel with the weights stored in a golden fixture."""
    from oracle import reference as orc
    V, E, P, J, H, L = (int(x) for x in d["dims"])
    m = orc.OracleModel(V, E, P, J, E, H, L, blank=0)
    sd = {k[len(prefix):]: torch.from_numpy(d[k]) for k in d.files if k.startswith(prefix)}
    missing = m.load_state_dict(sd, strict=False)
    assert not missing.unexpected_keys, missing
    m.eval()
    return m


class _Mha(torch.nn.Module):
    """parameter container with the reference MultiHeadedAttention's names (attention.py:35-46)."""

    def __init__(self, n_head, n_feat):
        super().__init__()
        self.h, self.d_k = n_head, n_feat // n_head
        self.linear_q = torch.nn.Linear(n_feat, n_feat)
        self.linear_k = torch.nn.Linear(n_feat, n_feat)
        self.linear_v = torch.nn.Linear(n_feat, n_feat)
        self.linear_out = torch.nn.Linear(n_feat, n_feat)


class HotwordBiasStandIn(torch.nn.Module):
    """The slice of ContextBias (context_bias.py:159-399) the greedy path reads, filled from a
    golden fixture: predictor-bias parameters under the reference's attribute names;
    forward_bias_hidden returns the stored phrase vectors (given / empty list) and
    forward_encoder_bias the stored biased encoder stream (the phrase extractor and the encoder-side
    attention are library modules outside the hot path)."""

    def __init__(self, d):
        super().__init__()
        P = int(d["dims"][1])
        self.predictor_bias = _Mha(int(d["heads"]), P)
        self.predictor_bias_bias_norm = torch.nn.LayerNorm(P)
        self.predictor_bias_combine = torch.nn.Linear(2 * P, P)
        self.predictor_bias_out_norm = torch.nn.LayerNorm(P)
        keys = ("cb.predictor_bias",)
        if "enc_empty" in d.files:   # gated fixture: ContextBias.forward_hw_pred* modules (context_bias.py:382-394)
            n_lab, D = d["cb.hw_output_layer.weight"].shape
            self.hw_bias = _Mha(int(d["heads"]), int(D))
            self.hw_bias_norm = torch.nn.LayerNorm(int(D))
            self.hw_output_layer = torch.nn.Linear(int(D), int(n_lab))
            self.hw_output_layer_enc = torch.nn.Linear(P, int(D))
            self.hw_output_layer_dec = torch.nn.Linear(P, int(D))
            keys += ("cb.hw_bias.", "cb.hw_bias_norm.", "cb.hw_output_layer")
        sd = {k[3:]: torch.from_numpy(d[k]) for k in d.files if k.startswith(keys)}
        self.load_state_dict(sd, strict=True)
        self.register_buffer("bias_hidden", torch.from_numpy(d["bias_hidden"]))
        self.register_buffer("bias_hidden_empty", torch.from_numpy(d["bias_hidden_empty"]))
        self.register_buffer("enc", torch.from_numpy(d["enc"]))
        self.register_buffer("enc_biased", torch.from_numpy(d["enc_biased"]))
        if "enc_empty" in d.files:
            self.register_buffer("enc_empty", torch.from_numpy(d["enc_empty"]))
            self.register_buffer("enc_bias", torch.from_numpy(d["enc_bias"]))

    def forward_bias_hidden(self, context_list, context_lengths):
        return self.bias_hidden_empty if context_list.numel() == 1 else self.bias_hidden

    def forward_encoder_bias(self, bias_hidden, encoder_out):
        assert torch.equal(encoder_out, self.enc)
        if bias_hidden is self.bias_hidden_empty:
            return self.enc_empty, None
        assert bias_hidden is self.bias_hidden
        return self.enc_biased, getattr(self, "enc_bias", None)

    def forward_hw_pred_both(self, h_enc_bias, h_dec_bias):
        """context_bias.py:388-394 for [N,1,E] inputs: one key per query, the attention weight is 1."""
        F = torch.nn.functional
        assert h_enc_bias.size(1) == 1
        h_enc = self.hw_output_layer_enc(h_enc_bias)
        a = self.hw_bias.linear_out(self.hw_bias.linear_v(h_enc))
        return self.hw_output_layer(self.hw_bias_norm(a))


def load_hotword_model(d):
    m = load_oracle_model(d)
    m.context_bias = HotwordBiasStandIn(d)
    m.eval()
    return m
