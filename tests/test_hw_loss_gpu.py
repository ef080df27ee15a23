"""GPU parity: hot-word cross-entropy loss (csrc/ce_loss.cu) and the hw_loss hook of
Transducer.forward (transducer.py:170-262).  fp32: loss rtol 1e-6, gradient atol 1e-7."""
import copy

import numpy as np
import pytest
import torch

from oracle import reference as orc
from test_transducer_gpu import TinyTransducer, _model

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _run(x_np, y_np, ignore_index=-100, dy=1.0):
    from wenet_celoss_b200 import hw_ce_loss
    x = torch.from_numpy(x_np).to(DEV).requires_grad_(True)
    loss = hw_ce_loss(x, torch.from_numpy(y_np).to(DEV), ignore_index)
    (loss * dy).backward()
    return loss.item(), x.grad.cpu().numpy()


def test_golden_three_loss_modes(golden):
    d = golden("hw_loss")
    for mode in d["modes"]:
        loss, grad = _run(d[f"{mode}.hw_output"], d[f"{mode}.labels"])
        np.testing.assert_allclose(loss, float(d[f"{mode}.hw_loss"]), rtol=1e-6)
        np.testing.assert_allclose(grad, d[f"{mode}.grad"], atol=1e-7)


@pytest.mark.parametrize("B,C,U", [(1, 2, 1), (7, 2, 33), (32, 31, 101), (64, 5, 400)])
def test_against_oracle_with_ignored_positions(B, C, U):
    g = torch.Generator().manual_seed(B * 100 + U)
    x = (torch.randn(B, C, U, generator=g) * 3).numpy()
    y = torch.randint(0, C, (B, U), generator=g)
    y[torch.rand(B, U, generator=g) < 0.2] = -100
    y[0, 0] = 0
    want, gwant = orc.hw_ce_loss(x, y.numpy())
    loss, grad = _run(x, y.numpy(), dy=0.4)
    np.testing.assert_allclose(loss, want, rtol=1e-6)
    np.testing.assert_allclose(grad, 0.4 * gwant, atol=1e-7)
    assert np.all(grad[np.broadcast_to((y.numpy() == -100)[:, None, :], grad.shape)] == 0)
    # torch's own fp32 CrossEntropyLoss (what the reference calls)
    ref = torch.nn.CrossEntropyLoss()(torch.from_numpy(x), y)
    np.testing.assert_allclose(loss, ref.item(), rtol=1e-5)


class _HwHeads(torch.nn.Module):
    """stand-in for the hot-word side of ContextBias (context_bias.py:341-399): identity feature
    biasing, linear classifier heads; only the call pattern of Transducer.forward matters here."""

    def __init__(self, D, n_labels, odim):
        super().__init__()
        self.dec = torch.nn.Linear(D, n_labels)
        self.both = torch.nn.Linear(D, n_labels)
        self.sep = torch.nn.Linear(D, odim)

    def forward_bias_hidden(self, context_list, context_lengths):
        return None

    def forward_encoder_bias(self, bias_hidden, x):
        return x, x * 0.5

    def forward_predictor_bias(self, bias_hidden, x):
        return x, x * 0.25

    def forward_hw_pred(self, bias_hidden, predictor_out):
        return self.dec(predictor_out)

    def forward_hw_pred_both(self, h_enc_bias, h_dec_bias):
        return self.both(h_dec_bias) + h_enc_bias.mean(dim=1, keepdim=True)[..., :self.both.out_features]

    def forward_hw_pred_both_sep(self, h_enc_bias, h_dec_bias):
        return None, self.sep(h_dec_bias)


@pytest.mark.parametrize("mode", ["pred", "both", "sep"])
def test_forward_hook(mode):
    """loss dict of the patched forward with hw_weight = 0.4: hw_loss equals the reference's
    slicing / end_blank / add_blank recipe (transducer.py:172-200) evaluated by the oracle."""
    from wenet_celoss_b200 import patch_transducer
    om = _model(seed=2)
    heads = _HwHeads(32, 3, 9)
    tm = TinyTransducer(copy.deepcopy(om))
    tm.context_bias, tm.hw_weight, tm.loss_mode = heads, 0.4, mode
    tm = patch_transducer(tm.to(DEV))
    g = torch.Generator().manual_seed(9)
    B, T, U = 3, 15, 5
    enc = torch.randn(B, T, 32, generator=g)
    tl, ul = torch.tensor([15, 12, 9]), torch.tensor([5, 3, 4])
    text = torch.randint(1, 200, (B, U), generator=g)
    hw = torch.randint(0, 3, (B, U), generator=g)
    for b in range(B):
        text[b, ul[b]:] = -1
        hw[b, ul[b]:] = -1
    out = tm(enc.to(DEV), tl.to(DEV), text.to(DEV), ul.to(DEV), None, None, hw.to(DEV))
    # oracle side on the CPU
    cpu_heads = copy.deepcopy(heads).cpu()
    pred_out = om.predictor(orc.add_blank(text, 0, -1))
    if mode == "pred":
        x = cpu_heads.forward_hw_pred(None, pred_out).permute(0, 2, 1)[..., :-1]
        y = torch.cat([hw, torch.zeros(B, 1, dtype=hw.dtype)], 1).masked_fill(
            torch.cat([hw, torch.zeros(B, 1, dtype=hw.dtype)], 1) == -1, 0)[..., :-1]
    elif mode == "both":
        x = cpu_heads.forward_hw_pred_both(enc * 0.5, pred_out * 0.25).permute(0, 2, 1)[..., :-1]
        y = hw.masked_fill(hw == -1, 0)
    else:
        x = cpu_heads.forward_hw_pred_both_sep(enc * 0.5, pred_out * 0.25)[1].permute(0, 2, 1)
        y = orc.add_blank(hw, 0, -1)
    want, _ = orc.hw_ce_loss(x.detach().numpy(), y.numpy())
    np.testing.assert_allclose(out["hw_loss"].item(), want, rtol=1e-5)
    base = out["loss"].item() - 0.4 * out["hw_loss"].item()
    np.testing.assert_allclose(base, 0.7 * out["loss_rnnt"].item() + 0.3 * out["loss_ctc"].item(), rtol=1e-5)
    out["loss"].backward()
    head = {"pred": tm.context_bias.dec, "both": tm.context_bias.both, "sep": tm.context_bias.sep}[mode]
    assert head.weight.grad is not None and float(head.weight.grad.abs().sum()) > 0
