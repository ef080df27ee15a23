"""GPU parity: rnnt_loss (C ABI -> sm_100a kernels) against the oracle.

Tolerances (north_star): fp32 loss and grads rtol 1e-4 (atol 1e-6 on grads);
fp16/bf16 logits rtol 1e-2 against the oracle run in fp32 on the rounded inputs.
"""
import numpy as np
import pytest
import torch
import torchaudio

from oracle import reference as orc
from oracle.c_oracle import rnnt_loss_c

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _run(logits, targets, tl, ul, blank=0, reduction="none", clamp=-1.0, dy=None):
    from wenet_celoss_b200 import rnnt_loss
    x = logits.to(DEV).requires_grad_(True)
    out = rnnt_loss(x, targets.to(DEV), tl.to(DEV), ul.to(DEV), blank=blank, clamp=clamp,
                    reduction=reduction)
    if dy is None:
        out.sum().backward()
    else:
        out.backward(dy.to(DEV).to(out.dtype))
    return out.detach().float().cpu().numpy(), x.grad.detach().float().cpu().numpy()


def _rand_case(B, T, U, V, seed, ragged=True, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(B, T, U + 1, V, generator=g) * scale
    targets = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32)
    if ragged:
        tl = torch.randint(max(1, T // 2), T + 1, (B,), generator=g, dtype=torch.int32)
        ul = torch.randint(U // 2, U + 1, (B,), generator=g, dtype=torch.int32)
        tl[0], ul[-1] = T, U
    else:
        tl = torch.full((B,), T, dtype=torch.int32)
        ul = torch.full((B,), U, dtype=torch.int32)
    return logits, targets, tl, ul


def test_golden_small(golden):
    d = golden("rnnt_joint_small")
    tgt, tl, ul = orc.rnnt_targets(torch.from_numpy(d["text"]), torch.from_numpy(d["text_lens"]),
                                   torch.from_numpy(d["enc_lens"]))
    B = d["logits"].shape[0]
    costs, grad = _run(torch.from_numpy(d["logits"]), tgt, tl, ul, dy=torch.full((B,), 1.0 / B))
    np.testing.assert_allclose(costs, d["costs"], rtol=1e-4)
    np.testing.assert_allclose(grad, d["grad_logits"], rtol=1e-4, atol=1e-6)


def test_golden_edge_cases(golden):
    d = golden("rnnt_edge")
    i32 = lambda a: torch.tensor(a, dtype=torch.int32)
    c, g = _run(torch.from_numpy(d["u0_logits"]), torch.zeros(2, 0, dtype=torch.int32), i32([4, 2]),
                i32([0, 0]))
    np.testing.assert_allclose(c, d["u0_costs"], rtol=1e-4)
    np.testing.assert_allclose(g, d["u0_grad"], rtol=1e-4, atol=1e-6)
    c, g = _run(torch.from_numpy(d["t1_logits"]), torch.from_numpy(d["t1_targets"]), i32([1, 1]),
                i32([2, 1]), blank=-1)
    np.testing.assert_allclose(c, d["t1_costs"], rtol=1e-4)
    np.testing.assert_allclose(g, d["t1_grad"], rtol=1e-4, atol=1e-6)
    c, g = _run(torch.from_numpy(d["tb_logits"]), torch.from_numpy(d["tb_targets"]),
                torch.from_numpy(d["tb_tlen"]), torch.from_numpy(d["tb_ulen"]), reduction="sum",
                clamp=0.05)
    np.testing.assert_allclose(c, d["tb_sum"], rtol=1e-4)
    np.testing.assert_allclose(g, d["tb_grad"], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("shape", [(4, 100, 20, 500), (2, 33, 7, 4233), (3, 17, 9, 1030), (2, 5, 3, 3)])
@pytest.mark.parametrize("ragged", [False, True])
def test_fp32_vs_torchaudio_and_c_oracle(shape, ragged):
    B, T, U, V = shape
    logits, targets, tl, ul = _rand_case(B, T, U, V, seed=B * 1000 + V, ragged=ragged, scale=2.0)
    dy = torch.linspace(0.5, 1.5, B)
    costs, grad = _run(logits, targets, tl, ul, dy=dy)
    x = logits.clone().requires_grad_(True)
    ref = torchaudio.functional.rnnt_loss(x, targets, tl, ul, blank=0, reduction="none")
    ref.backward(dy)
    ta = x.grad.numpy()
    c64, g64, _, _ = rnnt_loss_c(logits, targets, tl, ul, blank=0, dy=dy.double().numpy())
    np.testing.assert_allclose(costs, ref.detach().numpy(), rtol=1e-4)
    np.testing.assert_allclose(costs, c64, rtol=1e-4)
    # Our lattice is rescaled per anti-diagonal (fp64 offsets), so it tracks the fp64 oracle
    # to rtol 1e-4 / atol 1e-6.  torchaudio's fp32 log-space carries a cost*2^-23 noise
    # floor (alpha+beta+cost cancel at magnitude ~cost); allow for ITS distance to fp64.
    np.testing.assert_allclose(grad, g64, rtol=1e-4, atol=1e-6)
    floor = np.abs(ta - g64).max()
    np.testing.assert_allclose(grad, ta, rtol=1e-4, atol=1e-6 + 2.0 * floor)
    for b in range(B):  # padded cells: exactly zero
        assert np.all(grad[b, tl[b]:] == 0) and np.all(grad[b, :, ul[b] + 1:] == 0)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_half_precisions(dtype):
    B, T, U, V = 3, 40, 12, 1000
    logits, targets, tl, ul = _rand_case(B, T, U, V, seed=5, scale=2.0)
    lq = logits.to(dtype)
    costs, grad = _run(lq, targets, tl, ul)
    c64, g64, _, _ = rnnt_loss_c(lq.float(), targets, tl, ul, blank=0)
    np.testing.assert_allclose(costs, c64, rtol=1e-2)
    np.testing.assert_allclose(grad, g64, rtol=1e-2, atol=1e-3 if dtype == torch.bfloat16 else 2e-4)
    from wenet_celoss_b200 import rnnt_loss
    out = rnnt_loss(lq.to(DEV), targets.to(DEV), tl.to(DEV), ul.to(DEV), blank=0, reduction="none")
    assert out.dtype == (torch.float16 if dtype == torch.float16 else torch.float32)


def test_reductions_and_mean_matches_reference_call():
    from wenet_celoss_b200 import rnnt_loss
    logits, targets, tl, ul = _rand_case(4, 20, 6, 50, seed=11)
    args = (logits.to(DEV), targets.to(DEV), tl.to(DEV), ul.to(DEV))
    none = rnnt_loss(*args, blank=0, reduction="none")
    assert torch.allclose(rnnt_loss(*args, blank=0, reduction="mean"), none.mean())
    assert torch.allclose(rnnt_loss(*args, blank=0, reduction="sum"), none.sum())
    ref = torchaudio.functional.rnnt_loss(logits, targets, tl, ul, blank=0, reduction="mean")
    np.testing.assert_allclose(rnnt_loss(*args, blank=0, reduction="mean").item(), ref.item(), rtol=1e-4)
    # positional call exactly as wenet/transducer/transducer.py:142-147
    l2 = rnnt_loss(args[0], args[1], args[2], args[3], blank=0, reduction="mean")
    np.testing.assert_allclose(l2.item(), ref.item(), rtol=1e-4)


def test_validation_errors_match_torchaudio():
    from wenet_celoss_b200 import rnnt_loss
    logits, targets, tl, ul = (t.to(DEV) for t in _rand_case(2, 5, 3, 6, seed=1, ragged=False))
    cases = [
        (lambda: rnnt_loss(logits, targets.long(), tl, ul), "targets must be int32 type"),
        (lambda: rnnt_loss(logits, targets, tl.long(), ul), "logit_lengths must be int32 type"),
        (lambda: rnnt_loss(logits, targets, tl, ul.long()), "target_lengths must be int32 type"),
        (lambda: rnnt_loss(logits.double(), targets, tl, ul), "logits must be float32 or float16"),
        (lambda: rnnt_loss(logits.transpose(1, 2).contiguous().transpose(1, 2), targets, tl, ul),
         "logits must be contiguous"),
        (lambda: rnnt_loss(logits, targets, tl - 1, ul), "input length mismatch"),
        (lambda: rnnt_loss(logits, targets, tl, ul - 1), "output length mismatch"),
        (lambda: rnnt_loss(logits, targets, tl, ul, blank=6), "blank must be within"),
        (lambda: rnnt_loss(logits[0], targets, tl, ul), "logits must be 4-D"),
    ]
    for fn, msg in cases:
        with pytest.raises(RuntimeError, match=msg):
            fn()
    with pytest.raises(ValueError, match="reduction should be one of"):
        rnnt_loss(logits, targets, tl, ul, reduction="foo")
    with pytest.raises(RuntimeError, match="CUDA"):
        rnnt_loss(logits.cpu(), targets.cpu(), tl.cpu(), ul.cpu(), blank=0)


def test_properties_at_full_size_bf16():
    """BASELINE cfg3 shape (B=32,T=500,U=100,V=5000, bf16): size-independent checks."""
    from wenet_celoss_b200 import rnnt_loss
    B, T, U, V = 32, 500, 100, 5000
    g = torch.Generator(device=DEV).manual_seed(3)
    logits = torch.randn(B, T, U + 1, V, generator=g, device=DEV, dtype=torch.bfloat16)
    logits.requires_grad_(True)
    targets = torch.randint(1, V, (B, U), generator=g, device=DEV, dtype=torch.int32)
    tl = torch.randint(T // 2, T + 1, (B,), generator=g, device=DEV, dtype=torch.int32)
    ul = torch.randint(U // 2, U + 1, (B,), generator=g, device=DEV, dtype=torch.int32)
    tl[0], ul[1] = T, U
    costs = rnnt_loss(logits, targets, tl, ul, blank=0, reduction="none")
    costs.sum().backward()
    grad = logits.grad
    assert torch.all(costs > 0) and torch.all(torch.isfinite(costs))
    for b in (0, 1, 7, 31):  # padded cells exactly zero; valid cells sum to ~0 over V
        Tb, Ub = int(tl[b]), int(ul[b])
        assert torch.count_nonzero(grad[b, Tb:]) == 0
        assert torch.count_nonzero(grad[b, :, Ub + 1:]) == 0
        s = grad[b, :Tb, :Ub + 1].float().sum(-1)
        assert s.abs().max() < 2e-2
    # one utterance against the CPU oracle (fp32 maths on the bf16 values)
    b = 7
    Tb, Ub = int(tl[b]), int(ul[b])
    x = logits.detach()[b:b + 1, :Tb, :Ub + 1].float().cpu().contiguous()
    c64, g64, _, _ = rnnt_loss_c(x, targets[b:b + 1, :Ub].cpu(), [Tb], [Ub], blank=0)
    np.testing.assert_allclose(costs[b].item(), c64[0], rtol=1e-3)
    np.testing.assert_allclose(grad[b, :Tb, :Ub + 1].float().cpu().numpy(), g64[0], rtol=1e-2, atol=1e-3)
    # permuting the batch permutes the costs
    perm = torch.randperm(B, device=DEV)
    del grad
    logits.grad = None
    c2 = rnnt_loss(logits.detach()[perm].contiguous(), targets[perm].contiguous(), tl[perm].contiguous(),
                   ul[perm].contiguous(), blank=0, reduction="none")
    assert torch.equal(c2, costs.detach()[perm])
