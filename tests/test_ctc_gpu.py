"""GPU parity: CTC module / ctc_loss (C ABI -> sm_100a kernels) against the oracle.

Tolerances: fp32 loss rtol 1e-4, grads rtol 1e-4 + atol 1e-6 (fp32 log-space noise floor
handled as in test_rnnt_gpu); half precisions rtol 1e-2 on the rounded inputs.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import reference as orc
from oracle.c_oracle import ctc_loss_c

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(B, T, L, V, seed, repeats=True):
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(B, T, V, generator=g) * 2
    hl = torch.randint(max(2 * L + 1, T // 2), T + 1, (B,), generator=g, dtype=torch.int32)
    yl = torch.randint(0, L + 1, (B,), generator=g, dtype=torch.int32)
    hl[0], yl[0] = T, L
    if B > 1:
        yl[1] = 0
    ys = torch.randint(1, V, (B, L), generator=g)
    if repeats:
        for b in range(B):
            for i in range(1, L):
                if torch.rand((), generator=g) < 0.2:
                    ys[b, i] = ys[b, i - 1]
    for b in range(B):
        ys[b, yl[b]:] = -1
    return logits, ys, hl, yl


def _mine(logits, ys, hl, yl, scale):
    from wenet_celoss_b200 import ctc_loss
    x = logits.to(DEV).requires_grad_(True)
    nll = ctc_loss(x, ys.to(DEV), hl.to(DEV), yl.to(DEV), blank=0, reduction="none")
    (nll.float() * scale.to(DEV)).sum().backward()
    return nll.detach().float().cpu().numpy(), x.grad.detach().float().cpu().numpy()


def test_module_golden(golden):
    from wenet_celoss_b200 import CTC
    d = golden("ctc_small")
    V, E = int(d["dims"][0]), int(d["dims"][1])
    m = CTC(V, E).to(DEV)
    m.load_state_dict({k[len("w.ctc."):]: torch.from_numpy(d[k]) for k in d.files if k.startswith("w.ctc.")})
    hs = torch.from_numpy(d["hs"]).to(DEV).requires_grad_(True)
    loss = m(hs, torch.from_numpy(d["hlens"]).to(DEV), torch.from_numpy(d["ys"]).to(DEV),
             torch.from_numpy(d["ylens"]).to(DEV))
    np.testing.assert_allclose(loss.item(), d["loss"], rtol=1e-4)
    loss.backward()
    np.testing.assert_allclose(hs.grad.cpu().numpy(), d["grad_hs"], rtol=1e-4, atol=1e-6)
    for n, p in m.named_parameters():
        np.testing.assert_allclose(p.grad.cpu().numpy(), d["g.ctc." + n], rtol=1e-4, atol=2e-6, err_msg=n)
    with torch.no_grad():
        ref = orc.OracleCTC(V, E)
        ref.load_state_dict({k: v.cpu() for k, v in m.state_dict().items()})
        np.testing.assert_allclose(m.log_softmax(hs).cpu().numpy(),
                                   ref.log_softmax(hs.cpu()).numpy(), atol=1e-5)
        assert torch.equal(m.argmax(hs).cpu(), ref.argmax(hs.cpu()))


@pytest.mark.parametrize("shape", [(4, 60, 10, 500), (3, 50, 7, 4233), (2, 9, 2, 5), (5, 120, 40, 1030)])
def test_fp32_vs_aten_and_c_oracle(shape):
    B, T, L, V = shape
    logits, ys, hl, yl = _case(B, T, L, V, seed=T * 7 + V)
    scale = torch.linspace(0.5, 1.5, B)
    nll, grad = _mine(logits, ys, hl, yl, scale)
    x = logits.clone().requires_grad_(True)
    ref = F.ctc_loss(x.transpose(0, 1).log_softmax(2), ys, hl, yl, blank=0, reduction="none")
    (ref * scale).sum().backward()
    ta = x.grad.numpy()
    n64 = np.zeros(B)
    g64 = np.zeros_like(ta)
    for b in range(B):  # per-utterance scale through the C oracle
        n1, g1 = ctc_loss_c(logits[b:b + 1], ys[b:b + 1], hl[b:b + 1], yl[b:b + 1], blank=0,
                            scale=float(scale[b]))
        n64[b], g64[b] = n1[0], g1[0]
    np.testing.assert_allclose(nll, ref.detach().numpy(), rtol=1e-4)
    np.testing.assert_allclose(nll, n64, rtol=1e-4)
    # per-frame rescaled trellis (fp64 offsets): tracks the fp64 oracle directly (measured
    # ~1.6e-5 abs on O(1) elements).  ATen's fp32 CPU path sits up to 3e-4 from fp64 (its
    # log_softmax backward multiplies every row by a noisy 1 - sum(posterior)), so the
    # comparison against it allows for ATen's own distance to fp64.
    np.testing.assert_allclose(grad, g64, rtol=1e-4, atol=1e-6)
    floor = np.abs(ta - g64).max()
    floor_rel = (np.abs(ta - g64) / (np.abs(g64) + 1e-6)).max()
    np.testing.assert_allclose(grad, ta, rtol=1e-4 + 2.0 * floor_rel, atol=1e-6 + 2.0 * floor)
    for b in range(B):
        assert np.all(grad[b, hl[b]:] == 0)
    assert np.abs(grad.sum(-1)).max() < 1e-4  # sum_c grad = scale*(1 - sum_s posterior)


@pytest.mark.parametrize("shape,dtype", [((3, 50, 7, 4233), torch.float32), ((2, 301, 140, 1030), torch.float32),
                                         ((5, 17, 3, 7), torch.float32), ((3, 43, 9, 1001), torch.bfloat16),
                                         ((2, 33, 6, 250), torch.float16)])
def test_grad_kernel_shapes(shape, dtype):
    """ctc_grad_kernel writes the label entries inside the row stream (bitmap over the classes + rank into
    the class-ordered entries of the row).  Shapes: frames not a multiple of the 8-row group, rows not
    16-byte aligned (odd V: the per-element bitmap path), more than 256 extended states (the loops past
    the register-resident states), 8 elements per 16-byte segment; every label entry and every plain entry
    against the fp64 oracle."""
    B, T, L, V = shape
    logits, ys, hl, yl = _case(B, T, L, V, seed=T + V)
    lq = logits.to(dtype)
    scale = torch.linspace(0.5, 1.5, B)
    grad = _mine(lq, ys, hl, yl, scale)[1]
    g64 = np.zeros_like(grad)
    for b in range(B):
        g64[b] = ctc_loss_c(lq[b:b + 1].float(), ys[b:b + 1], hl[b:b + 1], yl[b:b + 1], blank=0,
                            scale=float(scale[b]))[1][0]
    if dtype == torch.float32:
        np.testing.assert_allclose(grad, g64, rtol=1e-4, atol=1e-6)
    else:
        np.testing.assert_allclose(grad, g64, rtol=1e-2, atol=6e-3 if dtype == torch.bfloat16 else 8e-4)
    for b in range(B):
        assert np.all(grad[b, hl[b]:] == 0)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_half_precisions(dtype):
    logits, ys, hl, yl = _case(3, 40, 8, 1000, seed=3)
    lq = logits.to(dtype)
    nll, grad = _mine(lq, ys, hl, yl, torch.ones(3))
    n64, g64 = ctc_loss_c(lq.float(), ys, hl, yl, blank=0, scale=1.0)
    np.testing.assert_allclose(nll, n64, rtol=1e-2)
    np.testing.assert_allclose(grad, g64, rtol=1e-2, atol=4e-3 if dtype == torch.bfloat16 else 5e-4)


def test_infeasible_alignment_is_inf_and_empty_target():
    from wenet_celoss_b200 import ctc_loss
    logits = torch.randn(2, 3, 6)
    ys = torch.tensor([[1, 1, 2], [-1, -1, -1]])  # needs >= 4 frames (repeat) but has 3
    hl = torch.tensor([3, 3], dtype=torch.int32)
    yl = torch.tensor([3, 0], dtype=torch.int32)
    nll = ctc_loss(logits.to(DEV), ys.to(DEV), hl.to(DEV), yl.to(DEV), reduction="none").cpu()
    assert torch.isinf(nll[0]) and nll[0] > 0
    lp = logits[1].log_softmax(-1)
    np.testing.assert_allclose(nll[1].item(), -lp[:, 0].sum().item(), rtol=1e-5)


def test_patch_reference_style_module():
    from wenet_celoss_b200 import patch_ctc
    ref = orc.OracleCTC(50, 16)
    ref.dropout_rate = 0.0
    hs = torch.randn(3, 20, 16)
    hl = torch.tensor([20, 15, 18])
    ys = torch.tensor([[3, 4, 4, 9], [7, -1, -1, -1], [1, 2, -1, -1]])
    yl = torch.tensor([4, 1, 2])
    want = orc.ctc_forward(hs, hl, ys, yl, ref)
    import copy
    m = patch_ctc(copy.deepcopy(ref).to(DEV))
    got = m(hs.to(DEV), hl.to(DEV), ys.to(DEV), yl.to(DEV))
    np.testing.assert_allclose(got.item(), want.item(), rtol=1e-4)


def test_properties_at_full_size():
    """BASELINE cfg5 shape (B=64, T=1000, V=5000, L<=100): size-independent checks."""
    from wenet_celoss_b200 import ctc_loss
    B, T, V, L = 64, 1000, 5000, 100
    g = torch.Generator(device=DEV).manual_seed(5)
    logits = torch.randn(B, T, V, generator=g, device=DEV).requires_grad_(True)
    yl = torch.randint(50, L + 1, (B,), generator=g, device=DEV, dtype=torch.int32)
    hl = torch.randint(600, T + 1, (B,), generator=g, device=DEV, dtype=torch.int32)
    ys = torch.randint(1, V, (B, L), generator=g, device=DEV)
    rep = torch.rand(B, L, generator=g, device=DEV) < 0.1
    ys[:, 1:] = torch.where(rep[:, 1:], ys[:, :-1], ys[:, 1:])
    ys = torch.where(torch.arange(L, device=DEV)[None] < yl[:, None], ys, torch.full_like(ys, -1))
    nll = ctc_loss(logits, ys, hl, yl, reduction="none")
    (nll.sum() / B).backward()
    grad = logits.grad
    assert torch.all(torch.isfinite(nll)) and torch.all(nll > 0)
    assert grad.sum(-1).abs().max() < 1e-5  # (1 - sum_s posterior)/B
    for b in (0, 17, 63):
        assert torch.count_nonzero(grad[b, int(hl[b]):]) == 0
    b = 17
    n64, g64 = ctc_loss_c(logits.detach()[b:b + 1].cpu(), ys[b:b + 1].cpu(), hl[b:b + 1].cpu(),
                          yl[b:b + 1].cpu(), blank=0, scale=1.0 / B)
    np.testing.assert_allclose(nll[b].item(), n64[0], rtol=1e-4)
    np.testing.assert_allclose(grad[b].cpu().numpy(), g64[0], rtol=1e-4, atol=1e-6)
