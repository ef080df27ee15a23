"""GPU parity: fused joint + RNN-T loss (tcgen05 kernels through the C ABI) vs the oracle.

The fused path computes the two big GEMMs with bf16 operands (fp32 accumulate), so the
bar is north_star's bf16 one: loss rtol 1e-2, gradients within 1e-2 (relative L2 and
max-abs relative to the largest entry) of the fp32 reference chain
TransducerJoint.forward + torchaudio.functional.rnnt_loss (transducer.py:132-147).
"""
import copy

import numpy as np
import pytest
import torch
import torchaudio

from oracle import reference as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(B, T, U, V, E, J, seed, ragged=True):
    g = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    joint = orc.OracleJoint(V, E, E, J)
    with torch.no_grad():
        joint.ffn_out.weight.mul_(2.0)  # logits with some spread
    enc = torch.randn(B, T, E, generator=g)
    pred = torch.randn(B, U + 1, E, generator=g)
    tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32)
    if ragged:
        tl = torch.randint(max(1, T // 2), T + 1, (B,), generator=g, dtype=torch.int32)
        ul = torch.randint(U // 2, U + 1, (B,), generator=g, dtype=torch.int32)
        tl[0], ul[-1] = T, U
    else:
        tl = torch.full((B,), T, dtype=torch.int32)
        ul = torch.full((B,), U, dtype=torch.int32)
    return joint, enc, pred, tgt, tl, ul


def _round_bf16_st(x):
    """x rounded to bf16 in the forward pass, identity in the backward pass (what a bf16 MMA operand is)."""
    return x + (x.detach().bfloat16().float() - x.detach())


def _reference(joint, enc, pred, tgt, tl, ul, reduction="mean", bf16_operands=False):
    """The reference chain TransducerJoint.forward + torchaudio rnnt_loss in fp32 on the host.
    ``bf16_operands``: the same chain with the two operands of the vocabulary GEMM (tanh output and
    ffn_out.weight) rounded to bf16, fp32 accumulation - the arithmetic the fused bf16 mode specifies."""
    j = copy.deepcopy(joint)
    enc = enc.clone().requires_grad_(True)
    pred = pred.clone().requires_grad_(True)
    if bf16_operands:
        h = torch.tanh(j.enc_ffn(enc)[:, :, None] + j.pred_ffn(pred)[:, None])
        logits = _round_bf16_st(h) @ _round_bf16_st(j.ffn_out.weight).t() + j.ffn_out.bias
    else:
        logits = orc.joint_logits(enc, pred, j)
    loss = torchaudio.functional.rnnt_loss(logits, tgt, tl, ul, blank=0, reduction=reduction)
    del logits
    loss.sum().backward()
    grads = {"enc": enc.grad, "pred": pred.grad}
    grads.update({n: p.grad for n, p in j.named_parameters()})
    return loss.detach(), grads


def _mine(joint, enc, pred, tgt, tl, ul, reduction="mean", precision="bf16", **kw):
    from wenet_celoss_b200 import fused_joint_rnnt_loss
    j = copy.deepcopy(joint).to(DEV)
    enc = enc.to(DEV).requires_grad_(True)
    pred = pred.to(DEV).requires_grad_(True)
    loss = fused_joint_rnnt_loss(j, enc, pred, tgt.to(DEV), tl.to(DEV), ul.to(DEV), blank=0,
                                 reduction=reduction, precision=precision, **kw)
    loss.sum().backward()
    grads = {"enc": enc.grad.cpu(), "pred": pred.grad.cpu()}
    grads.update({n: p.grad.cpu() for n, p in j.named_parameters()})
    return loss.detach().cpu(), grads


def _close(name, a, b, tol=1e-2):
    a, b = a.double(), b.double()
    rel_l2 = (a - b).norm() / b.norm().clamp_min(1e-30)
    rel_max = (a - b).abs().max() / b.abs().max().clamp_min(1e-30)
    assert rel_l2 < tol and rel_max < 2 * tol, f"{name}: rel_l2 {rel_l2:.3e} rel_max {rel_max:.3e}"


@pytest.mark.parametrize("shape", [(3, 30, 10, 700, 64, 512), (2, 17, 5, 300, 32, 128),
                                   (4, 40, 12, 4233, 64, 512), (1, 9, 0, 260, 32, 64)])
@pytest.mark.parametrize("ragged", [False, True])
def test_fused_matches_reference_chain(shape, ragged):
    B, T, U, V, E, J = shape
    case = _case(B, T, U, V, E, J, seed=V + T, ragged=ragged)
    ref_loss, ref_g = _reference(*case, reduction="none")
    loss, g = _mine(*case, reduction="none")
    np.testing.assert_allclose(loss.numpy(), ref_loss.numpy(), rtol=1e-2)
    for k in ref_g:
        _close(k, g[k], ref_g[k])
    joint, enc, pred, tgt, tl, ul = case
    for b in range(B):  # padded frames / labels receive exactly zero gradient
        assert torch.count_nonzero(g["enc"][b, tl[b]:]) == 0
        assert torch.count_nonzero(g["pred"][b, ul[b] + 1:]) == 0


def test_reductions_and_errors():
    from wenet_celoss_b200 import fused_joint_rnnt_loss
    joint, enc, pred, tgt, tl, ul = _case(3, 12, 4, 300, 32, 64, seed=2)
    j = joint.to(DEV)
    a = (enc.to(DEV), pred.to(DEV), tgt.to(DEV), tl.to(DEV), ul.to(DEV))
    none = fused_joint_rnnt_loss(j, *a, blank=0, reduction="none")
    assert torch.allclose(fused_joint_rnnt_loss(j, *a, blank=0, reduction="mean"), none.mean())
    assert torch.allclose(fused_joint_rnnt_loss(j, *a, blank=0, reduction="sum"), none.sum())
    with pytest.raises(RuntimeError, match="targets must be int32"):
        fused_joint_rnnt_loss(j, a[0], a[1], a[2].long(), a[3], a[4])
    with pytest.raises(RuntimeError, match="input length mismatch"):
        fused_joint_rnnt_loss(j, a[0], a[1], a[2], a[3] - 1, a[4])
    bad = orc.OracleJoint(300, 32, 32, 24).to(DEV)  # join dim not a multiple of 64
    with pytest.raises(NotImplementedError, match="join_dim"):
        fused_joint_rnnt_loss(bad, *a)


def test_matches_materialised_path_on_same_bf16_operands():
    """Same H (bf16) and W (bf16): the fused lse / log-probs equal the materialised ones."""
    from wenet_celoss_b200 import rnnt_loss, fused_joint_rnnt_loss
    joint, enc, pred, tgt, tl, ul = _case(2, 25, 8, 1000, 64, 512, seed=9)
    j = joint.to(DEV)
    enc, pred, tgt, tl, ul = (x.to(DEV) for x in (enc, pred, tgt, tl, ul))
    fused = fused_joint_rnnt_loss(j, enc, pred, tgt, tl, ul, blank=0, reduction="none", precision="bf16")
    h = torch.tanh(j.enc_ffn(enc)[:, :, None] + j.pred_ffn(pred)[:, None]).to(torch.bfloat16).float()
    logits = h @ j.ffn_out.weight.to(torch.bfloat16).float().t() + j.ffn_out.bias
    mat = rnnt_loss(logits.contiguous(), tgt, tl, ul, blank=0, reduction="none")
    np.testing.assert_allclose(fused.detach().cpu().numpy(), mat.detach().cpu().numpy(), rtol=2e-3)
