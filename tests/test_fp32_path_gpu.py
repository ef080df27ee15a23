"""GPU parity of the fp32-accurate (split-bf16) tensor-core path against the reference's fp32 chain.

BASELINE configs[1] runs TransducerJoint.forward + torchaudio.functional.rnnt_loss in fp32
(transducer.py:132-147) and CTC.forward runs ctc_lo + log_softmax + CTCLoss in fp32 (ctc.py:57-61);
north_star's bar for fp32 is rtol 1e-4.  The split path feeds every vocabulary GEMM with bf16 pairs
(x = hi + lo) and three tcgen05 products, see csrc/split_gemm.cu.

Tolerances written here: loss rtol 1e-4; gradients relative L2 and max-abs (relative to the largest
entry) 1e-4 against the host fp32 chain (whose own fp32 summation error is ~1e-6).
"""
import copy
import ctypes

import numpy as np
import pytest
import torch

from oracle import reference as orc
from oracle.c_oracle import ctc_loss_c, rnnt_loss_c
from test_joint_gpu import _case, _close, _mine, _reference

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _i64(v):
    return ctypes.c_longlong(int(v))


def _reference_exact_lattice(joint, enc, pred, tgt, tl, ul):
    """The reference chain with the lattice in fp64: fp32 joint (joint.py:45-70) -> C oracle of rnnt_loss
    (alpha / beta in double) -> autograd through the joint.  At T + U of a few hundred torchaudio's own fp32
    alpha / beta (|alpha| ~ 2000, one ulp = 1.2e-4) put 1e-4 .. 1e-3 of noise on the gradients, more than
    the 1e-4 bar; this chain has none, so it is the anchor for long utterances."""
    j = copy.deepcopy(joint)
    enc = enc.clone().requires_grad_(True)
    pred = pred.clone().requires_grad_(True)
    logits = orc.joint_logits(enc, pred, j)
    costs, grad, _, _ = rnnt_loss_c(logits.detach(), tgt, tl, ul, blank=0)
    logits.backward(torch.from_numpy(grad))
    grads = {"enc": enc.grad, "pred": pred.grad}
    grads.update({n: p.grad for n, p in j.named_parameters()})
    return torch.from_numpy(costs), grads


def _errs(g, ref):
    out = {}
    for k in ref:
        a, b = g[k].double(), ref[k].double()
        out[k] = float((a - b).norm() / b.norm().clamp_min(1e-30))
    return out


def _split_gemm_inputs(rows, K, N, seed):
    g = torch.Generator().manual_seed(seed)
    a = torch.randn(rows, K, generator=g)
    b = torch.randn(N, K, generator=g) * 0.2
    bias = torch.randn(N, generator=g)
    return a.to(DEV), b.to(DEV), bias.to(DEV)


@pytest.mark.parametrize("shape", [(1000, 256, 5000), (300, 512, 4233), (5, 64, 7), (70000, 256, 300)])
@pytest.mark.parametrize("max_ctas", [0, 2])
def test_split_gemm_lse_lsm_exp(shape, max_ctas):
    """x = A B^T + bias through the three epilogues, against fp64 on the same fp32 inputs."""
    from wenet_celoss_b200 import _lib
    from wenet_celoss_b200._lib import i32, ptr
    from wenet_celoss_b200.joint import _pad_bias, split_f32
    rows, K, N = shape
    if rows > 10000 and max_ctas:
        pytest.skip("large case runs on the full grid only")
    a, b, bias = _split_gemm_inputs(rows, K, N, seed=rows + N)
    a_hi, a_lo = split_f32(a)
    b_hi, b_lo = split_f32(b)
    np.testing.assert_allclose((a_hi.float() + a_lo.float()).cpu(), a.cpu(), rtol=2e-5, atol=1e-30)
    bias_pad = _pad_bias(bias, N)
    st = _lib.stream_handle(a.device)
    x = a.double() @ b.double().t() + bias.double()
    want_lse = torch.logsumexp(x, 1)

    lse = torch.empty(rows, dtype=torch.float32, device=DEV)
    _lib.call("b200_split_gemm_lse", ptr(a_hi), ptr(a_lo), _i64(rows), i32(K), i32(K), ptr(b_hi), ptr(b_lo),
              i32(N), i32(K), ptr(bias_pad), ptr(lse), i32(max_ctas), st)
    np.testing.assert_allclose(lse.cpu().numpy(), want_lse.cpu().numpy(), rtol=2e-5, atol=2e-5)

    if N % 4 == 0:
        out = torch.full((rows, N), float("nan"), dtype=torch.float32, device=DEV)
        _lib.call("b200_split_gemm_lsm", ptr(a_hi), ptr(a_lo), _i64(rows), i32(K), i32(K), ptr(b_hi),
                  ptr(b_lo), i32(N), i32(K), ptr(bias_pad), ptr(lse), ptr(out), _i64(N), i32(max_ctas), st)
        want = (x - want_lse[:, None]).float()
        np.testing.assert_allclose(out.cpu().numpy(), want.cpu().numpy(), rtol=1e-5, atol=2e-4)

    ldg = (N + 15) // 16 * 16
    coef = torch.zeros(rows, 4, dtype=torch.float32, device=DEV)
    coef[:, 0] = -want_lse.float()
    coef[:, 3] = torch.linspace(-1.0, 2.0, rows, device=DEV)  # both signs: scale in / out of the exponent
    coef[::7, 3] = 0.0
    g_hi = torch.full((rows, ldg), 7.0, dtype=torch.bfloat16, device=DEV)
    g_lo = torch.full((rows, ldg), 7.0, dtype=torch.bfloat16, device=DEV)
    _lib.call("b200_split_gemm_exp", ptr(a_hi), ptr(a_lo), _i64(rows), i32(K), i32(K), ptr(b_hi), ptr(b_lo),
              i32(N), i32(K), ptr(bias_pad), ptr(coef), ptr(g_hi), ptr(g_lo), _i64(ldg), i32(max_ctas), st)
    got = g_hi[:, :N].double() + g_lo[:, :N].double()
    want = coef[:, 3:4].double() * torch.softmax(x, 1)
    # operands carry 16 mantissa bits (2^-17 relative) and the lo.lo product is dropped: the logits are
    # off by ~3 * 2^-18 * |a||b| sqrt(K) ~ 1e-5 .. 4e-5 in absolute terms, which is the relative error of exp
    err = (got - want).abs().max() / want.abs().max()
    assert err < 1e-4, f"split exp: {err:.3e}"
    assert torch.count_nonzero(got[::7]) == 0

    # backward-side GEMMs on the split G: dW = G^T A, dA = G B, column sums
    if K in (256, 512):
        dW = torch.zeros(N, K, dtype=torch.float32, device=DEV)
        _lib.call("b200_joint_dw_split", ptr(g_hi), ptr(g_lo), i32(ldg), ptr(a_hi), ptr(a_lo), i32(K),
                  _i64(rows), i32(N), i32(K), ptr(dW), i32(max_ctas), st)
        w = got.t() @ a.double()
        err = (dW.double() - w).abs().max() / w.abs().max()
        assert err < 5e-5, f"split dW: {err:.3e}"
        dA = torch.full((rows, K), float("nan"), dtype=torch.float32, device=DEV)
        _lib.call("b200_split_gemm_nn", ptr(g_hi), ptr(g_lo), i32(ldg), ptr(b_hi), ptr(b_lo), _i64(rows),
                  i32(N), i32(K), ptr(dA), i32(max_ctas), st)
        w = got @ b.double()
        err = (dA.double() - w).abs().max() / w.abs().max()
        assert err < 5e-5, f"split dA: {err:.3e}"
    cs = torch.zeros(N, dtype=torch.float32, device=DEV)
    _lib.call("b200_colsum_split", ptr(g_hi), ptr(g_lo), _i64(rows), i32(N), _i64(ldg), ptr(cs), st)
    w = got.sum(0)
    assert (cs.double() - w).abs().max() / w.abs().max() < 3e-5


@pytest.mark.parametrize("shape", [(3, 30, 10, 700, 64, 512), (2, 17, 5, 300, 32, 128),
                                   (4, 40, 12, 4233, 64, 512), (1, 9, 0, 260, 32, 64),
                                   (2, 40, 63, 1100, 64, 256)])
@pytest.mark.parametrize("ragged", [False, True])
@pytest.mark.parametrize("nr", [None, False])
def test_fused_joint_fp32(shape, ragged, nr):
    """Both backward variants of the fp32 mode: the forward GEMM keeps exp(logit) as a split pair
    (default where the own dW / dH kernels apply, J in {256, 512}) / the recompute variant."""
    import wenet_celoss_b200.joint as jm
    B, T, U, V, E, J = shape
    case = _case(B, T, U, V, E, J, seed=V + T, ragged=ragged)
    ref_loss, ref_g = _reference(*case, reduction="none")
    loss, g = _mine(*case, reduction="none", precision="fp32", _no_recompute=nr)
    assert jm.LAST_BACKWARD_PATH == ("no_recompute" if nr is None and J in (256, 512) else "recompute")
    np.testing.assert_allclose(loss.numpy(), ref_loss.numpy(), rtol=1e-4)
    for k in ref_g:
        _close(k, g[k], ref_g[k], tol=1e-4)
    joint, enc, pred, tgt, tl, ul = case
    for b in range(B):
        assert torch.count_nonzero(g["enc"][b, tl[b]:]) == 0
        assert torch.count_nonzero(g["pred"][b, ul[b] + 1:]) == 0


@pytest.mark.parametrize("ragged", [False, True])
def test_cfg2_slice_fp32(ragged):
    """B=2 of BASELINE configs[1]: T=250, U=40, V=4233, J=512, fp32, rtol 1e-4 against the reference chain
    with an exact (fp64) lattice; torchaudio's fp32 lattice is further from it than this path is."""
    case = _case(2, 250, 40, 4233, 256, 512, seed=22, ragged=ragged)
    ref_loss, ref_g = _reference_exact_lattice(*case)
    ta_loss, ta_g = _reference(*case, reduction="none")
    for nr in (None, False):  # no-recompute (default) and recompute backward
        loss, g = _mine(*case, reduction="none", precision="fp32", _no_recompute=nr)
        mine, ta = _errs(g, ref_g), _errs(ta_g, ref_g)
        print("rel L2 vs exact-lattice chain   mine:", {k: f"{v:.1e}" for k, v in mine.items()})
        print("                          torchaudio:", {k: f"{v:.1e}" for k, v in ta.items()})
        np.testing.assert_allclose(loss.numpy(), ref_loss.numpy(), rtol=1e-4)
        np.testing.assert_allclose(loss.numpy(), ta_loss.numpy(), rtol=1e-4)
        for k in ref_g:
            _close(k, g[k], ref_g[k], tol=1e-4)


def test_fp32_auto_and_multichunk(monkeypatch):
    """precision='auto' takes the fp32 path for fp32 inputs outside autocast; chunked backward agrees."""
    import wenet_celoss_b200.joint as jm
    case = _case(5, 30, 9, 600, 64, 512, seed=5, ragged=True)
    l1, g1 = _mine(*case, reduction="none", precision="auto")
    l2, g2 = _mine(*case, reduction="none", precision="fp32")
    assert torch.equal(l1, l2)
    monkeypatch.setattr(jm, "SCRATCH_BYTES", 2 * 2 * 30 * 10 * 608 * 2)  # two utterances per chunk
    l3, g3 = _mine(*case, reduction="none", precision="fp32")
    assert jm.LAST_BACKWARD_PATH == "recompute"  # scratch too small for exp(logit): chunked recompute
    np.testing.assert_allclose(l3.numpy(), l2.numpy(), rtol=1e-6)
    for k in g2:
        _close(k, g3[k], g2[k], tol=5e-5)
    ref_loss, ref_g = _reference(*case, reduction="none")
    for k in ref_g:
        _close(k, g3[k], ref_g[k], tol=1e-4)


def _ctc_case(B, T, E, V, Lmax, seed, repeats=False):
    g = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    hs = torch.randn(B, T, E, generator=g)
    hl = torch.randint(T // 2, T + 1, (B,), generator=g, dtype=torch.int32)
    yl = torch.randint(max(1, Lmax // 2), Lmax + 1, (B,), generator=g, dtype=torch.int32)
    hl[0], yl[-1] = T, Lmax
    ys = torch.randint(1, 5 if repeats else V, (B, Lmax), generator=g, dtype=torch.int64)
    for b in range(B):
        ys[b, yl[b]:] = -1
    return hs, hl, ys, yl


@pytest.mark.parametrize("shape", [(4, 120, 256, 5000, 20), (3, 60, 512, 1000, 12), (2, 50, 256, 37, 9),
                                   (2, 600, 256, 700, 127), (2, 900, 256, 900, 200)])  # label blocks of 127: edge, two
@pytest.mark.parametrize("reduce", [True, False])
def test_ctc_module_fused_fp32(shape, reduce):
    """CTC.forward through split_gemm (no [B,T,V] logits) against the reference module's arithmetic
    in fp32 on the host: ctc_lo -> log_softmax -> CTCLoss(reduction='sum') / B (ctc.py:57-64)."""
    from wenet_celoss_b200 import CTC
    B, T, E, V, Lmax = shape
    hs, hl, ys, yl = _ctc_case(B, T, E, V, Lmax, seed=T + V, repeats=V < 100)
    torch.manual_seed(3)
    mod = CTC(V, E, reduce=reduce)
    with torch.no_grad():
        mod.ctc_lo.weight.mul_(3.0)
    lin = copy.deepcopy(mod.ctc_lo)
    x = hs.clone().requires_grad_(True)
    # the module's arithmetic with the trellis in fp64 (C oracle); torch's CPU CTCLoss agrees with it to
    # ~1e-4 on the loss but its fp32 alpha / beta put up to 2e-4 of noise on the gradients at T > 100
    logits = lin(x)
    nll, grad = ctc_loss_c(logits.detach(), ys, hl, yl, blank=0, scale=1.0 / B)
    logits.backward(torch.from_numpy(grad))
    want = torch.from_numpy(nll.sum(keepdims=True) if reduce else nll).float() / B
    lp = lin(hs).transpose(0, 1).log_softmax(2)
    crit = torch.nn.CTCLoss(reduction="sum" if reduce else "none")
    np.testing.assert_allclose((crit(lp, ys.clamp_min(0), hl, yl) / B).detach().numpy().reshape(-1),
                               want.numpy().reshape(-1), rtol=1e-4)

    m = mod.to(DEV)
    xg = hs.to(DEV).requires_grad_(True)
    got = m(xg, hl.to(DEV), ys.to(DEV), yl.to(DEV))
    got.sum().backward()
    np.testing.assert_allclose(got.detach().cpu().numpy().reshape(-1), want.numpy().reshape(-1), rtol=1e-4)
    _close("hs", xg.grad.cpu(), x.grad, tol=1e-4)
    _close("ctc_lo.weight", m.ctc_lo.weight.grad.cpu(), lin.weight.grad, tol=1e-4)
    _close("ctc_lo.bias", m.ctc_lo.bias.grad.cpu(), lin.bias.grad, tol=1e-4)
    for b in range(B):
        assert torch.count_nonzero(xg.grad[b, hl[b]:]) == 0
    # log_softmax / argmax of the module (ctc.py:66-84)
    with torch.no_grad():
        ls = m.log_softmax(hs.to(DEV)).cpu()
        np.testing.assert_allclose(ls.numpy(), lin(hs).log_softmax(2).detach().numpy(), rtol=1e-4, atol=2e-5)
