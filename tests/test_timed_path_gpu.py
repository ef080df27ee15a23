"""GPU parity on the code paths bench.py actually times ("next round" item 1).

The persistent joint / dW / dH kernels walk many 128-cell tiles per CTA at cfg3; the small
shapes of test_joint_gpu.py never make a CTA process a second tile.  Here:

  * ``_max_ctas`` caps the persistent grids so that every CTA walks >= 10 tiles (A-staging
    phase flips, TMEM accumulator phase wrap, H-store drain, dW K-split units, dH tile ranges);
  * cfg3- and cfg2-shaped slices (B=2, full T/U/V/J) run against the fp32 reference chain
    TransducerJoint.forward + torchaudio.functional.rnnt_loss (transducer.py:132-147) on the host;
  * the running-maximum epilogue (taken when max_v(|W_v|_1 + |b_v|) * log2(e) >= 100, or when no
    bound is supplied) is forced;
  * the multi-chunk backward loop is forced through a small SCRATCH_BYTES;
  * BOTH backward variants run on every shape: the no-recompute path (the forward kernel keeps
    exp(logit) in bf16; default when the scratch fits and the logit bound is small - what bench.py
    times) and the recompute path (``_no_recompute=False``; also the fallback for large bounds and
    chunked scratch);
  * greedy and beam-10 decode bench.py's own cfg4 model (V=5000, T=500) against the oracle.

Tolerance: bf16 MMA operands -> north_star's bf16 bar, rtol 1e-2 (relative L2 and max-abs
relative to the largest entry for gradients).
"""
import copy

import numpy as np
import pytest
import torch
import torchaudio

from oracle import reference as orc
from test_joint_gpu import _case, _close, _mine, _reference

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _mine_kw(case, reduction="none", **kw):
    from wenet_celoss_b200 import fused_joint_rnnt_loss
    joint, enc, pred, tgt, tl, ul = case
    kw.setdefault("precision", "bf16")
    j = copy.deepcopy(joint).to(DEV)
    enc = enc.to(DEV).requires_grad_(True)
    pred = pred.to(DEV).requires_grad_(True)
    loss = fused_joint_rnnt_loss(j, enc, pred, tgt.to(DEV), tl.to(DEV), ul.to(DEV), blank=0,
                                 reduction=reduction, **kw)
    loss.sum().backward()
    grads = {"enc": enc.grad.cpu(), "pred": pred.grad.cpu()}
    grads.update({n: p.grad.cpu() for n, p in j.named_parameters()})
    return loss.detach().cpu(), grads


def _check(case, tol_inputs=1e-2, expect_nr=True, **kw):
    """Both backward variants against the fp32 reference chain (``expect_nr``: the default variant
    must really have been the no-recompute one)."""
    from wenet_celoss_b200 import joint as jmod
    ref_loss, ref_g = _reference(*case, reduction="none")
    out = None
    for nr in (None, False):
        loss, g = _mine_kw(case, _no_recompute=nr, **kw)
        want_path = "no_recompute" if (nr is None and expect_nr) else "recompute"
        assert jmod.LAST_BACKWARD_PATH == want_path, (nr, jmod.LAST_BACKWARD_PATH)
        np.testing.assert_allclose(loss.numpy(), ref_loss.numpy(), rtol=1e-2)
        for k in ref_g:
            _close(k, g[k], ref_g[k], tol=1e-2 if k.startswith("ffn_out") else tol_inputs)
        _, _, _, _, tl, ul = case
        for b in range(len(tl)):  # padded frames / labels receive exactly zero gradient
            assert torch.count_nonzero(g["enc"][b, tl[b]:]) == 0
            assert torch.count_nonzero(g["pred"][b, ul[b] + 1:]) == 0
        out = out or (loss, g)
        if not expect_nr:
            break
    return out


@pytest.mark.parametrize("max_ctas", [2, 4])
@pytest.mark.parametrize("ragged", [False, True])
def test_many_tiles_per_cta(max_ctas, ragged):
    """9000 cells = 71 tiles of 128 over 2 / 4 persistent CTAs (>= 17 tiles each); the dW kernel
    runs 3 V tiles x 2 K splits = 3 units per pair with max_ctas = 4."""
    case = _case(3, 50, 59, 700, 64, 512, seed=11, ragged=ragged)
    _check(case, _max_ctas=max_ctas)


def test_many_tiles_j256():
    case = _case(2, 40, 63, 1100, 64, 256, seed=12, ragged=True)
    _check(case, _max_ctas=2)


@pytest.mark.parametrize("ragged", [False, True])
def test_cfg3_slice(ragged):
    """B=2 of the headline shape: T=500, U=100, V=5000, E=P=256, J=512 (2 GB of fp32 logits on the
    host for the reference chain).  790 tiles over 148 CTAs: 5-6 tiles per CTA, last wave partial.

    Two references.  (1) The reference chain with the vocabulary GEMM's operands rounded to bf16 (the
    arithmetic precision='bf16' specifies; fp32 everywhere else): every gradient within 1e-2 - this is
    the check of the kernels.  (2) The plain fp32 chain: loss within 1e-2, gradients within 2e-2.  At
    this lattice size the bf16 rounding of the logits (1e-3 absolute per cell) accumulates along the 600
    steps of a path into ~1 % errors of the cell occupancies alpha*beta/Z, which every gradient inherits
    (measured 0.8-1.3e-2 relative L2; scripts/dbg_prec.py traces it) - a property of bf16 operands at
    T+U = 600, shared by the reference under autocast, not of these kernels."""
    from wenet_celoss_b200 import joint as jmod
    case = _case(2, 500, 100, 5000, 256, 512, seed=13, ragged=ragged)
    loss, g = _mine_kw(case)
    assert jmod.LAST_BACKWARD_PATH == "no_recompute"
    loss_rc, g_rc = _mine_kw(case, _no_recompute=False)
    assert jmod.LAST_BACKWARD_PATH == "recompute"
    emu_loss, emu_g = _reference(*case, reduction="none", bf16_operands=True)
    for l_, g_ in ((loss, g), (loss_rc, g_rc)):
        np.testing.assert_allclose(l_.numpy(), emu_loss.numpy(), rtol=1e-3)
        for k in emu_g:
            _close(k, g_[k], emu_g[k], tol=1e-2)
    del emu_g
    ref_loss, ref_g = _reference(*case, reduction="none")
    _, _, _, _, tl, ul = case
    for l_, g_ in ((loss, g), (loss_rc, g_rc)):
        np.testing.assert_allclose(l_.numpy(), ref_loss.numpy(), rtol=1e-2)
        for k in ref_g:
            _close(k, g_[k], ref_g[k], tol=2e-2)
        for b in range(len(tl)):
            assert torch.count_nonzero(g_["enc"][b, tl[b]:]) == 0
            assert torch.count_nonzero(g_["pred"][b, ul[b] + 1:]) == 0


def test_cfg2_slice():
    """B=2 of cfg2: T=250, U=40, V=4233 (vocabulary tail tile of 9 classes), J=512."""
    case = _case(2, 250, 40, 4233, 256, 512, seed=14, ragged=True)
    _check(case)


def test_running_max_epilogue_large_weights():
    """ffn_out scaled until max_v(|W_v|_1 + |b_v|) * log2(e) > 100: the forward epilogue must take
    the running-maximum branch by itself."""
    from wenet_celoss_b200.joint import _pad_bias, joint_logit_bound
    case = _case(3, 30, 10, 700, 64, 512, seed=15, ragged=True)
    joint = case[0]
    with torch.no_grad():
        joint.ffn_out.weight.mul_(5.0)  # x2 in _case -> x10 of the default init
    w = joint.ffn_out.weight.to(DEV).to(torch.bfloat16).contiguous()
    bound = joint_logit_bound(w, _pad_bias(joint.ffn_out.bias.to(DEV), 700), 700)
    assert float(bound) * 1.4427 > 100.0, float(bound)
    _check(case, expect_nr=False)  # large bound: the recompute path is chosen by itself
    _check(case, expect_nr=False, _max_ctas=3)


def test_bound_between_no_recompute_and_running_max():
    """60 <= bound * log2(e) < 100: the training forward kernel (joint_kernel<FWDP>) keeps the
    maximum-free statistics but must NOT hand exp(logit) to the backward pass - the recompute
    variant is chosen (on the device and, through the asynchronous flag, on the host)."""
    from wenet_celoss_b200.joint import _pad_bias, joint_logit_bound
    case = _case(3, 30, 10, 700, 64, 512, seed=18, ragged=True)
    joint = case[0]
    with torch.no_grad():
        joint.ffn_out.weight.mul_(2.2)  # x2 in _case (bound * log2 e ~ 35) -> ~ 78
    w = joint.ffn_out.weight.to(DEV).to(torch.bfloat16).contiguous()
    bound = joint_logit_bound(w, _pad_bias(joint.ffn_out.bias.to(DEV), 700), 700)
    assert 60.0 < float(bound) * 1.4427 < 100.0, float(bound) * 1.4427
    _check(case, expect_nr=False)
    _check(case, expect_nr=False, _max_ctas=3)


@pytest.mark.parametrize("shape", [(2, 9, 0, 300, 32, 256), (3, 1, 4, 300, 32, 512), (1, 130, 1, 129, 32, 256)])
def test_edge_shapes_both_backward_variants(shape):
    """No labels at all (U = 0: the lattice is one column of blanks), a single frame, a vocabulary of one
    tile + 1 class, on the join dims the no-recompute backward covers."""
    B, T, U, V, E, J = shape
    case = _case(B, T, U, V, E, J, seed=T + V, ragged=U > 1)
    _check(case)


def test_weighted_costs_negative_and_zero_upstream_gradient():
    """d loss / d cost_b = w_b with a negative, a zero and a large weight: the per-cell factor of the
    no-recompute backward (rowscale = w_b * exp(c0)) carries sign and zero; the recompute kernel's
    'all factors positive' shortcut must not be taken."""
    from wenet_celoss_b200 import fused_joint_rnnt_loss
    from wenet_celoss_b200 import joint as jmod
    case = _case(4, 28, 9, 700, 64, 512, seed=21, ragged=True)
    joint, enc, pred, tgt, tl, ul = case
    w = torch.tensor([1.5, -0.75, 0.0, 3.0])
    j = copy.deepcopy(joint)
    er, pr = enc.clone().requires_grad_(True), pred.clone().requires_grad_(True)
    ref = torchaudio.functional.rnnt_loss(orc.joint_logits(er, pr, j), tgt, tl, ul, blank=0, reduction="none")
    (ref * w).sum().backward()
    ref_g = {"enc": er.grad, "pred": pr.grad, **{n: p.grad for n, p in j.named_parameters()}}
    for nr in (None, False):
        jd = copy.deepcopy(joint).to(DEV)
        ed, pd = enc.to(DEV).requires_grad_(True), pred.to(DEV).requires_grad_(True)
        costs = fused_joint_rnnt_loss(jd, ed, pd, tgt.to(DEV), tl.to(DEV), ul.to(DEV), blank=0, reduction="none",
                                      precision="bf16", _no_recompute=nr)
        (costs * w.to(DEV)).sum().backward()
        assert jmod.LAST_BACKWARD_PATH == ("no_recompute" if nr is None else "recompute")
        g = {"enc": ed.grad.cpu(), "pred": pd.grad.cpu(), **{n: p.grad.cpu() for n, p in jd.named_parameters()}}
        for k in ref_g:
            _close(k, g[k], ref_g[k], tol=1e-2)
        assert torch.count_nonzero(g["enc"][2]) == 0 and torch.count_nonzero(g["pred"][2]) == 0  # w = 0


def test_target_symbol_equal_to_blank():
    """Labels equal to the blank index: the emit and blank terms of d loss / d logits fall on the same
    column (torchaudio's if / else-if order, rnnt_loss.cu); both backward variants patch it once."""
    case = list(_case(3, 26, 9, 700, 64, 512, seed=19, ragged=True))
    tgt = case[3].clone()
    tgt[0, 2] = 0
    tgt[1, 0] = 0
    tgt[2, 5] = 0
    case[3] = tgt
    _check(tuple(case))


def test_running_max_epilogue_no_bound():
    """use_bound=False: same inputs as the fast path, results must agree with it and the reference."""
    case = _case(3, 30, 10, 700, 64, 512, seed=16, ragged=True)
    slow, gs = _check(case, expect_nr=False, _use_bound=False)
    fast, gf = _mine_kw(case, _no_recompute=False)  # same backward kernels, bounded-logit forward epilogue
    np.testing.assert_allclose(slow.numpy(), fast.numpy(), rtol=1e-5)
    for k in gs:
        _close(k, gs[k], gf[k], tol=1e-3)


def test_multi_chunk_backward(monkeypatch):
    """SCRATCH_BYTES forced small: 5 utterances in chunks of 2 + a ragged last chunk of 1;
    dW, d b_out and d p accumulate across chunks."""
    from wenet_celoss_b200 import joint as jmod
    case = _case(5, 24, 9, 700, 64, 512, seed=17, ragged=True)
    per_utt, ldg = 24 * 10, 704
    monkeypatch.setattr(jmod, "SCRATCH_BYTES", 2 * per_utt * ldg * 2 + 64)
    assert jmod._chunk_utts(5, per_utt, ldg, 148) <= 2
    _check(case, expect_nr=False)  # scratch too small for exp(logit) of the whole batch
    _check(case, expect_nr=False, _max_ctas=2)


# ----------------------------------------------------------------------------- cfg4 searches
def _bench_search_model():
    """bench.py's own cfg4 model and an OracleModel carrying the same weights."""
    import bench
    torch.manual_seed(4321)
    cfg = dict(bench.CFG3)
    m = bench._SearchModel(cfg).eval()
    om = orc.OracleModel(cfg["V"], cfg["E"], cfg["P"], cfg["J"], 256, 256, 2, blank=0).eval()
    missing = om.load_state_dict(m.state_dict(), strict=False)
    assert not missing.unexpected_keys and not missing.missing_keys, missing
    return cfg, m, om


def test_greedy_cfg4_shape_vs_oracle():
    """V=5000, T=500 (the bench's decode), 4 utterances: token lists identical.  An utterance whose
    oracle run saw a top-2 margin below 1e-5 is reported, not failed (sub-ulp tie)."""
    from wenet_celoss_b200 import greedy_search_batch
    cfg, m, om = _bench_search_model()
    g = torch.Generator().manual_seed(4321)
    B, T = 6, 500
    enc = torch.randn(B, T, cfg["E"], generator=g) * 2.0
    lens = torch.tensor([500, 500, 431, 377, 500, 266], dtype=torch.int32)
    got = greedy_search_batch(m.to(DEV), enc.to(DEV), lens.to(DEV), n_steps=64)
    ties, emitted = [], 0
    for b in range(B):
        want, margin = orc.greedy_search(om, enc[b:b + 1, :lens[b]], int(lens[b]), n_steps=64,
                                         record_margin=True)
        emitted += len(want)
        if got[b] != want:
            if margin < 1e-5:
                ties.append((b, margin))
                continue
            raise AssertionError(f"utt {b}: token lists differ (min top-2 margin {margin:.3e})")
    assert len(ties) <= 1, ties
    assert emitted > 50  # the synthetic model really emits symbols (bench: ~0.08 per frame)


def test_beam10_cfg4_shape_vs_oracle():
    """beam 10, V=5000, T=500, 2 utterances: n-best token lists identical, scores rtol 1e-5."""
    from wenet_celoss_b200 import prefix_beam_search_batch
    cfg, m, om = _bench_search_model()
    g = torch.Generator().manual_seed(99)
    B, T = 2, 500
    enc = torch.randn(B, T, cfg["E"], generator=g) * 2.0
    lens = torch.tensor([500, 389], dtype=torch.int32)
    beams = prefix_beam_search_batch(m.to(DEV), enc.to(DEV), lens.to(DEV), beam_size=10)
    for b in range(B):
        want = orc.prefix_beam_search(om, enc[b:b + 1, :lens[b]], beam_size=10)
        got_h, want_h = [s.hyp for s in beams[b]], [s.hyp for s in want]
        if got_h != want_h:
            # n-best ORDER may differ only between hypotheses whose scores tie to fp32 rounding
            ws = {tuple(s.hyp): s.score for s in want}
            assert sorted(map(tuple, got_h)) == sorted(map(tuple, want_h)), b
            for s in beams[b]:
                np.testing.assert_allclose(s.score, ws[tuple(s.hyp)], rtol=1e-5)
            continue
        np.testing.assert_allclose([s.score for s in beams[b]], [s.score for s in want], rtol=1e-5)
        assert len(want[0].hyp) > 3


@pytest.mark.parametrize("flat,weights", [(False, (0.3, 0.7)), (True, (0.3, 0.7)), (False, (0.0, 1.0)),
                                          (False, (1.0, 0.0))])
def test_beam_bounded_topk_equals_full_evaluation(flat, weights, monkeypatch):
    """The beam's top-k evaluates the fused score only where a cheap upper bound reaches the K-th best
    sampled score (csrc/beam_select.inc); B200_BEAM_TOPK_FULL=1 evaluates every vocabulary entry.  Both
    must give the same n-best lists and bit-identical scores: on the bench's peaked model, on a flat one
    (every entry survives the bound: more than the survivor cap, i.e. the fallback inside the default
    path), and with one of the two weights zero (log 0 in the bound)."""
    from wenet_celoss_b200 import prefix_beam_search_batch
    cfg, m, _ = _bench_search_model()
    if flat:
        with torch.no_grad():
            m.joint.ffn_out.weight.mul_(1e-3)
            m.joint.ffn_out.bias.zero_()
    g = torch.Generator().manual_seed(7)
    B, T = 3, 60
    enc = (torch.randn(B, T, cfg["E"], generator=g) * 2.0).to(DEV)
    lens = torch.tensor([60, 47, 60], dtype=torch.int32, device=DEV)
    m = m.to(DEV)
    cw, tw = weights
    out = {}
    for full in ("0", "1"):
        monkeypatch.setenv("B200_BEAM_TOPK_FULL", full)
        beams = prefix_beam_search_batch(m, enc, lens, beam_size=10, ctc_weight=cw, transducer_weight=tw)
        out[full] = [[(tuple(s.hyp), s.score) for s in bb] for bb in beams]
    assert out["0"] == out["1"]
    assert all(len(bb) > 0 for bb in out["0"])
