"""b200_ctc_forced_align_static: the allocation-free variant of the forced alignment (back-pointer scratch
in a static device array of 32 MiB), called through the C ABI and compared with b200_ctc_forced_align
(caller workspace) - integer outputs, so bit for bit - on two calls in a row that share the array.
Collected last on purpose (file name).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(B, T, V, Lmax, seed):
    g = torch.Generator().manual_seed(seed)
    lp = torch.log_softmax(torch.randn(B, T, V, generator=g) * 2.0, dim=2)
    hl = torch.randint(max(1, T // 2), T + 1, (B,), generator=g, dtype=torch.int32)
    hl[0] = T
    yl = torch.randint(1, Lmax + 1, (B,), generator=g, dtype=torch.int32)
    yl[-1] = Lmax
    ys = torch.randint(1, V, (B, Lmax), generator=g, dtype=torch.int64)
    if Lmax > 1:
        ys[0, 1] = ys[0, 0]  # a repeated label
        yl[0] = max(int(yl[0]), 2)
    return lp.to(DEV), ys.to(DEV), hl.to(DEV), yl.to(DEV)


def _both(lp, ys, hl, yl, blank=0):
    """(alignment of the workspace form, alignment of the static form); all buffers exist before the calls."""
    from wenet_celoss_b200 import _lib
    from wenet_celoss_b200._lib import i32, ptr
    B, T, V = lp.shape
    Lmax = ys.size(1)
    L = _lib.lib()
    L.b200_ctc_forced_align_workspace_bytes.restype = ctypes.c_size_t
    ws_bytes = L.b200_ctc_forced_align_workspace_bytes(B, T, Lmax)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)
    want = torch.full((B, T), -7, dtype=torch.int32, device=DEV)
    got = torch.full((B, T), -9, dtype=torch.int32, device=DEV)
    with torch.cuda.device(lp.device):
        st = _lib.stream_handle(lp.device)
        _lib.call("b200_ctc_forced_align", ptr(lp), ptr(ys), ptr(hl), ptr(yl), i32(B), i32(T), i32(V), i32(Lmax),
                  i32(blank), ptr(want), ptr(ws), ctypes.c_size_t(ws_bytes), st)
        before = torch.cuda.memory_allocated(lp.device)
        _lib.call("b200_ctc_forced_align_static", ptr(lp), ptr(ys), ptr(hl), ptr(yl), i32(B), i32(T), i32(V),
                  i32(Lmax), i32(blank), ptr(got), st)
        assert torch.cuda.memory_allocated(lp.device) == before
        torch.cuda.synchronize()
    return want.cpu(), got.cpu()


def test_static_scratch_matches_workspace_form_twice():
    # first call: cfg5-like width (201 states per frame), 8.0 MB of back-pointers
    want, got = _both(*_case(20, 1000, 500, 100, seed=11))
    assert torch.equal(want, got)
    assert int((got >= 0).sum()) > 0
    # second call, smaller and differently shaped rows: reads the array the first call left behind
    want, got = _both(*_case(5, 77, 64, 9, seed=12))
    assert torch.equal(want, got)
    # and the first shape again, other data
    want, got = _both(*_case(20, 1000, 500, 100, seed=13))
    assert torch.equal(want, got)


def test_static_scratch_limit_is_enforced():
    from wenet_celoss_b200 import _lib
    from wenet_celoss_b200._lib import i32, ptr
    L = _lib.lib()
    L.b200_ctc_forced_align_static_max_bytes.restype = ctypes.c_size_t
    L.b200_ctc_forced_align_workspace_bytes.restype = ctypes.c_size_t
    limit = L.b200_ctc_forced_align_static_max_bytes()
    assert limit == 32 << 20
    B, T, V, Lmax = 128, 1000, 8, 100  # 51.5 MB of back-pointers
    assert L.b200_ctc_forced_align_workspace_bytes(B, T, Lmax) > limit
    lp = torch.zeros(B, T, V, device=DEV)
    ys = torch.ones(B, Lmax, dtype=torch.int64, device=DEV)
    hl = torch.full((B,), T, dtype=torch.int32, device=DEV)
    yl = torch.full((B,), Lmax, dtype=torch.int32, device=DEV)
    out = torch.empty(B, T, dtype=torch.int32, device=DEV)
    with pytest.raises(_lib.B200Error, match="exceed the static scratch"):
        _lib.call("b200_ctc_forced_align_static", ptr(lp), ptr(ys), ptr(hl), ptr(yl), i32(B), i32(T), i32(V),
                  i32(Lmax), i32(0), ptr(out), _lib.stream_handle(lp.device))
