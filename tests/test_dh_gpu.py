"""joint_dh_kernel (2-CTA tcgen05: dH = G W, tanh', sums over u / t in the epilogue) against plain torch
on the same bf16 operands, through the C ABI."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _run(nb, T, U1, V, J, seed, lens=None, max_ctas=0):
    from wenet_celoss_b200 import _lib
    from wenet_celoss_b200._lib import i32, ptr
    g = torch.Generator().manual_seed(seed)
    n = nb * T * U1
    ldg = (V + 15) // 16 * 16
    G = torch.zeros(n, ldg, dtype=torch.bfloat16)
    G[:, :V] = (torch.randn(n, V, generator=g) * 0.1).to(torch.bfloat16)
    G[:, V:] = 7.0  # pad columns beyond V must never be read
    H = torch.tanh(torch.randn(n, J, generator=g)).to(torch.bfloat16)
    W = (torch.randn(V, J, generator=g) * 0.05).to(torch.bfloat16)
    tl = ul = None
    if lens is not None:
        tl, ul = lens
        Gv = G.view(nb, T, U1, ldg)
        for b in range(nb):  # the GRAD kernel writes exact zeros beyond the lengths
            Gv[b, tl[b]:] = 0
            Gv[b, :, ul[b] + 1:] = 0
    dz = (G[:, :V].float() @ W.float()) * (1 - H.float() ** 2)
    dz = dz.view(nb, T, U1, J)
    de_ref, dp_ref = dz.sum(2), dz.sum(1)
    Gd, Hd, Wd = G.to(DEV), H.to(DEV), W.to(DEV)
    de = torch.zeros(nb, T, J, device=DEV)
    dp = torch.zeros(nb, U1, J, device=DEV)
    tld = torch.tensor(tl, dtype=torch.int32, device=DEV) if tl is not None else None
    uld = torch.tensor(ul, dtype=torch.int32, device=DEV) if ul is not None else None
    _lib.call("b200_joint_dh", ptr(Gd), i32(ldg), ptr(Wd), ptr(Hd), i32(J), i32(nb), i32(T), i32(U1), i32(V),
              i32(J), ptr(tld), ptr(uld), ptr(de), ptr(dp), i32(max_ctas), _lib.stream_handle(torch.device(DEV)))
    torch.cuda.synchronize()
    for name, got, want in (("de", de.cpu(), de_ref), ("dp", dp.cpu(), dp_ref)):
        err = (got - want).abs().max() / want.abs().max()
        assert err < 2e-5, (name, float(err))


@pytest.mark.parametrize("shape", [
    (2, 40, 11, 700, 512),    # UB = 4 / 2 class shapes, vocabulary tail of 60
    (1, 70, 41, 300, 512),    # cfg2's U1 = 41
    (2, 33, 101, 260, 512),   # cfg3's U1 = 101, T not a multiple of the frame block
    (3, 9, 1, 128, 256),      # U = 0, J = 256 (one accumulator)
    (1, 300, 16, 1000, 256),
    (1, 20, 64, 520, 512),    # UB = 32
])
def test_dh_matches_torch(shape):
    _run(*shape, seed=sum(shape))


def test_dh_many_tiles_per_pair_and_ragged():
    _run(3, 64, 24, 600, 512, seed=5, max_ctas=2)
    _run(3, 64, 24, 600, 512, seed=6, lens=([64, 40, 17], [23, 9, 15]), max_ctas=4)
    _run(2, 250, 41, 4233, 512, seed=7, lens=([250, 140], [20, 40]))
