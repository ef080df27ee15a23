"""CPU: host-side logic of the product package (no kernel launches): label glue, batch
sharding and the one-collective gradient exchange (gloo, world_size 2)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import reference as orc


def test_add_blank_and_targets_match_oracle():
    from wenet_celoss_b200 import add_blank, rnnt_targets
    ys = torch.tensor([[1, 2, 3, 4, 5], [4, 5, 6, -1, -1], [7, -1, -1, -1, -1]])
    assert torch.equal(add_blank(ys, 0, -1), orc.add_blank(ys, 0, -1))
    lens, tl = torch.tensor([5, 3, 1]), torch.tensor([9, 8, 7])
    mine, ref = rnnt_targets(ys, lens, tl), orc.rnnt_targets(ys, lens, tl)
    for a, b in zip(mine, ref):
        assert a.dtype == torch.int32 and torch.equal(a, b)


def test_shard_batch():
    from wenet_celoss_b200.dist import shard_batch
    assert shard_batch(10, 1, 4) == [1, 5, 9]
    w = [5, 1, 9, 3, 7, 2, 8, 4]
    parts = [shard_batch(8, r, 2, w) for r in range(2)]
    assert sorted(parts[0] + parts[1]) == list(range(8))
    loads = [sum(w[i] for i in p) for p in parts]
    assert abs(loads[0] - loads[1]) <= max(w)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, xs, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from wenet_celoss_b200.dist import allreduce_step, shard_batch
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 3)
    idx = shard_batch(xs.size(0), rank, world)
    costs = (lin(xs[idx]) ** 2).sum(1)          # one "cost" per utterance
    costs.sum().backward()                      # gradient of the LOCAL sum
    loss, gb = allreduce_step(lin.parameters(), costs.detach().sum(), len(idx))
    out[rank] = (loss.item(), gb, lin.weight.grad.clone(), lin.bias.grad.clone())
    dist.destroy_process_group()


def test_allreduce_step_gloo_world2():
    xs = torch.randn(7, 4, generator=torch.Generator().manual_seed(1))
    mgr = mp.Manager()
    out = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(2, port, xs, out), nprocs=2, join=True)
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 3)
    ref = (lin(xs) ** 2).sum(1).mean()
    ref.backward()
    for r in range(2):
        loss, gb, gw, gbias = out[r]
        assert gb == 7
        np.testing.assert_allclose(loss, ref.item(), rtol=1e-6)
        np.testing.assert_allclose(gw.numpy(), lin.weight.grad.numpy(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(gbias.numpy(), lin.bias.grad.numpy(), rtol=1e-5, atol=1e-7)


def test_wrappers_refuse_cpu_tensors():
    import wenet_celoss_b200 as pkg
    x = torch.randn(1, 2, 2, 5)
    t = torch.ones(1, 1, dtype=torch.int32)
    l1, l2 = torch.tensor([2], dtype=torch.int32), torch.tensor([1], dtype=torch.int32)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.rnnt_loss(x, t, l1, l2, blank=0)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.ctc_loss(torch.randn(1, 3, 5), torch.tensor([[1]]), torch.tensor([3]), torch.tensor([1]))
    m = orc.OracleModel(20, 8, 8, 64, 8, 8, 1)
    with pytest.raises(RuntimeError, match="CUDA"):
        pkg.greedy_search_batch(m, torch.randn(1, 4, 8), torch.tensor([4]))


def test_cached_host_read_identity_and_version():
    """Length-validation reads are remembered per tensor OBJECT and version: an in-place update or a new
    tensor (even one reusing the same storage address) reads again."""
    from wenet_celoss_b200 import _lib
    a, b = torch.tensor([3, 4]), torch.tensor([1, 2])
    calls = []

    def read():
        calls.append(1)
        return [int(a.max()), int(b.max())]

    assert _lib.cached_host_read((a, b), read) == [4, 2]
    assert _lib.cached_host_read((a, b), read) == [4, 2] and len(calls) == 1
    a.add_(1)
    assert _lib.cached_host_read((a, b), read) == [5, 2] and len(calls) == 2
    ptr = a.data_ptr()
    del a
    a = torch.tensor([7, 1])  # may land on the freed address: must not hit
    assert _lib.cached_host_read((a, b), read) == [7, 2] and len(calls) == 3
    _ = ptr
