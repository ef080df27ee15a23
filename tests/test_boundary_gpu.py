"""Boundary behaviour the reference's callers rely on: torch.autocast around the model call
(wenet/utils/executor.py:91,97-103) and DistributedDataParallel (wenet/bin/train.py:141,231)."""
import copy
import os

import numpy as np
import pytest
import torch

from oracle import reference as orc
from test_joint_gpu import _case, _close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("amp_dtype", [torch.bfloat16, torch.float16])
def test_fused_joint_under_autocast(amp_dtype):
    """custom_fwd(cast_inputs=float32): the Linear layers run in the autocast dtype, the fused kernels get
    fp32 e / p; parameter gradients come back fp32 and close to the run without autocast."""
    from wenet_celoss_b200 import fused_joint_rnnt_loss
    joint, enc, pred, tgt, tl, ul = _case(3, 30, 10, 700, 64, 512, seed=31)
    a = (enc.to(DEV), pred.to(DEV), tgt.to(DEV), tl.to(DEV), ul.to(DEV))
    j0, j1 = copy.deepcopy(joint).to(DEV), copy.deepcopy(joint).to(DEV)
    l0 = fused_joint_rnnt_loss(j0, *a, blank=0, reduction="mean")
    l0.backward()
    with torch.autocast("cuda", dtype=amp_dtype):
        l1 = fused_joint_rnnt_loss(j1, *a, blank=0, reduction="mean")
    assert l1.dtype == torch.float32
    l1.backward()
    np.testing.assert_allclose(l1.item(), l0.item(), rtol=2e-2)
    for (n, p0), (_, p1) in zip(j0.named_parameters(), j1.named_parameters()):
        assert p1.grad.dtype == torch.float32
        _close(n, p1.grad.cpu(), p0.grad.cpu(), tol=5e-2)


def test_losses_under_autocast():
    from wenet_celoss_b200 import CTC, rnnt_loss
    g = torch.Generator().manual_seed(3)
    B, T, U, V = 2, 12, 4, 50
    logits = torch.randn(B, T, U + 1, V, generator=g).to(DEV).requires_grad_(True)
    tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(DEV)
    tl = torch.tensor([T, T - 3], dtype=torch.int32, device=DEV)
    ul = torch.tensor([U, U - 1], dtype=torch.int32, device=DEV)
    ref = rnnt_loss(logits, tgt, tl, ul, blank=0, reduction="mean")
    with torch.autocast("cuda", dtype=torch.bfloat16):
        got = rnnt_loss(logits, tgt, tl, ul, blank=0, reduction="mean")
    np.testing.assert_allclose(got.item(), ref.item(), rtol=1e-6)
    torch.manual_seed(0)
    ctc = CTC(V, 16).to(DEV)
    hs = torch.randn(B, T, 16, generator=g).to(DEV)
    ys = torch.tensor([[3, 4, 5, -1], [7, 7, 2, 9]], device=DEV)
    yl = torch.tensor([3, 4], device=DEV)
    ref = ctc(hs, tl, ys, yl)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        got = ctc(hs, tl, ys, yl)   # ctc_lo runs in bf16, the loss kernels read bf16 logits
    np.testing.assert_allclose(got.item(), ref.item(), rtol=3e-2)
    got.backward()
    assert ctc.ctc_lo.weight.grad.dtype == torch.float32


def _ddp_worker(rank, world, port, q):
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    from wenet_celoss_b200 import fused_joint_rnnt_loss
    from wenet_celoss_b200.dist import allreduce_step, shard_batch
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    joint, enc, pred, tgt, tl, ul = _case(4, 20, 6, 300, 32, 128, seed=41, ragged=False)
    idx = shard_batch(4, rank, world)
    a = [x[idx].to(dev) for x in (enc, pred, tgt, tl, ul)]

    class M(torch.nn.Module):  # DDP needs a Module whose forward returns the loss
        def __init__(self, j):
            super().__init__()
            self.joint = j

        def forward(self, e, p, t, l1, l2):
            return fused_joint_rnnt_loss(self.joint, e, p, t, l1, l2, blank=0, reduction="mean")
    ddp = DDP(M(copy.deepcopy(joint).to(dev)), device_ids=[rank], find_unused_parameters=True)
    ddp(*a).backward()  # DDP averages the rank means = global mean (equal shards)
    g_ddp = {n: p.grad.detach().cpu() for n, p in ddp.module.joint.named_parameters()}
    # the package's own exchange: gradient of the local SUM, one flat all-reduce, global mean
    j2 = copy.deepcopy(joint).to(dev)
    cost = fused_joint_rnnt_loss(j2, *a, blank=0, reduction="sum")
    cost.backward()
    loss, gb = allreduce_step(j2.parameters(), cost.detach(), len(idx))
    g_own = {n: p.grad.detach().cpu() for n, p in j2.named_parameters()}
    if rank == 0:
        q.put((g_ddp, g_own, float(loss), gb))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_ddp_and_allreduce_step_two_gpus():
    """NCCL, 2 ranks: the sharded gradients (through DDP and through dist.allreduce_step) equal the
    single-process full-batch ones."""
    import torch.multiprocessing as mp
    from wenet_celoss_b200 import fused_joint_rnnt_loss
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, 29531, q)) for r in range(2)]
    for p in procs:
        p.start()
    g_ddp, g_own, loss, gb = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    joint, enc, pred, tgt, tl, ul = _case(4, 20, 6, 300, 32, 128, seed=41, ragged=False)
    j = copy.deepcopy(joint).to(DEV)
    full = fused_joint_rnnt_loss(j, enc.to(DEV), pred.to(DEV), tgt.to(DEV), tl.to(DEV), ul.to(DEV), blank=0,
                                 reduction="mean")
    full.backward()
    assert gb == 4
    np.testing.assert_allclose(loss, full.item(), rtol=1e-5)
    for n, p in j.named_parameters():
        _close("ddp " + n, g_ddp[n], p.grad.cpu(), tol=2e-3)   # fp32 atomics: order differs between runs
        _close("own " + n, g_own[n], p.grad.cpu(), tol=2e-3)
