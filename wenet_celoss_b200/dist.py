"""Data-parallel exchange for the hot path: one flat all-reduce per step.

The reference shards utterances over ranks and lets DistributedDataParallel all-reduce every
parameter gradient (wenet/bin/train.py:141,231; data[rank::world_size], dataset.py:95-98).
For this path the only tensors that cross ranks are the joint / CTC weight gradients and the
scalar losses: they are packed into ONE fp32 buffer and reduced with a
single collective (NCCL over NVLink on GPUs, gloo in the CPU tests).  Utterance lattices,
d enc_out and d pred_out stay local.  ``reduction="mean"`` semantics: the caller passes the
local SUM of costs and the local batch size; the mean is taken over the GLOBAL batch.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_batch(n_utts: int, rank: int, world: int, weights: Sequence[float] = None) -> List[int]:
    """Utterance indices of this rank.  Without weights: the reference's ``rank::world`` slice.
    With per-utterance work estimates (e.g. T_b * U_b): sort by work and deal round-robin so
    ragged batches balance."""
    if weights is None:
        return list(range(rank, n_utts, world))
    order = sorted(range(n_utts), key=lambda i: -float(weights[i]))
    return sorted(order[rank::world])


def flatten_grads(params: Iterable[torch.nn.Parameter], extra: Sequence[torch.Tensor] = ()):
    """-> flat fp32 buffer [sum(numel) + len(extra)] and the split sizes."""
    params = list(params)
    parts = [(p.grad if p.grad is not None else torch.zeros_like(p)).detach().reshape(-1).float()
             for p in params]
    parts += [e.detach().reshape(-1).float() for e in extra]
    return torch.cat(parts), [p.numel() for p in parts]


def allreduce_step(params: Iterable[torch.nn.Parameter], cost_sum: torch.Tensor, local_batch: int,
                   group=None) -> Tuple[torch.Tensor, int]:
    """All-reduce (sum) the gradients of ``params`` and {cost_sum, batch} in one collective, then
    rescale: gradients were taken of the LOCAL sum of costs, the result is the gradient of the
    GLOBAL mean.  Returns (global mean loss, global batch)."""
    params = list(params)
    dev = cost_sum.device
    extra = [cost_sum.reshape(1), torch.tensor([float(local_batch)], device=dev)]
    flat, sizes = flatten_grads(params, extra)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    chunks = flat.split(sizes)
    global_batch = int(round(float(chunks[-1].item())))
    inv = 1.0 / max(global_batch, 1)
    for p, g in zip(params, chunks[:len(params)]):
        p.grad = (g * inv).view_as(p).to(p.dtype)
    return chunks[-2].reshape(()) * inv, global_batch
