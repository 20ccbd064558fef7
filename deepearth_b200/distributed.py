"""Points-sharded data parallelism for the encoder.

Points are independent and tables are read-only in forward, so the path shards by points with
replicated tables; the only exchange step is the all-reduce of the dense table gradients that the
reference gets implicitly from DDP (hpc/train_distrbuted.py:499-504).  Grid4DFunction hands autograd one
flat fp32 gradient buffer per encode, so this is ONE collective over NVLink instead of one per level.
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_points: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous [begin, end) slice of the batch for ``rank``; sizes differ by at most one point."""
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, rem = divmod(int(n_points), int(world_size))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def flat_grad_buffers(params: Iterable[torch.nn.Parameter]) -> List[torch.Tensor]:
    """Distinct underlying flat buffers behind the ``.grad`` views of ``params`` (usually exactly one
    per encode call, because Grid4DFunction returns views of one flat buffer)."""
    seen, out = set(), []
    for p in params:
        g = p.grad
        if g is None:
            continue
        st = g.untyped_storage()
        key = st.data_ptr()
        if key not in seen:
            seen.add(key)
            # a flat tensor over the whole storage: for Grid4DFunction's grads that is the single
            # zero-initialised buffer all level gradients are views of
            out.append(torch.empty(0, dtype=g.dtype, device=g.device).set_(st))
    return out


def allreduce_table_grads(params: Iterable[torch.nn.Parameter], group: Optional[dist.ProcessGroup] = None,
                          average: bool = True, async_op: bool = False):
    """All-reduce the table gradients in as few collectives as there are flat buffers (DDP semantics:
    mean over ranks when ``average``)."""
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed is not initialised")
    world = dist.get_world_size(group)
    works = []
    for buf in flat_grad_buffers(params):
        if average:
            buf.div_(world)
        works.append(dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group, async_op=async_op))
    return works if async_op else None
