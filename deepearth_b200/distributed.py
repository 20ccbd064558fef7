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


class GradSync:
    """Overlaps the table-gradient all-reduce with the backward scatter.

    The backward kernel walks level groups in order and each level's gradient is a contiguous slice of
    the flat buffer, so the call is cut into ``chunks`` consecutive runs of levels: as soon as one run's
    launch is queued, its slice is all-reduced on a side stream while the next run's reductions execute.
    Only the last slice's collective is exposed.  Semantics are DDP's (mean over ranks when ``average``).
    """

    def __init__(self, group: Optional[dist.ProcessGroup] = None, average: bool = True, chunks: int = 4):
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.group = group
        self.average = bool(average)
        self.chunks = max(1, int(chunks))
        self.world = dist.get_world_size(group)
        self._stream = None

    def stream(self, device) -> torch.cuda.Stream:
        if self._stream is None or self._stream.device != device:
            self._stream = torch.cuda.Stream(device=device)
        return self._stream

    def level_runs(self, n_levels: int) -> List[Tuple[int, int]]:
        k = min(self.chunks, n_levels)
        return [shard_bounds(n_levels, i, k) for i in range(k)]

    def grouped_ok(self, specs, n_points: int, n_features: int, linear: bool, levels_per_group: int, extra_flags: int,
                   group_width: int) -> bool:
        """True when a group-major gradient buffer of slice width ``group_width`` (functional.backward_group_width of the
        whole level list) can be walked run by run: every run must choose the same levels-per-group on its own."""
        import dataclasses
        from . import functional as Fn
        if not group_width or group_width % n_features:
            return False
        G = group_width // n_features
        for b, e in self.level_runs(len(specs)):
            if b % G or e % G:
                return False
            c0 = specs[b].out_col
            rebased = [dataclasses.replace(s, out_col=s.out_col - c0) for s in specs[b:e]]
            if Fn.backward_group_width(rebased, n_points, n_features, linear, levels_per_group, extra_flags) != group_width:
                return False
        return True

    def backward(self, specs, coords, grad_out, flat: torch.Tensor, grad_offsets, n_features: int, linear: bool,
                 mask_a=None, mask_b=None, levels_per_group: int = 0, extra_flags: int = 0, group_width: int = 0) -> None:
        """flat (zeroed) <- all-reduced table gradients of this rank's points.  ``group_width`` = W > 0: ``grad_out`` is the
        group-major buffer g[c // W][p][c % W] (see functional.encode_backward); each run walks its own planes."""
        from . import functional as Fn
        dev = flat.device
        main = torch.cuda.current_stream(dev)
        side = self.stream(dev)
        base = flat.data_ptr()
        ends = list(grad_offsets[1:]) + [flat.numel()]
        n = coords.shape[0]
        G = group_width // n_features if group_width else 0
        planes = grad_out.reshape(-1) if group_width else None
        for b, e in self.level_runs(len(specs)):
            go = planes[(b // G) * n * group_width:(e // G) * n * group_width] if group_width else grad_out
            Fn.encode_backward(specs[b:e], coords, go, [base + 4 * o for o in grad_offsets[b:e]], n_features,
                               linear, mask_a, mask_b, levels_per_group=levels_per_group, extra_flags=extra_flags,
                               group_width=group_width)
            ready = torch.cuda.Event()
            ready.record(main)
            side.wait_event(ready)
            piece = flat[grad_offsets[b]:ends[e - 1]]
            with torch.cuda.stream(side):
                if self.average:
                    piece.div_(self.world)
                dist.all_reduce(piece, op=dist.ReduceOp.SUM, group=self.group)
        done = torch.cuda.Event()
        done.record(side)
        main.wait_event(done)
        flat.record_stream(side)
