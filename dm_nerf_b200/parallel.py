"""Multi-GPU plumbing for the ray path: one process per GPU, torch.distributed.

* render: contiguous pixel blocks per rank, no communication until one final gather to rank 0;
* train : data-parallel over ray batches, one all-reduce of the flat parameter-gradient buffer.
The compute itself is passed in, so the host logic is testable with gloo on CPU."""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [start, stop) of ``total`` units for ``rank`` (first ``total % world`` ranks
    get one extra unit)."""
    base, extra = divmod(total, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def gather_rows(local: torch.Tensor, total: int, dst: int = 0) -> Optional[torch.Tensor]:
    """Gather row-sharded ``local`` ([n_r, ...], shards made by shard_range) onto ``dst`` as [total, ...]."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = [shard_range(total, r, world) for r in range(world)]
    longest = max(b - a for a, b in sizes)
    pad = local
    if local.shape[0] < longest:
        pad = torch.cat([local, local.new_zeros((longest - local.shape[0],) + tuple(local.shape[1:]))], 0)
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad.contiguous(), bufs, dst=dst)
    if rank != dst:
        return None
    return torch.cat([bufs[r][: sizes[r][1] - sizes[r][0]] for r in range(world)], 0)


def gather_frame(local: Dict[str, torch.Tensor], frame: Optional[Dict[str, torch.Tensor]], total: int,
                 keys: Sequence[str], dst: int = 0) -> Optional[Dict[str, torch.Tensor]]:
    """The final gather of a row-sharded frame without padding or concatenation: every rank sends its block of each
    output ONCE, point to point, and ``dst`` receives it straight into rows [start_r, stop_r) of its preallocated frame
    buffer (a contiguous slice: rows are the leading dimension).  All transfers form one batch (one NCCL group = the
    ncclSend/ncclRecv formulation of a gather).  ``frame`` may already hold ``dst``'s own block (when ``local`` are views
    of it nothing is copied).  Returns ``frame`` on ``dst``, None elsewhere."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        if frame is None:
            return dict(local)
        for k in keys:
            if frame[k].data_ptr() != local[k].data_ptr():
                frame[k][: local[k].shape[0]].copy_(local[k])
        return frame
    world, rank = dist.get_world_size(), dist.get_rank()
    ops_ = []
    if rank == dst:
        assert frame is not None, "gather_frame: the destination rank passes the frame buffers"
        a0, b0 = shard_range(total, dst, world)
        for k in keys:
            if frame[k][a0:b0].data_ptr() != local[k].data_ptr():
                frame[k][a0:b0].copy_(local[k])
            for r in range(world):
                if r != dst:
                    a, b = shard_range(total, r, world)
                    if b > a:
                        ops_.append(dist.P2POp(dist.irecv, frame[k][a:b], r))
    else:
        for k in keys:
            if local[k].shape[0] > 0:
                ops_.append(dist.P2POp(dist.isend, local[k].contiguous(), dst))
    if ops_:
        for w in dist.batch_isend_irecv(ops_):
            w.wait()
    return frame if rank == dst else None


def render_sharded(render_fn: Callable[..., Dict[str, torch.Tensor]], total: int,
                   keys: Sequence[str] = ("rgb", "depth", "labels"),
                   frame: Optional[Dict[str, torch.Tensor]] = None) -> Optional[Dict[str, torch.Tensor]]:
    """ONE frame split over the ranks (tester.py:62-76's chunk loop, sharded): rank r renders the contiguous pixel
    block ``shard_range(total, r, world)`` with ``render_fn(pix0, count)`` -- no data-path communication -- and rank 0
    ends up with the whole frame (``gather_frame``).  With ``frame`` (rank 0's preallocated buffers) rank 0 renders
    straight into its rows: ``render_fn(pix0, count, out=...)`` is then called with views of them."""
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    a, b = shard_range(total, rank, world)
    if frame is not None and rank == 0:
        local = render_fn(a, b - a, out={k: frame[k][a:b] for k in frame})
    else:
        local = render_fn(a, b - a)
    if rank == 0 and frame is None and world > 1:
        frame = {k: torch.empty((total,) + tuple(local[k].shape[1:]), dtype=local[k].dtype, device=local[k].device)
                 for k in keys}
    return gather_frame(local, frame, total, keys)


def allreduce_grads(params: List[torch.nn.Parameter], flat: Optional[torch.Tensor] = None) -> None:
    """Average the parameters' gradients over ranks with ONE all-reduce of a flat fp32 buffer
    (2 x 698,789 floats = 5.6 MB at K=32)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    grads = [p.grad for p in params if p.grad is not None]
    n = sum(g.numel() for g in grads)
    if flat is None or flat.numel() != n:
        flat = torch.empty(n, dtype=torch.float32, device=grads[0].device)
    torch._foreach_copy_(list(flat.split([g.numel() for g in grads])), [g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    flat.div_(dist.get_world_size())
    # copy back through views shaped like the gradients: reshape(-1) of a non-contiguous grad would be a temporary
    torch._foreach_copy_(grads, [c.view_as(g) for c, g in zip(flat.split([g.numel() for g in grads]), grads)])
