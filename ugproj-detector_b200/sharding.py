"""Frame sharding and detection-list gathering for multi-GPU runs.

Detection has no inter-frame state (the tracker does, and stays on rank 0's host), so rank r of N
takes the contiguous frame range [floor(r*F/N), floor((r+1)*F/N)) and the only exchange step is the
gather of per-frame rect lists to rank 0.  Lists travel as fixed slots {n; rect[cap]} so that one
collective moves a whole batch; the payload is KBs, the point is one launch, not bandwidth.
"""
from __future__ import annotations

import numpy as np


def shard_range(nframes: int, rank: int, world: int):
    """Contiguous frame range of `rank`: [lo, hi)."""
    return (rank * nframes) // world, ((rank + 1) * nframes) // world


def pack_slots(rects, cap: int) -> np.ndarray:
    """list of [n_i,4] int32 arrays -> [F, cap+1, 4] int32; slot 0 holds (n_i, 0, 0, 0)."""
    out = np.zeros((len(rects), cap + 1, 4), np.int32)
    for i, r in enumerate(rects):
        n = min(len(r), cap)
        out[i, 0, 0] = len(r)
        if n:
            out[i, 1:1 + n] = np.asarray(r, np.int32).reshape(-1, 4)[:n]
    return out


def unpack_slots(slots: np.ndarray):
    """Inverse of pack_slots (truncated lists stay truncated to cap)."""
    cap = slots.shape[1] - 1
    return [slots[i, 1:1 + min(int(slots[i, 0, 0]), cap)].copy() for i in range(slots.shape[0])]


def init_comm(det, rank: int, world: int):
    """Gives the detector its own NCCL communicator (ugd_comm_init): rank 0 draws the unique id through the library and
    torch.distributed -- the harness's plumbing, any backend -- carries the 128 bytes to the other ranks.  The gather
    itself then runs inside libugdet.so (ugd_detect_batch_sharded), not through torch."""
    import torch.distributed as dist
    from . import comm_unique_id
    box = [comm_unique_id() if rank == 0 else None]
    if world > 1:
        dist.broadcast_object_list(box, src=0)
    det.comm_init(box[0], rank, world)


def gather_to_rank0(local_slots, world: int, rank: int, max_frames: int):
    """torch.distributed gather of slot tensors to rank 0 (NCCL on GPUs, gloo on CPU).
    local_slots: torch int32 tensor [F_r, cap+1, 4] (F_r may differ by one between ranks; padded to
    max_frames for the collective).  Returns on rank 0 the list of per-rank tensors, else None."""
    import torch
    import torch.distributed as dist
    f = local_slots.shape[0]
    if f < max_frames:
        pad = torch.zeros((max_frames - f,) + tuple(local_slots.shape[1:]), dtype=local_slots.dtype,
                          device=local_slots.device)
        local_slots = torch.cat([local_slots, pad])
    bufs = [torch.empty_like(local_slots) for _ in range(world)] if rank == 0 else None
    dist.gather(local_slots, bufs, dst=0)
    return bufs
