"""Frame sharding across the GPUs of one box.

Frames are independent, so a batch of F frames is split into contiguous shards, one per rank (one process
per GPU, torch.distributed; NCCL over NVLink on the GPU box, gloo in the CPU tests).  There is no data-path
collective: height / temperature maps stay on the GPU that produced them.  The only exchange is one gather of
the fixed-size per-frame scalar records (184 B each) to rank 0 per batch.
"""
from __future__ import annotations

from typing import List, Optional, Tuple


def shard_range(n_frames: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of rank `rank`: the first n % world ranks take one extra frame."""
    if world < 1 or not (0 <= rank < world) or n_frames < 0:
        raise ValueError(f"bad shard request n={n_frames} rank={rank} world={world}")
    base, extra = divmod(int(n_frames), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(n_frames: int, world: int) -> List[int]:
    return [shard_range(n_frames, r, world)[1] - shard_range(n_frames, r, world)[0] for r in range(world)]


def gather_records(local_records, n_frames: int, dst: int = 0, group=None):
    """Gather per-rank record tensors ([n_local, record_bytes] uint8; CUDA under NCCL, CPU under gloo) on `dst`,
    in global frame order.  Returns the [n_frames, record_bytes] tensor on `dst`, None elsewhere.
    Shards may be ragged (n_frames % world != 0): records are padded to the largest shard for the collective."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        if local_records.shape[0] != n_frames:
            raise ValueError("single process: local records must cover every frame")
        return local_records
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = shard_sizes(n_frames, world)
    if local_records.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank}: {local_records.shape[0]} records for a shard of {sizes[rank]} frames")
    width, mx = int(local_records.shape[1]), max(sizes) if sizes else 0
    send = torch.zeros((mx, width), dtype=local_records.dtype, device=local_records.device)
    send[:sizes[rank]].copy_(local_records)
    bufs = [torch.empty_like(send) for _ in range(world)] if rank == dst else None
    dist.gather(send, bufs, dst=dst, group=group)
    if rank != dst:
        return None
    return torch.cat([b[:s] for b, s in zip(bufs, sizes)], dim=0)


def run_sharded(plan, make_frames, n_frames: int, step_frames: int = 64, group=None, masks: bool = False):
    """Process frames [0, n_frames) over all ranks.  `make_frames(lo, hi)` returns this rank's frames lo..hi-1 as a
    uint8 CUDA tensor [hi-lo,H,W,3] (generated or pre-staged in HBM).  Returns the gathered records (rank 0) or None."""
    import torch
    import torch.distributed as dist
    on = dist.is_available() and dist.is_initialized()
    world, rank = (dist.get_world_size(group), dist.get_rank(group)) if on else (1, 0)
    lo, hi = shard_range(n_frames, rank, world)
    parts = []
    for a in range(lo, hi, step_frames):
        b = min(hi, a + step_frames)
        frames = make_frames(a, b)
        res = plan.run(frames, masks=masks, reuse_outputs=False)
        if hasattr(plan, "records_host"):
            plan.records_host(res, frames)        # re-runs frames that found no T10 / T12 slot (ST_RERUN_ALONE)
        parts.append(res["records"])
    width = parts[0].shape[1] if parts else 184
    local = torch.cat(parts, dim=0) if parts else torch.zeros((0, width), dtype=torch.uint8, device="cuda")
    return gather_records(local, n_frames, 0, group)
