"""Host-side multi-GPU plumbing: which frames / tiles a rank renders and how the per-rank float
accumulation buffers are combined at the read-back boundary.  torch.distributed is plumbing only; the backend
is chosen by the caller ("nccl" on GPUs, "gloo" in the CPU tests)."""
from __future__ import annotations


def frame_range(rank: int, world: int):
    """Sample-range sharding: rank g renders global 1-based frames g+1, g+1+G, ... -> (first, stride)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} of {world}")
    return rank + 1, world


def frames_of_rank(rank: int, world: int, frames_per_rank: int):
    first, stride = frame_range(rank, world)
    return [first + k * stride for k in range(frames_per_rank)]


def owns_tile(rank: int, world: int, x: int, y: int, width: int, tile_w: int = 32, tile_h: int = 32) -> bool:
    """Interleaved tiles, the rule of cudapt_set_tile_partition: (tile + tile_row) % world == rank."""
    if world <= 1:
        return True
    tiles_x = (width + tile_w - 1) // tile_w
    tx, ty = x // tile_w, y // tile_h
    return (ty * tiles_x + tx + ty) % world == rank


def reduce_accum(accum, dst: int = 0):
    """Sum the ranks' accumulation buffers (RGBA float, alpha = samples) onto rank dst.  One collective per
    read-back; `accum` is a torch tensor (a zero-copy view of cudapt_accum_device_ptr on GPUs)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def total_frame_count(frames_per_rank: int, world: int, mode: str) -> int:
    """Samples per pixel held by the reduced buffer: sample ranges add up, tiles do not."""
    return frames_per_rank * world if mode == "frames" else frames_per_rank
