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


def layout(rank: int, world: int, mode: str):
    """(tile_rank, tile_world, group, n_groups) of a rank: 'samples' = one sample range per rank,
    'tiles' = one interleaved tile set per rank, 'combined' = tile ranks inside a group x sample ranges across groups
    (4 tile ranks per group from 8 ranks up, else 2)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} of {world}")
    if world == 1 or mode == "samples":
        return 0, 1, rank, world
    if mode == "tiles":
        return rank, world, 0, 1
    if mode != "combined":
        raise ValueError(mode)
    tile_world = 4 if world % 4 == 0 and world > 4 else 2
    if world % tile_world:
        raise ValueError(f"combined partition needs a multiple of {tile_world} ranks")
    return rank % tile_world, tile_world, rank // tile_world, world // tile_world
