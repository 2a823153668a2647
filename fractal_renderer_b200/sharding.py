"""Host-side plumbing for one-process-per-GPU runs (torchrun): which rows a rank owns and how the shards become ONE image.
The data path has NO collective (shards are independent).  Two ways to the whole frame:

  * `render_into_shared_frame` — the production path: the frame lives in POSIX shared memory that every rank maps and pins
    (`SharedFrame`, frb_shared_frame_*); each rank's frb_render_bands writes its interleaved bands straight into it and a
    barrier in the frame's header says "complete".  No gather step, no second copy; this is what bench.py times at N > 1.
  * `render_distributed` — compact shards sent point to point to rank 0 and assembled there; works with any
    torch.distributed backend (`gloo` in the CPU tests) for callers whose ranks do not share a host."""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np

from . import Shard
from .calc import Config

BAND_ROWS = 16  # interleaved bands of <= 16 rows keep the 8-GPU imbalance <= 1.015


def rank_shard(rank: int, world: int, band_rows: int = BAND_ROWS) -> Shard:
    return Shard(rank, world, band_rows) if world > 1 else Shard()


def frame_of_rank(n_frames: int, rank: int, world: int) -> list[int]:
    """Frames of an animation owned by `rank`: round robin (imbalance 1.018 on the 240-frame zoom vs 2.63 for blocks)."""
    return list(range(rank, n_frames, world))


def assemble(parts: list[np.ndarray], height: int, band_rows: int = BAND_ROWS) -> np.ndarray:
    """parts[r] = compact shard image of rank r, shape (rows_r, W, C).  Returns the (height, W, C) frame."""
    world = len(parts)
    w, c = parts[0].shape[1], parts[0].shape[2]
    out = np.empty((height, w, c), parts[0].dtype)
    for r, part in enumerate(parts):
        rows = rank_shard(r, world, band_rows).image_rows(height)
        assert len(rows) == part.shape[0], (r, len(rows), part.shape)
        out[rows] = part
    return out


def render_into_shared_frame(config: Config, frame, rank: int, world: int, render_bands: Optional[Callable] = None, fmt: int = 3,
                             device: int = 0, band_rows: int = BAND_ROWS) -> np.ndarray:
    """Every rank calls this with the same `frame` (a SharedFrame of >= H*W*fmt bytes): renders the rank's shard into its place
    and returns the (H, W, fmt) view of the complete image once ALL ranks have written theirs.  `render_bands(config, shard,
    frame_array, fmt, device)` defaults to the GPU call (frb_render_bands); tests inject a CPU stand-in."""
    if render_bands is None:
        from . import render_bands as gpu_render_bands

        render_bands = gpu_render_bands
    render_bands(config, rank_shard(rank, world, band_rows), frame.array, fmt, device)
    frame.barrier()
    return frame.array[: config.height * config.width * fmt].reshape(config.height, config.width, fmt)


def render_distributed(config: Config, render_shard: Callable[[Config, Shard], np.ndarray], dist=None,
                       band_rows: int = BAND_ROWS) -> Optional[np.ndarray]:
    """Every rank renders its shard with `render_shard` (the GPU call in production, anything in tests) and rank 0
    returns the assembled frame (other ranks return None)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return render_shard(config, Shard())
    import torch

    rank, world = dist.get_rank(), dist.get_world_size()
    mine = np.ascontiguousarray(render_shard(config, rank_shard(rank, world, band_rows)))
    t = torch.from_numpy(mine)
    # shards differ in size (ragged last band), so point-to-point instead of gather
    if rank == 0:
        parts = [mine]
        for r in range(1, world):
            buf = torch.empty((rank_shard(r, world, band_rows).rows(config.height),) + tuple(mine.shape[1:]), dtype=t.dtype)
            dist.recv(buf, src=r)
            parts.append(buf.numpy())
        return assemble(parts, config.height, band_rows)
    dist.send(t, dst=0)
    return None
