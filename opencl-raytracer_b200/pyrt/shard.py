"""Row-block sharding of a frame over ranks: blocks of `block_rows` rows dealt round-robin,
rank r owns blocks b with b % n_shards == r.  Host logic; mirrors rt_set_shard / locate_pixel in the CUDA library."""
from __future__ import annotations

import numpy as np


def shard_rows(height: int, block_rows: int, shard: int, n_shards: int) -> np.ndarray:
    """Global row indices owned by `shard`, ascending."""
    if block_rows <= 0 or n_shards <= 0 or not (0 <= shard < n_shards):
        raise ValueError("bad shard")
    rows = np.arange(height, dtype=np.int64)
    return rows[(rows // block_rows) % n_shards == shard]


def scatter_rows(frame: np.ndarray, width: int, rows: np.ndarray, shard_pixels: np.ndarray) -> None:
    """Place a shard's pixels (rows in ascending order, row-major) into the full frame (N, C)."""
    idx = (rows[:, None] * width + np.arange(width, dtype=np.int64)[None, :]).reshape(-1)
    frame[idx] = shard_pixels
