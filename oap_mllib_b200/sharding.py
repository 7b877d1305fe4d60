"""Host-side helpers of the multi-GPU path: row sharding (one contiguous block per rank, as the
reference gives one merged table per executor, KMeansDALImpl.scala:48-50) and the layout of the
per-iteration all-reduce payload [k*ld sums | k counts | cost] (FP64) that replaces the reference's
allgatherv + root-side merge (native/KMeansDALImpl.cpp:80-130).  No arithmetic of the Lloyd step
happens here; the layout is shared with csrc/kmb200.cu (Workspace::red)."""
from __future__ import annotations

import numpy as np


def shard_bounds(rows: int, world: int):
    """[(row0, row1)] per rank: contiguous, disjoint, covering, sizes differ by at most one row."""
    if rows < 0 or world < 1:
        raise ValueError("rows >= 0 and world >= 1 required")
    edges = [(rows * r) // world for r in range(world + 1)]
    return [(edges[r], edges[r + 1]) for r in range(world)]


def padded_ld(d: int) -> int:
    return (d + 31) // 32 * 32


def payload_size(k: int, d: int) -> int:
    return k * padded_ld(d) + k + 1


def pack_partial(sums: np.ndarray, counts: np.ndarray, cost: float) -> np.ndarray:
    k, d = sums.shape
    ld = padded_ld(d)
    buf = np.zeros(payload_size(k, d), dtype=np.float64)
    buf[:k * ld].reshape(k, ld)[:, :d] = sums
    buf[k * ld:k * ld + k] = counts           # exact in FP64 below 2^53
    buf[k * ld + k] = cost
    return buf


def unpack_partial(buf: np.ndarray, k: int, d: int):
    ld = padded_ld(d)
    sums = buf[:k * ld].reshape(k, ld)[:, :d].copy()
    counts = np.rint(buf[k * ld:k * ld + k]).astype(np.int64)
    return sums, counts, float(buf[k * ld + k])
