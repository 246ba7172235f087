"""Multi-GPU partition of the RtReference path: by sample.

Every (pixel, frameIndex) sample is independent (RNG seed = (px, py, frameIndex), main.rgen:229), so rank g of G
renders its own block of frame indices of the same image with the scene replicated, and ONE reduce of the
(sum.rgb, count) image combines the ranks. This module is the host-side arithmetic of that scheme; the device
side is prt_export_sum_device / prt_import_sum_device (k_mean_to_sum / k_sum_to_mean).
"""
from __future__ import annotations

import numpy as np

FRAME_PERIOD = 4096  # RtRef.cpp:31


def rank_frames(rank: int, world_size: int, spp: int) -> tuple[int, int]:
    """(first frameIndex, count) of `rank`: frames rank*spp+1 .. (rank+1)*spp. The reference's counter starts at 1
    (RtRef.cpp:170) and wraps at 4096, so all ranks together must stay inside one period."""
    if not (0 <= rank < world_size):
        raise ValueError("rank outside the world")
    if spp <= 0:
        raise ValueError("spp must be positive")
    if world_size * spp > FRAME_PERIOD - 1:
        raise ValueError("spp x ranks exceeds the reference's frame-index period (RtRef.cpp:31)")
    return rank * spp + 1, spp


def mean_to_sum(mean_rgba: np.ndarray) -> np.ndarray:
    """Running-mean image (rgb = mean, a = sample count; main.rgen:285-298) -> (sum.rgb, count)."""
    out = np.array(mean_rgba, dtype=np.float32, copy=True)
    out[..., :3] *= out[..., 3:4]
    return out


def sum_to_mean(sum_rgba: np.ndarray) -> np.ndarray:
    out = np.array(sum_rgba, dtype=np.float32, copy=True)
    n = out[..., 3:4]
    inv = np.where(n > 0, np.float32(1.0) / np.where(n > 0, n, 1), np.float32(0.0)).astype(np.float32)
    out[..., :3] *= inv
    return out


def reduce_sum_image(sum_tensor, dst: int = 0):
    """The path's single collective: an in-place SUM reduce of the (sum.rgb, count) image to `dst`."""
    import torch.distributed as dist
    dist.reduce(sum_tensor, dst=dst, op=dist.ReduceOp.SUM)
    return sum_tensor
