"""Drop-in for the compute part of the reference's scripts/post_process_vol.py (the viewer is out of scope): the
occupied voxels of an extracted volume as a coloured point set, optionally sliced by the scene's bounding box.

    volume_points(occupancy_volume, rgb_volume, aabb_meta=None) -> (points [N,3] fp64 in [-1,1], rgbs [N,3] fp32)
        post_process_vol.py:89-97 (argwhere, colour gather, normalisation) fused with slice_aabb (:37-50), one kernel
        pair on the device (csrc/volpoints.cu)
    slice_aabb(points, rgbs, aabb_meta)     the reference's function on device tensors (boundary helper)
    save_volume(points, rgbs, aabb_meta, path)   the same dict keys as post_process_vol.py:10-14
"""
from __future__ import annotations

import torch

from . import ops


def volume_points(occupancy_volume, rgb_volume, aabb_meta=None):
    lo = hi = None
    if aabb_meta is not None:
        lo, hi = list(aabb_meta["aabb_min"]), list(aabb_meta["aabb_max"])
    return ops.volume_points(occupancy_volume, rgb_volume, lo, hi)


def slice_aabb(points, rgbs, aabb_meta):
    """scripts/post_process_vol.py:37-50 (closed box) on tensors that are already point lists."""
    lo = torch.as_tensor(list(aabb_meta["aabb_min"]), dtype=points.dtype, device=points.device)
    hi = torch.as_tensor(list(aabb_meta["aabb_max"]), dtype=points.dtype, device=points.device)
    mask = ((points >= lo) & (points <= hi)).all(dim=1)
    return points[mask], rgbs[mask]


def save_volume(points, rgbs, aabb_meta, sliced_vol_path):
    """scripts/post_process_vol.py:10-14: numpy arrays under the reference's keys."""
    torch.save({"points_normalized": points.cpu().numpy(), "rgbs": rgbs.cpu().numpy(), "aabb_meta": aabb_meta}, sliced_vol_path)
