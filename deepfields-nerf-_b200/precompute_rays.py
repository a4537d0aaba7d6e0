"""Drop-in for the reference's scripts/precompute_rays.py: the same function names and argument meaning, computed
on the GPU (csrc/raygen.cu) and returned as device tensors instead of numpy arrays / an .npz file.

    get_rays(H, W, K, camera_poses, colmap=True, normalize_directions=True) -> (rays_o, rays_d)   # [M,H,W,3] fp32
    ray_cube_intersection_numpy(ray_origins, ray_directions, cube_min=-1.0, cube_max=1.0) -> (intersects, t_near, t_far)
    precompute_rays_on_device(H, W, K, camera_poses, images_u8) -> {"rays_o","rays_d","rgbs","t_near","t_far"}  # hits only, in order

The reference keeps its rays in fp64 until dataset.py:52-53 casts them to fp32; here the fp64 arithmetic happens in
the kernel and the fp32 values are what leaves it.
"""
from __future__ import annotations

import torch

from . import ops


def get_rays(H, W, K, camera_poses, colmap=True, normalize_directions=True, device="cuda"):
    """scripts/precompute_rays.py:67-119."""
    return ops.generate_rays(int(H), int(W), K, camera_poses, colmap, normalize_directions, device=device)


def ray_cube_intersection(ray_origins, ray_directions, cube_min=-1.0, cube_max=1.0):
    """scripts/precompute_rays.py:10-64 on device tensors (numpy inputs are uploaded)."""
    dev = ray_origins.device if isinstance(ray_origins, torch.Tensor) else torch.device("cuda")
    o = torch.as_tensor(ray_origins, dtype=torch.float32).to(dev)
    d = torch.as_tensor(ray_directions, dtype=torch.float32).to(dev)
    return ops.ray_cube_intersect(o, d, cube_min, cube_max)


ray_cube_intersection_numpy = ray_cube_intersection   # the reference's name


def precompute_rays_on_device(H, W, K, camera_poses, images=None, cube_min=-1.0, cube_max=1.0, device="cuda"):
    """scripts/precompute_rays.py:206-228 without the .npz round trip: the dict train.RayStore.from_tensors takes."""
    if images is not None and not isinstance(images, torch.Tensor):
        images = torch.as_tensor(images)
    return ops.precompute_rays(int(H), int(W), K, camera_poses, images, cube_min, cube_max, device=device)
