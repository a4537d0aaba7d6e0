"""Drop-in for the reference's scripts/sampling.py (StratifiedSampler), backed by CUDA kernels."""
from __future__ import annotations

import torch

from . import ops


class StratifiedSampler:
    """Same constructor / call signatures as scripts/sampling.py:4-49."""

    def __init__(self, N_samples, device):
        self.N_samples = N_samples
        self.device = device
        self.t_vals = torch.linspace(0, 1, N_samples, device=device)  # sampling.py:8

    def sample_z(self, near, far, t_rand=None):
        """Jittered depths only (the fused render path never materialises points)."""
        B = near.shape[0]
        if t_rand is None:
            t_rand = torch.rand((B, self.N_samples), device=self.device)  # sampling.py:20
        return ops.stratified_sample(near, far, self.t_vals, t_rand)

    def __call__(self, rays_o, rays_d, near, far, t_rand=None):
        z_vals = self.sample_z(near, far, t_rand)
        pts_flat, input_dir = ops.points_on_rays(rays_o, rays_d, z_vals)
        return pts_flat, input_dir, z_vals

    def fine_points(self, rays_o, rays_d, z_samples, z_vals, near=None, far=None, N_samples=None, N_importance=None):
        """sampling.py:34-49 (near/far/N_samples/N_importance are unused there too)."""
        z_all = ops.merge_sort_z(z_vals, z_samples)
        pts_flat, input_dir = ops.points_on_rays(rays_o, rays_d, z_all)
        return pts_flat, input_dir, z_all
