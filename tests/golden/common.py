"""Deterministic synthetic inputs shared by the golden generator, the tests and bench.py.

Everything is drawn from numpy's frozen legacy MT19937 stream (np.random.RandomState), so the
same seed reproduces bit-identical inputs on any machine / torch version; the golden files then
only need to hold reference OUTPUTS.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch


def make_state_dict(seed: int, hidden: int = 256, pos_L: int = 10, dir_L: int = 4):
    """State dict with the reference's key set / shapes (scripts/model.py:87-137) and an
    nn.Linear-like init (U(+-1/sqrt(in)); sigma head N(0,0.01) / bias 0.1, model.py:126-127)."""
    rs = np.random.RandomState(seed)
    pd, dd = 3 + 6 * pos_L, 3 + 6 * dir_L
    sd = OrderedDict()
    sd["pos_encoding.freq_bands"] = (2.0 ** torch.arange(pos_L)) * torch.pi
    sd["dir_encoding.freq_bands"] = (2.0 ** torch.arange(dir_L)) * torch.pi

    def lin(name, o, i):
        bound = 1.0 / math.sqrt(i)
        sd[name + ".weight"] = torch.from_numpy(rs.uniform(-bound, bound, (o, i)).astype(np.float32))
        sd[name + ".bias"] = torch.from_numpy(rs.uniform(-bound, bound, (o,)).astype(np.float32))

    for j, i in zip((0, 2, 4, 6), (pd, hidden, hidden, hidden)):
        lin(f"pos_layers.{j}", hidden, i)
    for j, i in zip((0, 2, 4, 6), (hidden + pd, hidden, hidden, hidden)):
        lin(f"skip_layers.{j}", hidden, i)
    sd["density_layer.weight"] = torch.from_numpy(rs.normal(0.0, 0.01, (1, hidden)).astype(np.float32))
    sd["density_layer.bias"] = torch.full((1,), 0.1)
    lin("feature_layer", hidden, hidden)
    lin("color_layer.0", hidden // 2, hidden + dd)
    lin("color_layer.2", 3, hidden // 2)
    return sd


def peaked(sd, scale: float = 30.0):
    """'Peaked sigma' variant: scales the sigma head so rays saturate and the
    pdf / early-termination paths see non-uniform weights."""
    out = OrderedDict((k, v.clone()) for k, v in sd.items())
    out["density_layer.weight"] *= scale * 10.0
    out["density_layer.bias"] *= scale
    return out


def make_rays(seed: int, B: int):
    """Cameras on a shell of radius 2.5-4 looking at points in [-0.8,0.8]^3; unit directions;
    t_near/t_far from the slab test against [-1,1]^3 (the distribution precompute_rays.py:10-120
    produces).  Returns float32 CPU tensors: rays_o [B,3], rays_d [B,3], near [B,1], far [B,1],
    rgb [B,3]."""
    rs = np.random.RandomState(seed)
    v = rs.normal(size=(B, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    o = v * rs.uniform(2.5, 4.0, (B, 1))
    tgt = rs.uniform(-0.8, 0.8, (B, 3))
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o32, d32 = o.astype(np.float32), d.astype(np.float32)
    inv = 1.0 / d32.astype(np.float64)
    t0 = (-1.0 - o32) * inv
    t1 = (1.0 - o32) * inv
    near = np.maximum(np.minimum(t0, t1).max(axis=1), 0.0)
    far = np.maximum(t0, t1).min(axis=1)
    assert (far > near).all()
    rgb = rs.uniform(0.0, 1.0, (B, 3)).astype(np.float32)
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a.astype(np.float32)))
    return f(o32), f(d32), f(near)[:, None], f(far)[:, None], f(rgb)


def make_draws(seed: int, B: int, S: int, I: int):
    rs = np.random.RandomState(seed)
    t_rand = torch.from_numpy(rs.uniform(0, 1, (B, S)).astype(np.float32))
    u = torch.from_numpy(rs.uniform(0, 1, (B, I)).astype(np.float32))
    # float32 rounding of a double in [0,1) can land on 1.0; the reference's torch.rand cannot
    t_rand.clamp_(max=1.0 - 2.0 ** -24)
    u.clamp_(max=1.0 - 2.0 ** -24)
    return t_rand, u


def make_noise(seed: int, n: int):
    rs = np.random.RandomState(seed)
    return torch.from_numpy(rs.normal(size=(n, 1)).astype(np.float32))


def make_cameras(seed: int, M: int, H: int, W: int, focal: float = None):
    """Deterministic pinhole cameras for the ray-generation tests: intrinsics K (3x3, fp64), camera-to-world poses
    [M,4,4] fp64 on a shell of radius 2.5-4 around the origin, looking roughly (not exactly) at it so that part of every
    image misses the [-1,1]^3 cube, and u8 images [M,H,W,3]."""
    import numpy as np
    rng = np.random.default_rng(seed)
    f = float(focal) if focal is not None else 0.9 * W
    K = np.array([[f, 0.0, W / 2.0 - 0.25], [0.0, f * 1.01, H / 2.0 + 0.5], [0.0, 0.0, 1.0]], dtype=np.float64)
    poses = np.zeros((M, 4, 4), dtype=np.float64)
    for m in range(M):
        c = rng.normal(size=3)
        c = c / np.linalg.norm(c) * rng.uniform(2.5, 4.0)
        target = rng.uniform(-0.9, 0.9, size=3)
        back = c - target
        back /= np.linalg.norm(back)                     # camera +z points away from the scene (OpenGL convention)
        up = np.array([0.0, 0.0, 1.0]) if abs(back[2]) < 0.95 else np.array([0.0, 1.0, 0.0])
        right = np.cross(up, back)
        right /= np.linalg.norm(right)
        poses[m, :3, 0], poses[m, :3, 1], poses[m, :3, 2], poses[m, :3, 3] = right, np.cross(back, right), back, c
        poses[m, 3, 3] = 1.0
    images = rng.integers(0, 256, size=(M, H, W, 3), dtype=np.uint8)
    return H, W, K, poses, images


def edge_case_rays():
    """fp32 rays that exercise the special cases of the slab test: zero / tiny / negative-zero direction components,
    origins inside the cube, on a slab plane (0/0 -> NaN in numpy), behind the cube, grazing an edge."""
    import numpy as np
    o = np.array([[0.0, 0.0, 0.0], [2.0, 0.1, -0.2], [2.0, 0.1, -0.2], [0.75, 0.0, 0.0], [-0.5, 0.2, 0.3],
                  [0.1, 0.2, 3.0], [0.1, 0.2, 3.0], [3.0, 3.0, 3.0], [0.0, 0.74, -2.0], [0.2, -0.1, 0.3],
                  [5.0, 0.0, 0.0], [0.3, 0.3, 0.3]], dtype=np.float32)
    d = np.array([[0.0, 0.0, 1.0], [-1.0, 0.0, 0.0], [1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [-0.0, -1.0, 0.0],
                  [0.0, 0.0, -1.0], [1e-9, -1e-9, -1.0], [-0.577, -0.577, -0.577], [0.0, 1e-3, 1.0], [1e-9, 1e-12, -1e-9],
                  [-1.0, 0.2, 0.0], [0.0, 0.0, 0.0]], dtype=np.float32)
    return o, d


def make_volume(seed: int, R: int):
    """Deterministic u8 occupancy [R,R,R] (a blob plus speckle, ~20 % set) and u8 colours [R,R,R,3] (maximum < 255)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    g = np.linspace(-1, 1, R)
    x, y, z = np.meshgrid(g, g, g, indexing="ij")
    occ = ((x * x + 0.6 * y * y + 1.4 * z * z < 0.35) | (rng.random((R, R, R)) < 0.03)).astype(np.uint8)
    rgb = rng.integers(0, 241, size=(R, R, R, 3), dtype=np.uint8)
    return occ, rgb
