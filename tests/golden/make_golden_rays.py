"""Generates tests/golden/rays_golden.npz by RUNNING THE REFERENCE's scripts/precompute_rays.py functions
(imported read-only from /root/reference/scripts) on the deterministic cameras of common.make_cameras.

Run once in the build container:  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_rays.py
omegaconf is absent and only used by the reference's CLI entry point -> empty stub module.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.dont_write_bytecode = True
REF = "/root/reference/scripts"

from common import edge_case_rays, make_cameras  # noqa: E402


def main():
    om = types.ModuleType("omegaconf")
    om.OmegaConf = type("OmegaConf", (), {})
    sys.modules.setdefault("omegaconf", om)
    sys.path.insert(0, REF)
    import precompute_rays as R  # the reference

    H, W, K, poses, images = make_cameras(7, M=3, H=12, W=16)
    rays_o, rays_d = R.get_rays(H, W, K, poses)                     # fp64 [M,H,W,3]
    o, d = rays_o.reshape(-1, 3), rays_d.reshape(-1, 3)
    hits, t_near, t_far = R.ray_cube_intersection_numpy(o, d)
    rgb = (images.astype(np.float32) / 255.0).reshape(-1, 3)        # load_poses_from_json, precompute_rays.py:131-132
    eo, ed = edge_case_rays()
    with np.errstate(divide="ignore", invalid="ignore"):
        ehits, enear, efar = R.ray_cube_intersection_numpy(eo, ed, -0.5, 0.75)
    out = dict(rays_o=np.ascontiguousarray(rays_o), rays_d=rays_d, hits=hits, t_near=t_near, t_far=t_far,
               kept_o=o[hits], kept_d=d[hits], kept_rgb=rgb[hits], kept_near=t_near[hits], kept_far=t_far[hits],
               edge_hits=ehits, edge_near=enear, edge_far=efar)
    path = os.path.join(HERE, "rays_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: (v.shape, str(v.dtype)) for k, v in out.items()}, "hits", int(hits.sum()), "of", hits.size)


if __name__ == "__main__":
    main()
