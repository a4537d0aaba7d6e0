"""Generates tests/golden/volume_points_golden.npz with the reference's scripts/post_process_vol.py: slice_aabb is
imported from it; the statements of its main() that build the point set (:82, :89-95) are not a function there, so they
are executed here as written.  omegaconf and vispy are absent and unused by these lines -> empty stub modules.

Run once in the build container:  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_volume.py
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.dont_write_bytecode = True
REF = "/root/reference/scripts"

from common import make_volume  # noqa: E402

AABB = {"aabb_min": [-0.55, -0.4, -0.7], "aabb_max": [0.6, 1.0, 0.22580645161290325]}   # one bound exactly on a voxel coordinate


def main():
    om = types.ModuleType("omegaconf")
    om.OmegaConf = type("OmegaConf", (), {})
    sys.modules.setdefault("omegaconf", om)
    vp = types.ModuleType("vispy")
    vp.scene, vp.app = types.ModuleType("vispy.scene"), types.ModuleType("vispy.app")
    sys.modules.setdefault("vispy", vp)
    sys.path.insert(0, REF)
    import post_process_vol as R  # the reference

    occ, rgb = make_volume(5, 32)
    volume, rgb_volume = torch.from_numpy(occ), torch.from_numpy(rgb)
    # post_process_vol.py:82, :89-95, as written
    rgb_volume = rgb_volume / rgb_volume.max()
    points = np.argwhere(volume.numpy())
    rgbs = rgb_volume[points[:, 0], points[:, 1], points[:, 2]].numpy()
    res = volume.shape[0]
    points = (points / (res - 1)) * 2 - 1
    sp, sr = R.slice_aabb(points, rgbs, AABB)
    out = dict(points=points, rgbs=rgbs, sliced_points=sp, sliced_rgbs=sr)
    path = os.path.join(HERE, "volume_points_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: (v.shape, str(v.dtype)) for k, v in out.items()})


if __name__ == "__main__":
    main()
