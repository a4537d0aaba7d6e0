"""Golden vectors for the dense-grid query and the volume.pth writer, produced by RUNNING THE REFERENCE
(imported read-only from /root/reference/scripts) -- the configurations bench.py times at 512^3.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden_extract.py

  * extract_occupancy_volume.query_density_and_rgb_batch (:118-172, the runnable twin of extract_vol.py:85-120) on the
    64^3 grid of generate_3d_grid (:90-115) and on rank 3's x-slab (of 8) of the 128^3 grid, fp32 on CPU, fine-net
    weights = peaked(make_state_dict(32)) so that sigma is not flat.  sigma is kept in fp32, rgb in fp16 (5e-4 of
    rounding against the 2e-2 bar of the bf16 MLP) and as the reference's own u8 packing (rgb * 255).byte().
  * extract_vol.save_volume (:153-206) on a 6^3 volume: the saved dict is re-loaded and committed (arrays + metadata), so the
    writer test compares keys, dtypes, shapes and values with what the reference itself wrote.  extract_vol.py imports
    omegaconf, skimage, matplotlib and tinycudann at module level only -> empty stubs.
  * the (key, shape, dtype) list of OptimizedVanillaNeRF.state_dict() (model.py:66-137), which nerf_final.pth must load into.
"""
from __future__ import annotations

import io
import json
import os
import sys
import tempfile
import types
from contextlib import redirect_stdout

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.dont_write_bytecode = True
REF = "/root/reference/scripts"

from common import make_state_dict, peaked  # noqa: E402


def import_reference():
    om = types.ModuleType("omegaconf")
    om.OmegaConf = type("OmegaConf", (), {})
    sys.modules.setdefault("omegaconf", om)
    sys.modules.setdefault("tinycudann", types.ModuleType("tinycudann"))
    sk, skf = types.ModuleType("skimage"), types.ModuleType("skimage.filters")
    skf.threshold_otsu = None
    sk.filters = skf
    sys.modules.setdefault("skimage", sk)
    sys.modules.setdefault("skimage.filters", skf)
    for name in ("matplotlib", "matplotlib.pyplot", "mpl_toolkits", "mpl_toolkits.mplot3d"):
        m = types.ModuleType(name)
        m.Axes3D = None
        sys.modules.setdefault(name, m)
    sys.path.insert(0, REF)
    import model as ref_model
    import extract_occupancy_volume as ref_eov
    import extract_vol as ref_ev
    return ref_model, ref_eov, ref_ev


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    ref_model, ref_eov, ref_ev = import_reference()
    out = {}
    sd = peaked(make_state_dict(32, 256), 3.0)
    net = ref_model.OptimizedVanillaNeRF(10, 4, 256)
    net.load_state_dict(sd)
    net.eval()

    def query(points):
        with redirect_stdout(io.StringIO()):
            dens, cols = ref_eov.query_density_and_rgb_batch(net, points, batch_size=8192, device="cpu", show_progress=False)
        return dens.reshape(-1), cols

    with redirect_stdout(io.StringIO()):
        g64 = ref_eov.generate_3d_grid((-1.0, 1.0), 64)
        g128 = ref_eov.generate_3d_grid((-1.0, 1.0), 128)
    s64, c64 = query(g64)
    out["r64_sigma"] = s64.numpy().astype(np.float32)
    out["r64_rgb_f16"] = c64.numpy().astype(np.float16)
    out["r64_rgb_u8"] = (c64 * 255).byte().numpy()                       # extract_vol.py:198
    R, world, rank = 128, 8, 3
    planes = (R + world - 1) // world
    a, b = rank * planes * R * R, (rank + 1) * planes * R * R
    s128, c128 = query(g128[a:b])
    out["r128_slab_range"] = np.array([a, b], np.int64)
    out["r128_slab_sigma"] = s128.numpy().astype(np.float32)
    out["r128_slab_rgb_f16"] = c128.numpy().astype(np.float16)

    # ---- save_volume (extract_vol.py:153-206) ----
    Rv = 6
    rs = np.random.RandomState(61)
    occ = torch.from_numpy(rs.uniform(0, 1, Rv ** 3) < 0.4)
    rgbv = torch.from_numpy(rs.uniform(0, 1, (Rv ** 3, 3)).astype(np.float32))
    meta_in = {"checkpoint_path": "ckpt/nerf_final.pth", "device_used": "cuda", "batch_size": 8192,
               "density_stats": {"min": 0.0, "max": 9.5, "mean": 1.25, "std": 2.5},
               "view_direction": [0.0, 0.0, -1.0],
               "rgb_stats": {"min": 0.01, "max": 0.99, "mean": 0.5, "std": 0.25}}
    with tempfile.TemporaryDirectory() as td, redirect_stdout(io.StringIO()):
        path = os.path.join(td, "volume.pth")
        ref_ev.save_volume(occ, rgbv, path, Rv, (-1.0, 1.0), 0.375, dict(meta_in))
        vol = torch.load(path, map_location="cpu", weights_only=False)
    out["sv_occ_in"], out["sv_rgb_in"] = occ.numpy(), rgbv.numpy()
    out["sv_occupancy_volume"] = vol["occupancy_volume"].numpy()
    out["sv_rgb_volume"] = vol["rgb_volume"].numpy()
    spec = {
        "save_volume": {
            "top_keys": list(vol.keys()),
            "occupancy_volume": {"dtype": str(vol["occupancy_volume"].dtype), "shape": list(vol["occupancy_volume"].shape)},
            "rgb_volume": {"dtype": str(vol["rgb_volume"].dtype), "shape": list(vol["rgb_volume"].shape)},
            "metadata": {k: v for k, v in vol["metadata"].items() if k != "timestamp"},
            "metadata_keys": list(vol["metadata"].keys()),
            "metadata_in": meta_in, "resolution": Rv, "bounds": [-1.0, 1.0], "threshold": 0.375,
        },
        "state_dict": [[k, list(v.shape), str(v.dtype)] for k, v in ref_model.OptimizedVanillaNeRF(10, 4, 256).state_dict().items()],
    }
    # unsupported suffix -> ValueError (extract_vol.py:205-206)
    try:
        ref_ev.save_volume(occ, rgbv, "/tmp/volume.npy", Rv, (-1.0, 1.0), 0.375, dict(meta_in))
        spec["save_volume"]["bad_suffix_raises"] = None
    except ValueError as e:
        spec["save_volume"]["bad_suffix_raises"] = "ValueError: " + str(e)

    path = os.path.join(HERE, "extract_golden.npz")
    np.savez_compressed(path, **out)
    with open(os.path.join(HERE, "extract_golden.json"), "w") as fh:
        json.dump(spec, fh, indent=1)
    print("wrote", path, f"{os.path.getsize(path) / 1e6:.2f} MB,", len(out), "arrays")


if __name__ == "__main__":
    main()
