"""deepfields-nerf-_b200: B200-native NeRF render/train hot path, drop-in for the Python surface of
mohamedhaiham94/DeepFields-NeRF- (scripts/sampling.py, model.py, train.py, extract_vol.py).

Import as `deepfields_nerf_b200` (see deepfields_nerf_b200.py at the repository root)."""
from . import _lib, ops  # noqa: F401
from .sampling import StratifiedSampler  # noqa: F401
from .model import OptimizedPositionalEncoding, OptimizedVanillaNeRF, get_nerf_models, InstantNGP  # noqa: F401
from .train import (  # noqa: F401
    sample_pdf, raw2outputs, process_rays_in_chunks, forward_pass, FusedTrainer, NGPTrainer, RayStore, onecycle_tables, train,
)
from .extract_vol import (  # noqa: F401
    generate_3d_grid, volume_extraction, apply_otsu_threshold, save_volume, extract_volume_fused, gather_slabs, global_stats,
    slab_range,
)
from .precompute_rays import get_rays, ray_cube_intersection, ray_cube_intersection_numpy, precompute_rays_on_device  # noqa: F401
from .post_process_vol import volume_points, slice_aabb  # noqa: F401
from .render import render_rays, render_image, row_range, gather_rows  # noqa: F401
from .config import load_config  # noqa: F401
from . import torch_ops  # noqa: F401  (registers torch.ops.dfn.*)

__all__ = [
    "StratifiedSampler", "OptimizedPositionalEncoding", "OptimizedVanillaNeRF", "get_nerf_models", "InstantNGP", "sample_pdf",
    "raw2outputs", "process_rays_in_chunks", "forward_pass", "FusedTrainer", "NGPTrainer", "RayStore", "onecycle_tables", "train", "generate_3d_grid",
    "volume_extraction", "apply_otsu_threshold", "save_volume", "extract_volume_fused", "load_config", "ops", "get_rays", "ray_cube_intersection", "ray_cube_intersection_numpy", "precompute_rays_on_device", "volume_points", "slice_aabb", "render_rays", "render_image", "row_range", "gather_rows", "gather_slabs", "global_stats", "slab_range",
]
