"""ctypes binding of libdeepfields_b200.so (C ABI declared in include/deepfields_b200.h).

There is no CPU fallback: if the library is missing, or a non-CUDA tensor reaches an op, the call raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdeepfields_b200%s.so" % os.environ.get("DFN_LIB_SUFFIX", ""))   # suffix: bring-up builds only

_p = C.c_void_p
_i = C.c_int
_ll = C.c_longlong
_f = C.c_float
_u = C.c_uint


class PointSource(C.Structure):
    """struct dfn_point_source"""

    _fields_ = [
        ("mode", _i), ("pos", _p), ("view", _p), ("rays_o", _p), ("rays_d", _p), ("z", _p), ("S", _i),
        ("lin", _p), ("R", _i), ("p_offset", _ll),
    ]


class HashGrid(C.Structure):
    """struct dfn_hashgrid"""

    _fields_ = [
        ("scale", _f * 16), ("res", _u * 16), ("size", _u * 16), ("offset", _u * 16), ("n_entries", _ll),
        ("log2_hashmap_size", _i),
    ]


# name -> (restype, argtypes); mirrors include/deepfields_b200.h one to one
SIGNATURES = {
    "dfn_last_error": (C.c_char_p, []),
    "dfn_abi_version": (_i, []),
    "dfn_set_sm_limit": (_i, [_i]),
    "dfn_stratified_sample": (_i, [_p, _p, _p, _p, _i, _i, _p, _p]),
    "dfn_points_on_rays": (_i, [_p, _p, _p, _i, _i, _p, _p, _p]),
    "dfn_sample_pdf": (_i, [_p, _i, _p, _ll, _i, _p, _i, _i, _i, _p, _p, _p, _p]),
    "dfn_sample_pdf_with_cdf": (_i, [_p, _p, _p, _i, _i, _i, _p, _p, _p]),
    "dfn_merge_sort_z": (_i, [_p, _i, _p, _i, _i, _p, _p]),
    "dfn_sample_merge": (_i, [_p, _p, _p, _i, _i, _i, _p, _p, _p]),
    "dfn_stratified_sample_rng": (_i, [_p, _p, _p, C.c_ulonglong, _p, _i, _i, _p, _p]),
    "dfn_sample_merge_rng": (_i, [_p, _p, C.c_ulonglong, _p, _i, _i, _i, _p, _p, _p]),
    "dfn_philox_dump": (_i, [C.c_ulonglong, _p, _u, _ll, _p, _p, _p]),
    "dfn_mlp_forward_rng": (_i, [C.POINTER(PointSource), _ll, _p, _p, C.c_ulonglong, _p, _u, _p, _p, _p, _p]),
    "dfn_posenc": (_i, [_p, _p, _ll, _i, _i, _p, _p]),
    "dfn_composite_forward": (_i, [_p, _p, _p, _i, _i, _i, _i, _f, _p, _p, _p, _p, _p]),
    "dfn_composite_backward": (_i, [_p, _p, _p, _i, _i, _i, _i, _f, _p, _p, _p, _p, _p]),
    "dfn_render_loss": (_i, [_p, _p, _p, _p, _i, _i, _i, _i, _f, _f, _p, _p, _p, _p, _p, _p]),
    "dfn_mlp_fwd_pack_bytes": (_ll, []),
    "dfn_mlp_bwd_pack_bytes": (_ll, []),
    "dfn_mlp_consts_bytes": (_ll, []),
    "dfn_mlp_stash_bytes": (_ll, [_ll]),
    "dfn_mlp_dz_bytes": (_ll, [_ll]),
    "dfn_mlp_pack_weights": (_i, [_p, _p, _p, _p, _p]),
    "dfn_mlp_pack_map_bytes": (_ll, []),
    "dfn_mlp_build_pack_map": (_i, [_p, _p]),
    "dfn_mlp_pack_weights_mapped": (_i, [_p, _p, _p, _p, _p, _p]),
    "dfn_mlp_forward": (_i, [C.POINTER(PointSource), _ll, _p, _p, _p, _p, _p, _p, _p]),
    "dfn_mlp_backward": (_i, [C.POINTER(PointSource), _ll, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "dfn_mlp_backward_dgrad": (_i, [_ll, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "dfn_mlp_backward_wgrad": (_i, [C.POINTER(PointSource), _ll, _p, _p, _p, _p]),
    "dfn_mlp_bwd_fused_ws_bytes": (_ll, [_ll]),
    "dfn_mlp_backward_fused": (_i, [C.POINTER(PointSource), _ll, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "dfn_umma_selftest": (_i, [_p, _p, _i, _i, _i, _i, _u, _u, _u, _u, _u, _u, _p, _p]),
    "dfn_grid_points": (_i, [_p, _i, _ll, _ll, _p, _p]),
    "dfn_histogram": (_i, [_p, _ll, _f, _f, _i, _p, _p]),
    "dfn_pack_volume": (_i, [_p, _p, _ll, _f, _p, _p, _p]),
    "dfn_generate_rays": (_i, [_p, _p, _i, _i, _i, _i, _i, _p, _p, _p]),
    "dfn_ray_cube_intersect": (_i, [_p, _p, _ll, _f, _f, _p, _p, _p, _p]),
    "dfn_precompute_rays_workspace_bytes": (_ll, [_ll]),
    "dfn_precompute_rays": (_i, [_p, _p, _i, _i, _i, _i, _i, _p, _f, _f, _p, _p, _p, _p, _p, _p, _p, _p]),
    "dfn_volume_points_workspace_bytes": (_ll, [_ll]),
    "dfn_volume_points_count": (_i, [_p, _p, _i, _p, _p, _p, _p, _p]),
    "dfn_volume_points_emit": (_i, [_p, _p, _i, _p, _p, _p, _p]),
    "dfn_volume_points": (_i, [_p, _p, _i, _p, _p, _p, _p, _p, _p, _p]),
    "dfn_ert_update": (_i, [_p, _i, _p, _p, _p, _i, _i, _i, _f, _f, _p, _p, _p, _p, _p]),
    "dfn_compact_indices_workspace_bytes": (_ll, [_ll]),
    "dfn_compact_indices": (_i, [_p, _p, _ll, _p, _p, _p, _p]),
    "dfn_compact_indices_static_max_n": (_ll, []),
    "dfn_compact_indices_static": (_i, [_p, _p, _ll, _p, _p, _p]),
    "dfn_hashgrid_init": (_ll, [C.POINTER(HashGrid), _i]),
    "dfn_hashgrid_forward": (_i, [_p, _ll, _p, C.POINTER(HashGrid), _p, _p]),
    "dfn_hashgrid_backward": (_i, [_p, _ll, _p, C.POINTER(HashGrid), _f, _p, _p]),
    "dfn_sh4_encode": (_i, [_p, _ll, _p, _p]),
    "dfn_ngp_forward": (_i, [_p, _p, _i, _ll, _p, C.POINTER(HashGrid), _p, _p, _p, _p, _p]),
    "dfn_ngp_forward_f32": (_i, [_p, _p, _i, _i, _ll, _p, C.POINTER(HashGrid), _p, _p, _p, _p, _p]),
    "dfn_ngp_backward": (_i, [_p, _p, _i, _i, _ll, _p, C.POINTER(HashGrid), _p, _p, _p, _p, _f, _p, _p, _p, _p]),
    "dfn_adamw_step_shadow": (_i, [_p, _p, _p, _p, _ll, _p, _p, _p, _i, _f, _f, _f, _f, _p, _p, _p]),
    "dfn_nonfinite_flag": (_i, [_p, _ll, _p, _p]),
    "dfn_adamw_step": (_i, [_p, _p, _p, _p, _ll, _p, _p, _p, _i, _f, _f, _f, _f, _p]),
}

_lib = None


def load() -> C.CDLL:
    """Loads the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` "
                "(deepfields-nerf-_b200/_build.py).  There is no CPU fallback."
            )
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here == header/library mismatch
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def call(name: str, *args) -> None:
    """Calls an int-returning entry point and raises RuntimeError with dfn_last_error() on failure."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {lib.dfn_last_error().decode()}")
