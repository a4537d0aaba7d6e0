"""Tensor-level wrappers over the C ABI: argument checking, output allocation (torch caching
allocator) and launch on torch's current CUDA stream.  No op has a CPU path."""
from __future__ import annotations

import ctypes as C
import os
import threading
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import PointSource

Tensor = torch.Tensor
MLP_NUM_PARAMS = 595844
TILE = 128


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _chk(t: Tensor, dtype, name: str) -> Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{name}: expected a CUDA tensor (this library has no CPU fallback)")
    if t.dtype != dtype:
        raise TypeError(f"{name}: expected {dtype}, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


def _f32(t: Tensor, name: str) -> Tensor:
    return _chk(t, torch.float32, name)


def _ptr(t: Optional[Tensor]):
    return None if t is None else t.data_ptr()


# ---- sampling ------------------------------------------------------------------------------------
def stratified_sample(near: Tensor, far: Tensor, t_vals: Tensor, t_rand: Tensor) -> Tensor:
    """scripts/sampling.py:13-21 with the torch.rand draw passed in.  near/far [B] or [B,1]."""
    t_rand = _f32(t_rand, "t_rand")
    B, S = t_rand.shape
    near, far = _f32(near.reshape(-1), "near"), _f32(far.reshape(-1), "far")
    assert near.numel() == B and far.numel() == B and t_vals.numel() == S
    z = torch.empty_like(t_rand)
    _lib.call("dfn_stratified_sample", near.data_ptr(), far.data_ptr(), _f32(t_vals, "t_vals").data_ptr(),
              t_rand.data_ptr(), B, S, z.data_ptr(), _stream())
    return z


_sm_limit_now = threading.local()   # mirror of the library's thread-local budget (the C ABI has a setter only)


class sm_limit:
    """`with ops.sm_limit(n):` the MLP launches inside size their persistent grids to n SMs (dfn_set_sm_limit, per calling
    thread; 0 = whole device).  Nests: leaving the block restores the budget that was in force when it was entered."""

    def __init__(self, n_sms: int):
        self.n = int(n_sms)

    def __enter__(self):
        self.prev = getattr(_sm_limit_now, "n", 0)
        _lib.load().dfn_set_sm_limit(self.n)
        _sm_limit_now.n = self.n

    def __exit__(self, *exc):
        _lib.load().dfn_set_sm_limit(self.prev)
        _sm_limit_now.n = self.prev


RNG_JITTER, RNG_PDF, RNG_NOISE_COARSE, RNG_NOISE_FINE = 1, 2, 3, 4   # stream ids of csrc/rng.cuh


def _step_ptr(step: Tensor) -> int:
    assert step.dtype == torch.int32 and step.is_cuda and step.numel() == 1
    return step.data_ptr()


def stratified_sample_rng(near: Tensor, far: Tensor, t_vals: Tensor, seed: int, step: Tensor) -> Tensor:
    """stratified_sample with the jitter drawn in the kernel (Philox: f(seed, step counter on the device, sample))."""
    near, far = _f32(near.reshape(-1), "near"), _f32(far.reshape(-1), "far")
    B, S = near.numel(), t_vals.numel()
    z = torch.empty(B, S, device=near.device, dtype=torch.float32)
    _lib.call("dfn_stratified_sample_rng", near.data_ptr(), far.data_ptr(), _f32(t_vals, "t_vals").data_ptr(), int(seed),
              _step_ptr(step), B, S, z.data_ptr(), _stream())
    return z


def philox_dump(seed: int, step: Tensor, rng_stream: int, n_items: int, normals: bool = False) -> Tensor:
    """The uniforms [n,4] (or normals [n]) the in-kernel draws produce for items 0..n-1 (tests)."""
    out = torch.empty(n_items if normals else (n_items, 4), device=step.device, dtype=torch.float32)
    _lib.call("dfn_philox_dump", int(seed), _step_ptr(step), int(rng_stream), int(n_items), None if normals else out.data_ptr(),
              out.data_ptr() if normals else None, _stream())
    return out


def points_on_rays(rays_o: Tensor, rays_d: Tensor, z: Tensor, want_dirs: bool = True):
    """scripts/sampling.py:24-31.  Returns pts [B*S,3] (and dirs [B*S,3])."""
    z = _f32(z, "z")
    B, S = z.shape
    o, d = _f32(rays_o, "rays_o"), _f32(rays_d, "rays_d")
    pts = torch.empty(B * S, 3, device=z.device, dtype=torch.float32)
    dirs = torch.empty_like(pts) if want_dirs else None
    _lib.call("dfn_points_on_rays", o.data_ptr(), d.data_ptr(), z.data_ptr(), B, S, pts.data_ptr(), _ptr(dirs), _stream())
    return (pts, dirs) if want_dirs else pts


def sample_pdf(bins: Tensor, weights: Tensor, u: Tensor, return_inds: bool = False, return_cdf: bool = False,
               bins_mid_of_z: bool = False):
    """scripts/train.py:31-75 (u = the torch.rand draw of :53).  `weights` may be a last-dim slice of a
    contiguous [B,W] tensor (train.py:158 passes weights[...,1:-1]) -- it is read in place."""
    u = _f32(u, "u")
    B, NS = u.shape
    bins = _f32(bins, "bins")
    NB = bins.shape[-1] - (1 if bins_mid_of_z else 0)
    if weights.dtype != torch.float32 or not weights.is_cuda:
        raise RuntimeError("sample_pdf: weights must be a float32 CUDA tensor")
    if weights.dim() != 2 or weights.stride(1) != 1:
        weights = weights.contiguous()
    assert weights.shape == (B, NB - 1), (weights.shape, B, NB)
    samples = torch.empty(B, NS, device=u.device, dtype=torch.float32)
    inds = torch.empty(B, NS, device=u.device, dtype=torch.int64) if return_inds else None
    cdf = torch.empty(B, NB, device=u.device, dtype=torch.float32) if return_cdf else None
    _lib.call("dfn_sample_pdf", bins.data_ptr(), int(bins_mid_of_z), weights.data_ptr(), weights.stride(0), 0,
              u.data_ptr(), B, NB, NS, samples.data_ptr(), _ptr(inds), _ptr(cdf), _stream())
    out = (samples,) + ((inds,) if return_inds else ()) + ((cdf,) if return_cdf else ())
    return out[0] if len(out) == 1 else out


def sample_pdf_with_cdf(bins: Tensor, cdf: Tensor, u: Tensor) -> Tuple[Tensor, Tensor]:
    """scripts/train.py:56-73 on a given cdf; returns (samples, inds int64)."""
    u, bins, cdf = _f32(u, "u"), _f32(bins, "bins"), _f32(cdf, "cdf")
    B, NS = u.shape
    NB = bins.shape[-1]
    samples = torch.empty(B, NS, device=u.device, dtype=torch.float32)
    inds = torch.empty(B, NS, device=u.device, dtype=torch.int64)
    _lib.call("dfn_sample_pdf_with_cdf", bins.data_ptr(), cdf.data_ptr(), u.data_ptr(), B, NB, NS, samples.data_ptr(),
              inds.data_ptr(), _stream())
    return samples, inds


def merge_sort_z(z_coarse: Tensor, z_fine: Tensor) -> Tensor:
    """torch.sort(torch.cat([z_vals, z_samples], -1), -1).values  (scripts/sampling.py:40)."""
    a, b = _f32(z_coarse, "z_coarse"), _f32(z_fine, "z_fine")
    B = a.shape[0]
    out = torch.empty(B, a.shape[1] + b.shape[1], device=a.device, dtype=torch.float32)
    _lib.call("dfn_merge_sort_z", a.data_ptr(), a.shape[1], b.data_ptr(), b.shape[1], B, out.data_ptr(), _stream())
    return out


def sample_merge(z: Tensor, weights: Tensor, u: Optional[Tensor], want_new: bool = False, n_new: Optional[int] = None,
                 rng: Optional[Tuple[int, Tensor]] = None):
    """sample_pdf(mid-points of z, weights[:, 1:-1], u) + sort(cat(z, samples)) in one launch (scripts/train.py:157-158,
    scripts/sampling.py:40); the new depths never reach HBM unless `want_new`.  Returns z_all [B,NA+NS] (, z_new [B,NS]).
    u = None with rng = (seed, step counter tensor): the n_new draws per ray are made in the kernel (Philox)."""
    z, weights = _f32(z, "z"), _f32(weights, "weights")
    B, NA = z.shape
    NS = int(n_new) if u is None else u.shape[1]
    assert weights.shape == (B, NA)
    z_all = torch.empty(B, NA + NS, device=z.device, dtype=torch.float32)
    z_new = torch.empty(B, NS, device=z.device, dtype=torch.float32) if want_new else None
    if u is None:
        seed, step = rng
        _lib.call("dfn_sample_merge_rng", z.data_ptr(), weights.data_ptr(), int(seed), _step_ptr(step), B, NA, NS, z_all.data_ptr(),
                  _ptr(z_new), _stream())
    else:
        u = _f32(u, "u")
        assert u.shape[0] == B
        _lib.call("dfn_sample_merge", z.data_ptr(), weights.data_ptr(), u.data_ptr(), B, NA, NS, z_all.data_ptr(), _ptr(z_new), _stream())
    return (z_all, z_new) if want_new else z_all


# ---- encoding ------------------------------------------------------------------------------------
def posenc(x: Tensor, freq_bands: Tensor, include_input: bool = True) -> Tensor:
    """scripts/model.py:32-53."""
    shape = x.shape
    x2 = _f32(x.reshape(-1, 3), "x")
    L = freq_bands.numel()
    D = (3 if include_input else 0) + 6 * L
    out = torch.empty(x2.shape[0], D, device=x.device, dtype=torch.float32)
    _lib.call("dfn_posenc", x2.data_ptr(), _f32(freq_bands, "freq_bands").data_ptr(), x2.shape[0], L,
              int(include_input), out.data_ptr(), _stream())
    return out.reshape(*shape[:-1], D)


# ---- compositing ---------------------------------------------------------------------------------
def _comp_inputs(rgb: Tensor, sigma: Tensor, z: Tensor):
    z = _f32(z, "z_vals")
    B, S = z.shape
    if rgb.dtype != sigma.dtype:
        rgb, sigma = rgb.float(), sigma.float()
    if rgb.dtype not in (torch.float32, torch.bfloat16):
        rgb, sigma = rgb.float(), sigma.float()
    rgb = _chk(rgb.reshape(B, S, 3), rgb.dtype, "rgb")
    sigma = _chk(sigma.reshape(B, S), sigma.dtype, "sigma")
    return rgb, sigma, z, B, S, int(rgb.dtype == torch.bfloat16)


def composite_forward(rgb, sigma, z, white_bg=False, dist_last=1e10, want_extras=False):
    rgb, sigma, z, B, S, bf = _comp_inputs(rgb, sigma, z)
    rgb_map = torch.empty(B, 3, device=z.device, dtype=torch.float32)
    weights = torch.empty(B, S, device=z.device, dtype=torch.float32)
    acc = torch.empty(B, device=z.device, dtype=torch.float32) if want_extras else None
    depth = torch.empty(B, device=z.device, dtype=torch.float32) if want_extras else None
    _lib.call("dfn_composite_forward", rgb.data_ptr(), sigma.data_ptr(), z.data_ptr(), B, S, bf, int(white_bg),
              float(dist_last), rgb_map.data_ptr(), weights.data_ptr(), _ptr(acc), _ptr(depth), _stream())
    return (rgb_map, weights, acc, depth) if want_extras else (rgb_map, weights)


def composite_backward(rgb, sigma, z, d_rgb_map, d_weights=None, white_bg=False, dist_last=1e10):
    rgb, sigma, z, B, S, bf = _comp_inputs(rgb, sigma, z)
    d_rgb = torch.empty(B, S, 3, device=z.device, dtype=torch.float32)
    d_sigma = torch.empty(B, S, device=z.device, dtype=torch.float32)
    dm = _f32(d_rgb_map, "d_rgb_map")
    dw = _f32(d_weights, "d_weights") if d_weights is not None else None
    _lib.call("dfn_composite_backward", rgb.data_ptr(), sigma.data_ptr(), z.data_ptr(), B, S, bf, int(white_bg),
              float(dist_last), dm.data_ptr(), _ptr(dw), d_rgb.data_ptr(), d_sigma.data_ptr(), _stream())
    return d_rgb, d_sigma


def render_loss(rgb, sigma, z, target, loss_sum: Tensor, white_bg=False, dist_last=1e10, loss_scale=1.0,
                want_weights=True):
    """Fused raw2outputs + MSELoss(mean) + backward.  loss_sum[0] += sum of squared errors.
    Returns (rgb_map, weights, d_rgb [B,S,3], d_sigma [B,S])."""
    rgb, sigma, z, B, S, bf = _comp_inputs(rgb, sigma, z)
    target = _f32(target, "target")
    rgb_map = torch.empty(B, 3, device=z.device, dtype=torch.float32)
    weights = torch.empty(B, S, device=z.device, dtype=torch.float32) if want_weights else None
    d_rgb = torch.empty(B, S, 3, device=z.device, dtype=torch.float32)
    d_sigma = torch.empty(B, S, device=z.device, dtype=torch.float32)
    _lib.call("dfn_render_loss", rgb.data_ptr(), sigma.data_ptr(), z.data_ptr(), target.data_ptr(), B, S, bf,
              int(white_bg), float(dist_last), float(loss_scale), rgb_map.data_ptr(), _ptr(weights),
              _f32(loss_sum, "loss_sum").data_ptr(), d_rgb.data_ptr(), d_sigma.data_ptr(), _stream())
    return rgb_map, weights, d_rgb, d_sigma


# ---- MLP -------------------------------------------------------------------------------------------
class PackedMLP:
    """bf16 operand images of one OptimizedVanillaNeRF (hidden 256, L 10/4) built from the flat fp32
    parameter vector (reference named_parameters() order)."""

    _maps: Dict[str, Tensor] = {}   # per device: the int32 source map of the packed images (architecture-only, built once)

    def __init__(self, device):
        lib = _lib.load()
        self.fwd = torch.empty(lib.dfn_mlp_fwd_pack_bytes(), dtype=torch.uint8, device=device)
        self.bwd = torch.empty(lib.dfn_mlp_bwd_pack_bytes(), dtype=torch.uint8, device=device)
        self.consts = torch.empty(lib.dfn_mlp_consts_bytes() // 4, dtype=torch.float32, device=device)

    def _map(self) -> Tensor:
        key = str(self.fwd.device)
        m = PackedMLP._maps.get(key)
        if m is None:
            m = torch.empty(_lib.load().dfn_mlp_pack_map_bytes() // 4, dtype=torch.int32, device=self.fwd.device)
            _lib.call("dfn_mlp_build_pack_map", m.data_ptr(), _stream())
            PackedMLP._maps[key] = m
        return m

    def pack(self, flat_params: Tensor) -> None:
        p = _f32(flat_params, "flat_params")
        assert p.numel() == MLP_NUM_PARAMS, p.numel()
        _lib.call("dfn_mlp_pack_weights_mapped", p.data_ptr(), self._map().data_ptr(), self.fwd.data_ptr(), self.bwd.data_ptr(),
                  self.consts.data_ptr(), _stream())


def _source(pos=None, view=None, rays=None, grid=None):
    """Builds the dfn_point_source struct; returns (struct, P, keep-alive list)."""
    s = PointSource()
    keep = []
    if pos is not None:
        pos, view = _f32(pos.reshape(-1, 3), "pos"), _f32(view.reshape(-1, 3), "view")
        assert pos.shape == view.shape
        s.mode, s.pos, s.view = 0, pos.data_ptr(), view.data_ptr()
        keep += [pos, view]
        P = pos.shape[0]
    elif rays is not None:
        o, d, z = (_f32(t, n) for t, n in zip(rays, ("rays_o", "rays_d", "z")))
        B, S = z.shape
        s.mode, s.rays_o, s.rays_d, s.z, s.S = 1, o.data_ptr(), d.data_ptr(), z.data_ptr(), S
        keep += [o, d, z]
        P = B * S
    else:
        lin, view, start, count = grid
        lin, view = _f32(lin, "lin"), _f32(view.reshape(3), "view")
        s.mode, s.lin, s.view, s.R, s.p_offset = 2, lin.data_ptr(), view.data_ptr(), lin.numel(), int(start)
        keep += [lin, view]
        P = int(count)
    return s, P, keep


def mlp_num_tiles(P: int) -> int:
    """128-point tiles the MLP kernels lay the stash / dz scratch out for (whole 256-point pairs, mlp_common.cuh)."""
    return 2 * ((P + 2 * TILE - 1) // (2 * TILE))


def mlp_stash_bytes(P: int) -> int:
    return _lib.load().dfn_mlp_stash_bytes(P)


def mlp_dz_bytes(P: int) -> int:
    return _lib.load().dfn_mlp_dz_bytes(P)


def mlp_forward(packed: PackedMLP, *, pos=None, view=None, rays=None, grid=None, noise: Optional[Tensor] = None,
                stash: Optional[Tensor] = None, out: Optional[Tuple[Tensor, Tensor]] = None,
                noise_rng: Optional[Tuple[int, Tensor, int]] = None):
    """OptimizedVanillaNeRF.forward (scripts/model.py:173-187).  Returns (rgb [P,3], sigma [P]) fp32.
    noise_rng = (seed, step counter tensor, stream id): the sigma noise of model.py:160-162 is drawn in the kernel."""
    s, P, keep = _source(pos, view, rays, grid)
    dev = packed.fwd.device
    if out is None:
        rgb = torch.empty(P, 3, device=dev, dtype=torch.float32)
        sigma = torch.empty(P, device=dev, dtype=torch.float32)
    else:
        rgb, sigma = out
    if noise is not None:
        noise = _f32(noise.reshape(-1), "noise")
        assert noise.numel() == P
    if stash is not None:
        assert stash.is_cuda and stash.dtype == torch.uint8 and stash.numel() >= mlp_stash_bytes(P)
    if noise is None and noise_rng is not None:
        seed, step, sid = noise_rng
        _lib.call("dfn_mlp_forward_rng", C.byref(s), P, packed.fwd.data_ptr(), packed.consts.data_ptr(), int(seed), _step_ptr(step),
                  int(sid), rgb.data_ptr(), sigma.data_ptr(), _ptr(stash), _stream())
        return rgb, sigma
    _lib.call("dfn_mlp_forward", C.byref(s), P, packed.fwd.data_ptr(), packed.consts.data_ptr(), _ptr(noise),
              rgb.data_ptr(), sigma.data_ptr(), _ptr(stash), _stream())
    return rgb, sigma


def mlp_bwd_fused_ws_bytes(P: int) -> int:
    return _lib.load().dfn_mlp_bwd_fused_ws_bytes(P)


def bwd_fused_default() -> bool:
    """The one-launch backward (csrc/mlp_bwd_fused.cu) is an EXPERIMENT that measured slower than the two launches
    (DESIGN.md section 5); DFN_BWD_FUSED=1 selects it."""
    return os.environ.get("DFN_BWD_FUSED", "0") == "1"


def mlp_backward_fused(packed: PackedMLP, rgb: Tensor, sigma: Tensor, d_rgb: Tensor, d_sigma: Tensor, stash: Tensor,
                       ws: Tensor, grads: Tensor, *, pos=None, view=None, rays=None, grid=None) -> None:
    """mlp_backward as ONE launch: dgrad pairs and wgrad units run concurrently, dZ tiles pass through an L2-resident ring
    in `ws` (mlp_bwd_fused_ws_bytes(P) bytes) instead of an HBM scratch."""
    s, P, keep = _source(pos, view, rays, grid)
    assert grads.numel() == MLP_NUM_PARAMS and grads.dtype == torch.float32 and grads.is_cuda
    assert ws.is_cuda and ws.dtype == torch.uint8 and ws.numel() >= mlp_bwd_fused_ws_bytes(P)
    _lib.call("dfn_mlp_backward_fused", C.byref(s), P, packed.bwd.data_ptr(), packed.consts.data_ptr(),
              _f32(rgb.reshape(-1, 3), "rgb").data_ptr(), _f32(sigma.reshape(-1), "sigma").data_ptr(),
              _f32(d_rgb.reshape(-1, 3), "d_rgb").data_ptr(), _f32(d_sigma.reshape(-1), "d_sigma").data_ptr(),
              stash.data_ptr(), ws.data_ptr(), grads.data_ptr(), _stream())


def mlp_backward(packed: PackedMLP, rgb: Tensor, sigma: Tensor, d_rgb: Tensor, d_sigma: Tensor, stash: Tensor,
                 dz: Tensor, grads: Tensor, *, pos=None, view=None, rays=None, grid=None, mark=None, dgrad_sms: int = 0) -> None:
    """Accumulates d loss / d params into `grads` (flat fp32, reference parameter order).  Two launches: the
    data-gradient chain, then the weight-gradient reduction; `mark(name)` (optional) is called between them.
    dgrad_sms > 0: the data-gradient launch leaves the other SMs free (a collective running beside it needs somewhere to live)."""
    s, P, keep = _source(pos, view, rays, grid)
    assert grads.numel() == MLP_NUM_PARAMS and grads.dtype == torch.float32 and grads.is_cuda
    assert dz.numel() >= mlp_dz_bytes(P)
    with sm_limit(dgrad_sms if dgrad_sms > 0 else getattr(_sm_limit_now, "n", 0)):
        _lib.call("dfn_mlp_backward_dgrad", P, packed.bwd.data_ptr(), packed.consts.data_ptr(),
                  _f32(rgb.reshape(-1, 3), "rgb").data_ptr(), _f32(sigma.reshape(-1), "sigma").data_ptr(),
                  _f32(d_rgb.reshape(-1, 3), "d_rgb").data_ptr(), _f32(d_sigma.reshape(-1), "d_sigma").data_ptr(),
                  stash.data_ptr(), dz.data_ptr(), _stream())
    if mark is not None:
        mark("dgrad")
    _lib.call("dfn_mlp_backward_wgrad", C.byref(s), P, stash.data_ptr(), dz.data_ptr(), grads.data_ptr(), _stream())
    if mark is not None:
        mark("wgrad")


def umma_selftest(a_img: Tensor, b_img: Tensor, N: int, K: int, a_mn: int, b_mn: int, a_lbo: int, a_sbo: int,
                  b_lbo: int, b_sbo: int, a_kstep: int, b_kstep: int) -> Tensor:
    d = torch.empty(128, N, device=a_img.device, dtype=torch.float32)
    _lib.call("dfn_umma_selftest", a_img.data_ptr(), b_img.data_ptr(), N, K, a_mn, b_mn, a_lbo, a_sbo, b_lbo, b_sbo,
              a_kstep, b_kstep, d.data_ptr(), _stream())
    return d


# ---- dense grid / volume ---------------------------------------------------------------------------
def grid_points(lin: Tensor, start: int, count: int) -> Tensor:
    lin = _f32(lin, "lin")
    out = torch.empty(count, 3, device=lin.device, dtype=torch.float32)
    _lib.call("dfn_grid_points", lin.data_ptr(), lin.numel(), int(start), int(count), out.data_ptr(), _stream())
    return out


def histogram(v: Tensor, lo: float, hi: float, nbins: int = 256) -> Tensor:
    v = _f32(v.reshape(-1), "values")
    hist = torch.zeros(nbins, device=v.device, dtype=torch.int64)
    _lib.call("dfn_histogram", v.data_ptr(), v.numel(), float(lo), float(hi), nbins, hist.data_ptr(), _stream())
    return hist


def pack_volume(sigma: Tensor, rgb: Tensor, thr: float) -> Tuple[Tensor, Tensor]:
    sigma, rgb = _f32(sigma.reshape(-1), "sigma"), _f32(rgb.reshape(-1, 3), "rgb")
    n = sigma.numel()
    occ = torch.empty(n, device=sigma.device, dtype=torch.uint8)
    col = torch.empty(n, 3, device=sigma.device, dtype=torch.uint8)
    _lib.call("dfn_pack_volume", sigma.data_ptr(), rgb.data_ptr(), n, float(thr), occ.data_ptr(), col.data_ptr(), _stream())
    return occ, col


# ---- ray generation (scripts/precompute_rays.py) ---------------------------------------------------
def _camera_args(K, poses, device):
    import ctypes
    import numpy as np
    kinv = np.ascontiguousarray(np.linalg.inv(np.asarray(K, dtype=np.float64)))   # precompute_rays.py:80 (host, 3x3)
    if kinv.shape != (3, 3):
        raise ValueError("K must be 3x3")
    poses = torch.as_tensor(np.asarray(poses, dtype=np.float64) if not isinstance(poses, Tensor) else poses)
    poses = poses.to(device=device, dtype=torch.float64).contiguous()
    if poses.dim() != 3 or tuple(poses.shape[1:]) != (4, 4):
        raise ValueError("camera_poses must be [M,4,4]")
    return kinv, kinv.ctypes.data_as(ctypes.c_void_p), poses


def generate_rays(H: int, W: int, K, poses, colmap: bool = True, normalize: bool = True, device="cuda") -> Tuple[Tensor, Tensor]:
    """get_rays: fp32 [M,H,W,3] origins and directions on the device."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("deepfields_nerf_b200: generate_rays needs a CUDA device (there is no CPU fallback)")
    kinv, kptr, poses = _camera_args(K, poses, device)
    M = poses.shape[0]
    o = torch.empty(M, H, W, 3, device=device, dtype=torch.float32)
    d = torch.empty_like(o)
    with torch.cuda.device(device):
        _lib.call("dfn_generate_rays", kptr, poses.data_ptr(), M, int(H), int(W), int(bool(colmap)), int(bool(normalize)),
                  o.data_ptr(), d.data_ptr(), _stream())
    return o, d


def ray_cube_intersect(rays_o: Tensor, rays_d: Tensor, cube_min: float = -1.0, cube_max: float = 1.0):
    rays_o, rays_d = _f32(rays_o.reshape(-1, 3), "ray_origins"), _f32(rays_d.reshape(-1, 3), "ray_directions")
    n = rays_o.shape[0]
    if rays_d.shape[0] != n:
        raise ValueError("ray_origins and ray_directions differ in length")
    hit = torch.empty(n, device=rays_o.device, dtype=torch.uint8)
    t_near = torch.empty(n, device=rays_o.device, dtype=torch.float32)
    t_far = torch.empty_like(t_near)
    _lib.call("dfn_ray_cube_intersect", rays_o.data_ptr(), rays_d.data_ptr(), n, float(cube_min), float(cube_max), hit.data_ptr(),
              t_near.data_ptr(), t_far.data_ptr(), _stream())
    return hit.bool(), t_near, t_far


def precompute_rays(H: int, W: int, K, poses, images: Optional[Tensor] = None, cube_min: float = -1.0, cube_max: float = 1.0,
                    colmap: bool = True, normalize: bool = True, device="cuda") -> Dict[str, Tensor]:
    """Generate + intersect + keep the hits in order (one pass over HBM); images = u8 [M,H,W,3] on the device or None."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("deepfields_nerf_b200: precompute_rays needs a CUDA device (there is no CPU fallback)")
    kinv, kptr, poses = _camera_args(K, poses, device)
    M = poses.shape[0]
    n = M * int(H) * int(W)
    if images is not None:
        if images.dtype != torch.uint8 or images.numel() != n * 3:
            raise ValueError("images must be uint8 [M,H,W,3]")
        images = images.to(device).contiguous()
    o = torch.empty(n, 3, device=device, dtype=torch.float32)
    d = torch.empty_like(o)
    rgb = torch.empty_like(o) if images is not None else None
    t_near = torch.empty(n, device=device, dtype=torch.float32)
    t_far = torch.empty_like(t_near)
    count = torch.zeros(1, device=device, dtype=torch.int64)
    ws = torch.empty(int(_lib.load().dfn_precompute_rays_workspace_bytes(n)), device=device, dtype=torch.uint8)
    with torch.cuda.device(device):
        _lib.call("dfn_precompute_rays", kptr, poses.data_ptr(), M, int(H), int(W), int(bool(colmap)), int(bool(normalize)),
                  images.data_ptr() if images is not None else None, float(cube_min), float(cube_max), o.data_ptr(), d.data_ptr(),
                  rgb.data_ptr() if rgb is not None else None, t_near.data_ptr(), t_far.data_ptr(), count.data_ptr(), ws.data_ptr(),
                  _stream())
    k = int(count.item())
    out = {"rays_o": o[:k], "rays_d": d[:k], "t_near": t_near[:k], "t_far": t_far[:k]}
    if rgb is not None:
        out["rgbs"] = rgb[:k]
    return out


# ---- occupied voxels as points (scripts/post_process_vol.py) -----------------------------------------
def volume_points(occupancy: Tensor, rgb_volume: Tensor, aabb_min=None, aabb_max=None) -> Tuple[Tensor, Tensor]:
    """argwhere + colour gather + normalised coordinates (+ closed AABB slice): points [N,3] fp64, rgbs [N,3] fp32."""
    import ctypes
    import numpy as np
    if not occupancy.is_cuda or not rgb_volume.is_cuda:
        raise RuntimeError("deepfields_nerf_b200: volume_points needs CUDA tensors (there is no CPU fallback)")
    if occupancy.dim() != 3 or len(set(occupancy.shape)) != 1:
        raise ValueError("occupancy_volume must be [R,R,R]")
    R = occupancy.shape[0]
    if tuple(rgb_volume.shape) != (R, R, R, 3) or rgb_volume.dtype != torch.uint8:
        raise ValueError("rgb_volume must be uint8 [R,R,R,3]")
    occ = occupancy.to(torch.uint8).contiguous()
    rgb = rgb_volume.contiguous()
    n = R ** 3
    lo = hi = None
    if (aabb_min is None) != (aabb_max is None):
        raise ValueError("aabb_min and aabb_max go together")
    if aabb_min is not None:
        lo = np.ascontiguousarray(np.asarray(aabb_min, dtype=np.float64).reshape(3))
        hi = np.ascontiguousarray(np.asarray(aabb_max, dtype=np.float64).reshape(3))
    dev = occ.device
    count = torch.zeros(1, device=dev, dtype=torch.int64)
    ws = torch.empty(int(_lib.load().dfn_volume_points_workspace_bytes(n)), device=dev, dtype=torch.uint8)
    with torch.cuda.device(dev):
        _lib.call("dfn_volume_points_count", occ.data_ptr(), rgb.data_ptr(), R,
                  lo.ctypes.data_as(ctypes.c_void_p) if lo is not None else None,
                  hi.ctypes.data_as(ctypes.c_void_p) if hi is not None else None, count.data_ptr(), ws.data_ptr(), _stream())
        k = int(count.item())                      # the outputs are sized by the count, not by R^3
        points = torch.empty(k, 3, device=dev, dtype=torch.float64)
        rgbs = torch.empty(k, 3, device=dev, dtype=torch.float32)
        if k:
            _lib.call("dfn_volume_points_emit", occ.data_ptr(), rgb.data_ptr(), R, points.data_ptr(), rgbs.data_ptr(), ws.data_ptr(),
                      _stream())
    return points, rgbs


# ---- early ray termination (inference render) --------------------------------------------------------
def ert_update(active: Optional[Tensor], seg_rgb: Tensor, seg_sigma: Tensor, z: Tensor, s0: int, rgb_full: Tensor,
               sigma_full: Tensor, T_run: Tensor, eps: float, dist_last: float = 1e10) -> Tensor:
    """Scatter one depth segment of the active rays into the full arrays, update their transmittance; returns alive [n] u8."""
    z = _f32(z, "z_vals")
    B, S = z.shape
    n, C = seg_sigma.shape
    seg_rgb, seg_sigma = _f32(seg_rgb.reshape(n, C, 3), "segment rgb"), _f32(seg_sigma, "segment sigma")
    if active is not None:
        active = _chk(active, torch.int32, "active ray list")
        assert active.numel() == n
    assert rgb_full.shape == (B, S, 3) and sigma_full.shape == (B, S) and T_run.shape == (B,)
    alive = torch.empty(n, device=z.device, dtype=torch.uint8)
    _lib.call("dfn_ert_update", _ptr(active), n, seg_rgb.data_ptr(), seg_sigma.data_ptr(), z.data_ptr(), S, int(s0), C,
              float(dist_last), float(eps), _f32(rgb_full, "rgb").data_ptr(), _f32(sigma_full, "sigma").data_ptr(),
              _f32(T_run, "T").data_ptr(), alive.data_ptr(), _stream())
    return alive


def compact_indices(flags: Tensor, idx_in: Optional[Tensor] = None) -> Tensor:
    """idx_in[flags != 0] in order (idx_in None: the positions), int32."""
    flags = _chk(flags.reshape(-1), torch.uint8, "flags")
    n = flags.numel()
    if idx_in is not None:
        idx_in = _chk(idx_in, torch.int32, "index list")
        assert idx_in.numel() == n
    out = torch.empty(n, device=flags.device, dtype=torch.int32)
    count = torch.zeros(1, device=flags.device, dtype=torch.int64)
    ws = torch.empty(max(int(_lib.load().dfn_compact_indices_workspace_bytes(n)), 16), device=flags.device, dtype=torch.uint8)
    _lib.call("dfn_compact_indices", flags.data_ptr(), _ptr(idx_in), n, out.data_ptr(), count.data_ptr(), ws.data_ptr(), _stream())
    return out[:int(count.item())]


# ---- optimiser -------------------------------------------------------------------------------------
def adamw_step(p: Tensor, g: Tensor, m: Tensor, v: Tensor, lr_table: Tensor, beta1_table: Tensor, step: Tensor,
               beta2=0.999, eps=1e-8, weight_decay=1e-6, grad_scale=1.0) -> None:
    assert step.dtype == torch.int32 and step.is_cuda
    _lib.call("dfn_adamw_step", _f32(p, "p").data_ptr(), _f32(g, "g").data_ptr(), _f32(m, "m").data_ptr(),
              _f32(v, "v").data_ptr(), p.numel(), _f32(lr_table, "lr_table").data_ptr(),
              _f32(beta1_table, "beta1_table").data_ptr(), step.data_ptr(), lr_table.numel(), float(beta2), float(eps),
              float(weight_decay), float(grad_scale), _stream())


def adamw_step_shadow(p: Tensor, g: Tensor, m: Tensor, v: Tensor, lr_table: Tensor, beta1_table: Tensor, step: Tensor,
                      shadow: Tensor, skip_flag: Optional[Tensor] = None, beta2=0.999, eps=1e-8, weight_decay=1e-6,
                      grad_scale=1.0) -> None:
    """adamw_step that also refreshes `shadow` (fp16 copy of p) and is skipped when skip_flag[0] != 0 (nonfinite_flag)."""
    assert step.dtype == torch.int32 and step.is_cuda and shadow.dtype == torch.float16 and shadow.numel() == p.numel()
    _lib.call("dfn_adamw_step_shadow", _f32(p, "p").data_ptr(), _f32(g, "g").data_ptr(), _f32(m, "m").data_ptr(),
              _f32(v, "v").data_ptr(), p.numel(), _f32(lr_table, "lr_table").data_ptr(),
              _f32(beta1_table, "beta1_table").data_ptr(), step.data_ptr(), lr_table.numel(), float(beta2), float(eps),
              float(weight_decay), float(grad_scale), shadow.data_ptr(), _ptr(skip_flag), _stream())


def nonfinite_flag(g: Tensor, flag: Tensor) -> None:
    """flag[0] (int32, device) = 1 iff g holds an inf / NaN: GradScaler's guard in front of optimizer.step."""
    assert flag.dtype == torch.int32 and flag.is_cuda
    _lib.call("dfn_nonfinite_flag", _f32(g, "g").data_ptr(), g.numel(), flag.data_ptr(), _stream())


# ---- Instant-NGP variant (scripts/instantngp.py) ---------------------------------------------------
NGP_LEVELS = 16
NGP_DMLP_PARAMS = 64 * 32 + 16 * 64
NGP_CMLP_PARAMS = 64 * 32 + 64 * 64 + 16 * 64


def hashgrid_levels(log2_hashmap_size: int) -> "_lib.HashGrid":
    """Level table of the reference's HashGrid config (instantngp.py:23-33).  Host only."""
    g = _lib.HashGrid()
    n = _lib.load().dfn_hashgrid_init(C.byref(g), int(log2_hashmap_size))
    if n < 0:
        raise ValueError(_lib.load().dfn_last_error().decode())
    return g


def _f16(t: Tensor, name: str) -> Tensor:
    return _chk(t, torch.float16, name)


def _pos(t: Tensor, name: str) -> Tensor:
    t = _f32(t, name)
    if t.dim() != 2 or t.shape[1] != 3:
        raise ValueError(f"{name}: expected [P,3], got {tuple(t.shape)}")
    return t


def hashgrid_forward(pos: Tensor, table: Tensor, grid) -> Tensor:
    """tcnn HashGrid encode (instantngp.py:77): pos [P,3] fp32, table fp16 [n_entries*2] -> fp16 [P,32]."""
    pos, table = _pos(pos, "pos"), _f16(table, "table")
    if table.numel() != 2 * grid.n_entries:
        raise ValueError(f"table: expected {2 * grid.n_entries} values, got {table.numel()}")
    out = torch.empty(pos.shape[0], 2 * NGP_LEVELS, device=pos.device, dtype=torch.float16)
    _lib.call("dfn_hashgrid_forward", pos.data_ptr(), pos.shape[0], table.data_ptr(), C.byref(grid), out.data_ptr(), _stream())
    return out


def hashgrid_backward(pos: Tensor, d_out: Tensor, grid, grad_scale: float = 1.0, d_table: Optional[Tensor] = None) -> Tensor:
    """Scatter-add backward of the encode: returns fp32 [n_entries*2] (accumulates into d_table when given)."""
    pos, d_out = _pos(pos, "pos"), _f16(d_out, "d_out")
    if d_out.shape != (pos.shape[0], 2 * NGP_LEVELS):
        raise ValueError(f"d_out: expected {(pos.shape[0], 2 * NGP_LEVELS)}, got {tuple(d_out.shape)}")
    if d_table is None:
        d_table = torch.zeros(2 * grid.n_entries, device=pos.device, dtype=torch.float32)
    else:
        d_table = _f32(d_table, "d_table")
    _lib.call("dfn_hashgrid_backward", pos.data_ptr(), pos.shape[0], d_out.data_ptr(), C.byref(grid), float(grad_scale),
              d_table.data_ptr(), _stream())
    return d_table


def sh4_encode(dirs: Tensor) -> Tensor:
    """tcnn SphericalHarmonics degree 4 (instantngp.py:34-40, :98): dirs [P,3] fp32 in [0,1] -> fp16 [P,16]."""
    dirs = _pos(dirs, "dirs")
    out = torch.empty(dirs.shape[0], 16, device=dirs.device, dtype=torch.float16)
    _lib.call("dfn_sh4_encode", dirs.data_ptr(), dirs.shape[0], out.data_ptr(), _stream())
    return out


def ngp_forward(pos: Tensor, dirs: Optional[Tensor], table: Tensor, grid, dmlp_w: Tensor, cmlp_w: Tensor,
                want_features: bool = True, want_rgb: bool = True):
    """InstantNGP.forward / query_density as one launch (inference).  dirs [P,3] or a single [3] / [1,3] view.
    Returns (rgb fp16 [P,3] | None, features fp16 [P,16] | None)."""
    pos, table = _pos(pos, "pos"), _f16(table, "table")
    dmlp_w, cmlp_w = _f16(dmlp_w, "dmlp_w"), _f16(cmlp_w, "cmlp_w")
    if table.numel() != 2 * grid.n_entries or dmlp_w.numel() != NGP_DMLP_PARAMS or cmlp_w.numel() != NGP_CMLP_PARAMS:
        raise ValueError("ngp_forward: parameter sizes do not match the reference configuration")
    P = pos.shape[0]
    stride = 3
    if want_rgb:
        if dirs is None:
            raise ValueError("ngp_forward: rgb needs view directions")
        dirs = _f32(dirs, "dirs")
        if dirs.numel() == 3:
            stride, dirs = 0, dirs.reshape(3)
        elif dirs.shape != (P, 3):
            raise ValueError(f"dirs: expected {(P, 3)} or a single direction, got {tuple(dirs.shape)}")
    feat = torch.empty(P, 16, device=pos.device, dtype=torch.float16) if want_features else None
    rgb = torch.empty(P, 3, device=pos.device, dtype=torch.float16) if want_rgb else None
    _lib.call("dfn_ngp_forward", pos.data_ptr(), _ptr(dirs) if want_rgb else None, stride, P, table.data_ptr(), C.byref(grid),
              dmlp_w.data_ptr(), cmlp_w.data_ptr(), _ptr(feat), _ptr(rgb), _stream())
    return rgb, feat


def _dirs_args(dirs: Tensor, P: int, dirs_group: int):
    dirs = _f32(dirs, "dirs")
    if dirs.numel() == 3:
        return dirs.reshape(3), 0, 1
    assert dirs.dim() == 2 and dirs.shape[1] == 3 and dirs.shape[0] * dirs_group == P, (dirs.shape, P, dirs_group)
    return dirs, 3, int(dirs_group)


def ngp_forward_f32(pos: Tensor, dirs: Tensor, table: Tensor, grid, dmlp_w: Tensor, cmlp_w: Tensor, dirs_group: int = 1,
                    out: Optional[Tuple[Tensor, Tensor]] = None) -> Tuple[Tensor, Tensor]:
    """InstantNGP.forward as ONE launch with fp32 outputs for the compositing kernels: rgb [P,3], sigma [P] (fp16 values).
    dirs: [P / dirs_group, 3] (dirs_group consecutive points share a direction: the samples of a ray) or one direction [3]."""
    pos, table = _pos(pos, "pos"), _f16(table, "table")
    P = pos.shape[0]
    d, stride, group = _dirs_args(dirs, P, dirs_group)
    rgb, sigma = out if out is not None else (torch.empty(P, 3, device=pos.device), torch.empty(P, device=pos.device))
    _lib.call("dfn_ngp_forward_f32", pos.data_ptr(), d.data_ptr(), stride, group, P, table.data_ptr(), C.byref(grid),
              _f16(dmlp_w, "dmlp_w").data_ptr(), _f16(cmlp_w, "cmlp_w").data_ptr(), rgb.data_ptr(), sigma.data_ptr(), _stream())
    return rgb, sigma


def ngp_backward(pos: Tensor, dirs: Tensor, table: Tensor, grid, dmlp_w: Tensor, cmlp_w: Tensor, d_rgb: Tensor, d_sigma: Tensor,
                 d_table: Tensor, d_dmlp: Tensor, d_cmlp: Tensor, dirs_group: int = 1, loss_scale: float = 128.0) -> None:
    """Accumulates d loss / d (table, density MLP, colour MLP) (fp32) from d loss / d rgb [P,3], d loss / d sigma [P]: ONE
    launch, the forward is recomputed in registers (csrc/ngp_train.cu)."""
    pos, table = _pos(pos, "pos"), _f16(table, "table")
    P = pos.shape[0]
    d, stride, group = _dirs_args(dirs, P, dirs_group)
    assert d_table.numel() == 2 * grid.n_entries and d_dmlp.numel() == NGP_DMLP_PARAMS and d_cmlp.numel() == NGP_CMLP_PARAMS
    _lib.call("dfn_ngp_backward", pos.data_ptr(), d.data_ptr(), stride, group, P, table.data_ptr(), C.byref(grid),
              _f16(dmlp_w, "dmlp_w").data_ptr(), _f16(cmlp_w, "cmlp_w").data_ptr(), _f32(d_rgb.reshape(-1, 3), "d_rgb").data_ptr(),
              _f32(d_sigma.reshape(-1), "d_sigma").data_ptr(), float(loss_scale), _f32(d_table, "d_table").data_ptr(),
              _f32(d_dmlp, "d_dmlp").data_ptr(), _f32(d_cmlp, "d_cmlp").data_ptr(), _stream())
