"""CPU oracle for the NeRF render/train hot path.  TEST INFRASTRUCTURE ONLY.

This file is a CPU restatement (torch-on-CPU, fp32 unless stated) of the arithmetic of the
reference's hot path.  Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it; the product path under
``deepfields-nerf-_b200/`` never does (it fails loudly when the CUDA library is missing).

Parity status: PINNED for the vanilla path -- every function below is checked against golden
vectors produced by importing the reference itself (``tests/golden/make_golden.py``, run in the
build container where ``/root/reference`` is mounted; outputs committed under ``tests/golden/``).
The Instant-NGP restatement (``oracle/ngp_oracle.py``) is "parity unpinned": its arithmetic lives in
tiny-cuda-nn (commit 74ea173d, un-vendored) and cannot be executed here.

All random draws are explicit arguments (the reference draws internally; order documented in
SURVEY.md section 8c): ``t_rand`` [B,S] for the stratified jitter, ``u`` [B,I] for the inverse-CDF
draw, ``noise`` [P,1] for the sigma noise.

Reference citations are relative to /root/reference/.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# ----------------------------------------------------------------------------------------------
# (a1,a2) stratified sampling -- scripts/sampling.py:5-32
# ----------------------------------------------------------------------------------------------
def stratified_z(near: Tensor, far: Tensor, t_rand: Tensor) -> Tensor:
    """Jittered depths.  sampling.py:8 (t grid), :13 (lerp), :16-18 (bin edges), :21 (jitter)."""
    S = t_rand.shape[-1]
    t = torch.linspace(0.0, 1.0, S, dtype=torch.float32, device=near.device)
    z = near * (1.0 - t[None, :]) + far * t[None, :]
    centre = 0.5 * (z[:, 1:] + z[:, :-1])
    hi = torch.cat([centre, z[:, -1:]], dim=1)
    lo = torch.cat([z[:, :1], centre], dim=1)
    return lo + (hi - lo) * t_rand


def points_on_rays(rays_o: Tensor, rays_d: Tensor, z: Tensor) -> Tuple[Tensor, Tensor]:
    """pts = o + d*z flattened ray-major; dirs broadcast per sample.  sampling.py:24-31 / :43-48."""
    pts = rays_o[:, None, :] + rays_d[:, None, :] * z[:, :, None]
    dirs = rays_d[:, None, :].expand(pts.shape)
    return pts.reshape(-1, 3), dirs.reshape(-1, 3).contiguous()


# ----------------------------------------------------------------------------------------------
# (a3) inverse-CDF sampling -- scripts/train.py:31-75
# ----------------------------------------------------------------------------------------------
def pdf_to_cdf(weights: Tensor) -> Tensor:
    """train.py:44-49.  torch CPU cumsum accumulates in double and rounds every prefix."""
    w = weights + 1e-5
    pdf = w / w.sum(dim=-1, keepdim=True)
    c = torch.cumsum(pdf, dim=-1)
    return torch.cat([torch.zeros_like(c[:, :1]), c], dim=-1)


def invert_cdf(bins: Tensor, cdf: Tensor, u: Tensor) -> Tuple[Tensor, Tensor]:
    """train.py:56-73 given a cdf.  Returns (samples, inds) with inds the raw searchsorted result."""
    n = cdf.shape[-1]
    inds = torch.searchsorted(cdf, u.contiguous(), right=True)
    lo_i = (inds - 1).clamp(min=0)
    hi_i = inds.clamp(max=n - 1)
    c_lo = torch.gather(cdf, 1, lo_i)
    c_hi = torch.gather(cdf, 1, hi_i)
    b_lo = torch.gather(bins, 1, lo_i)
    b_hi = torch.gather(bins, 1, hi_i)
    span = c_hi - c_lo
    span = torch.where(span < 1e-5, torch.ones_like(span), span)
    frac = (u - c_lo) / span
    return b_lo + frac * (b_hi - b_lo), inds


def sample_pdf(bins: Tensor, weights: Tensor, u: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    cdf = pdf_to_cdf(weights)
    samples, inds = invert_cdf(bins, cdf, u)
    return samples, inds, cdf


# ----------------------------------------------------------------------------------------------
# (a4) merge of coarse and fine depths -- scripts/sampling.py:40
# ----------------------------------------------------------------------------------------------
def merge_depths(z_coarse: Tensor, z_fine: Tensor) -> Tensor:
    return torch.sort(torch.cat([z_coarse, z_fine], dim=-1), dim=-1).values


# ----------------------------------------------------------------------------------------------
# (a5) positional encoding -- scripts/model.py:27-53
# ----------------------------------------------------------------------------------------------
def posenc(x: Tensor, L: int, include_input: bool = True) -> Tensor:
    """[x | sin(x_d * 2^k pi) (d-major, k-minor) | cos(...)]   (model.py:30, :48-53)."""
    freqs = ((2.0 ** torch.arange(L)) * torch.pi).to(x.device)  # fp32, as the registered buffer
    ang = (x[..., None] * freqs).flatten(-2)                   # [..., 3L]
    parts = ([x] if include_input else []) + [torch.sin(ang), torch.cos(ang)]
    return torch.cat(parts, dim=-1)


# ----------------------------------------------------------------------------------------------
# (a6-a8) vanilla NeRF MLP -- scripts/model.py:99-187
# ----------------------------------------------------------------------------------------------
def _lin(x: Tensor, w: Tensor, b: Tensor, low: Optional[torch.dtype]) -> Tensor:
    """One nn.Linear.  ``low`` = bf16 reproduces autocast: operands rounded to bf16, fp32
    accumulate, result rounded to bf16 (torch CPU/CUDA autocast semantics for F.linear)."""
    if low is None:
        return F.linear(x, w, b)
    y = F.linear(x.to(low).float(), w.to(low).float(), b.to(low).float())
    return y.to(low)


def mlp_forward(
    sd: Dict[str, Tensor],
    pos: Tensor,
    view: Tensor,
    noise: Optional[Tensor] = None,
    low: Optional[torch.dtype] = None,
    pos_L: int = 10,
    dir_L: int = 4,
) -> Tuple[Tensor, Tensor]:
    """OptimizedVanillaNeRF.forward (model.py:173-187) from a state dict.

    noise: sigma noise added before the ReLU (model.py:160-166); None == eval mode.
    low:   None -> fp32 (train_no_amp.py semantics); torch.bfloat16 -> autocast semantics.
    """
    pe = posenc(pos, pos_L)
    h = pe
    for i in (0, 2, 4, 6):                                       # model.py:99-108
        h = torch.relu(_lin(h, sd[f"pos_layers.{i}.weight"], sd[f"pos_layers.{i}.bias"], low))
    h = torch.cat([h, pe.to(h.dtype)], dim=-1)                   # model.py:154
    for i in (0, 2, 4, 6):                                       # model.py:111-120
        h = torch.relu(_lin(h, sd[f"skip_layers.{i}.weight"], sd[f"skip_layers.{i}.bias"], low))
    s = _lin(h, sd["density_layer.weight"], sd["density_layer.bias"], low)   # model.py:158
    if noise is not None:
        s = s + noise.to(s.dtype)
    sigma = torch.relu(s)                                        # model.py:166
    feat = _lin(h, sd["feature_layer.weight"], sd["feature_layer.bias"], low)  # no activation, :169
    de = posenc(view, dir_L)
    c = torch.cat([feat, de.to(feat.dtype)], dim=-1)             # model.py:182
    c = torch.relu(_lin(c, sd["color_layer.0.weight"], sd["color_layer.0.bias"], low))
    rgb = torch.sigmoid(_lin(c, sd["color_layer.2.weight"], sd["color_layer.2.bias"], low))
    return rgb, sigma


# ----------------------------------------------------------------------------------------------
# (a10) alpha compositing -- scripts/train.py:78-106 (train_no_amp.py:84 uses dist_last=1e6)
# ----------------------------------------------------------------------------------------------
def composite(
    rgb: Tensor, sigma: Tensor, z: Tensor, white_bg: bool = False, dist_last: float = 1e10
) -> Dict[str, Tensor]:
    """Returns rgb_map, weights (the reference's outputs) plus acc and depth (extras whose oracle
    is sum(w) / sum(w*z))."""
    rgb = rgb.float()
    s = sigma.float().reshape(z.shape)
    delta = torch.cat([z[:, 1:] - z[:, :-1], torch.full_like(z[:, :1], dist_last)], dim=1)
    alpha = 1.0 - torch.exp(-s * delta)
    keep = 1.0 - alpha + 1e-10
    trans = torch.cumprod(torch.cat([torch.ones_like(keep[:, :1]), keep], dim=1), dim=1)[:, :-1]
    w = alpha * trans
    acc = w.sum(dim=1)
    rgb_map = (w[:, :, None] * rgb).sum(dim=1)
    if white_bg:
        rgb_map = rgb_map + (1.0 - acc[:, None])
    return {"rgb_map": rgb_map, "weights": w, "acc": acc, "depth": (w * z).sum(dim=1)}


# ----------------------------------------------------------------------------------------------
# (a12) coarse -> fine render + losses -- scripts/train.py:129-186
# ----------------------------------------------------------------------------------------------
def render_losses(
    sd_coarse: Dict[str, Tensor],
    sd_fine: Dict[str, Tensor],
    rays_o: Tensor,
    rays_d: Tensor,
    near: Tensor,
    far: Tensor,
    target: Tensor,
    t_rand: Tensor,
    u: Tensor,
    noise_coarse: Optional[Tensor] = None,
    noise_fine: Optional[Tensor] = None,
    low: Optional[torch.dtype] = None,
    white_bg: bool = False,
    dist_last: float = 1e10,
) -> Dict[str, Tensor]:
    B, S = t_rand.shape
    z = stratified_z(near, far, t_rand)
    pts, dirs = points_on_rays(rays_o, rays_d, z)
    rgb_c, sig_c = mlp_forward(sd_coarse, pts, dirs, noise_coarse, low)
    out_c = composite(rgb_c.reshape(B, S, 3), sig_c.reshape(B, S), z, white_bg, dist_last)
    mids = 0.5 * (z[:, :-1] + z[:, 1:])                                    # train.py:157
    z_new, inds, cdf = sample_pdf(mids, out_c["weights"][:, 1:-1], u)      # train.py:158
    z_new = z_new.detach()
    z_all = merge_depths(z, z_new)
    pts_f, dirs_f = points_on_rays(rays_o, rays_d, z_all)
    rgb_f, sig_f = mlp_forward(sd_fine, pts_f, dirs_f, noise_fine, low)
    T = z_all.shape[1]
    out_f = composite(rgb_f.reshape(B, T, 3), sig_f.reshape(B, T), z_all, white_bg, dist_last)
    lc = F.mse_loss(out_c["rgb_map"], target)
    lf = F.mse_loss(out_f["rgb_map"], target)
    return {
        "loss": lc + lf, "loss_coarse": lc, "loss_fine": lf,
        "z": z, "z_all": z_all, "z_new": z_new, "inds": inds, "cdf": cdf,
        "rgb_map_coarse": out_c["rgb_map"], "rgb_map_fine": out_f["rgb_map"],
        "weights_coarse": out_c["weights"], "weights_fine": out_f["weights"],
    }


# ----------------------------------------------------------------------------------------------
# (a15,a16) dense grid query -- scripts/extract_vol.py:57-120
# ----------------------------------------------------------------------------------------------
def grid_points(bounds: Tuple[float, float], R: int) -> Tensor:
    """ij meshgrid, x slowest / z fastest (extract_vol.py:71-77)."""
    c = torch.linspace(bounds[0], bounds[1], R)
    gx, gy, gz = torch.meshgrid(c, c, c, indexing="ij")
    return torch.stack([gx.reshape(-1), gy.reshape(-1), gz.reshape(-1)], dim=1)


def dense_query(sd, pts: Tensor, view: Tensor, low=None, batch: int = 65536):
    """volume_extraction (extract_vol.py:85-120): eval-mode model on every point, one view."""
    dens, cols = [], []
    for i in range(0, pts.shape[0], batch):
        p = pts[i : i + batch]
        rgb, sig = mlp_forward(sd, p, view[None, :].expand(p.shape[0], 3), None, low)
        dens.append(sig.float())
        cols.append(rgb.float())
    return torch.cat(dens), torch.cat(cols)


# ----------------------------------------------------------------------------------------------
# (a17) Otsu threshold + packing -- scripts/extract_vol.py:123-206
# scikit-image (version unpinned in requirements.txt:8, not installed here) -> "parity unpinned";
# restated from its published algorithm: 256-bin histogram over [min,max], bin centres,
# maximise between-class variance w0*w1*(m0-m1)^2, return the centre of the arg-max bin.
# ----------------------------------------------------------------------------------------------
def otsu_threshold(values: Tensor, nbins: int = 256) -> float:
    v = values.flatten().double()
    lo, hi = float(v.min()), float(v.max())
    if hi == lo:
        return lo
    hist = torch.histc(v, bins=nbins, min=lo, max=hi)
    edges = torch.linspace(lo, hi, nbins + 1, dtype=torch.float64)
    centres = 0.5 * (edges[:-1] + edges[1:])
    w0 = torch.cumsum(hist, 0)
    w1 = torch.cumsum(hist.flip(0), 0).flip(0)
    m0 = torch.cumsum(hist * centres, 0) / w0.clamp(min=1e-300)
    m1 = (torch.cumsum((hist * centres).flip(0), 0) / w1.flip(0).clamp(min=1e-300)).flip(0)
    var12 = w0[:-1] * w1[1:] * (m0[:-1] - m1[1:]) ** 2
    return float(centres[int(torch.argmax(var12))])


def pack_volume(sigma: Tensor, rgb: Tensor, thr: float, R: int):
    """occupancy = sigma > thr as u8; rgb*255 truncated to u8 (extract_vol.py:142,197-198)."""
    occ = (sigma.flatten() > thr).reshape(R, R, R).to(torch.uint8)
    col = (rgb.reshape(R, R, R, 3) * 255).to(torch.uint8)
    return occ, col


# ----------------------------------------------------------------------------------------------
# AdamW + OneCycleLR (train.py:244-259) -- restated for the fused optimiser kernel's parity test
# ----------------------------------------------------------------------------------------------
def onecycle(step: int, total: int, base_lr: float) -> Tuple[float, float]:
    """OneCycleLR(max_lr=10*lr, pct_start=.1, cos, div_factor=10, final_div_factor=100) as
    train.py:250-259 builds it.  Returns (lr, beta1) in force at optimiser step ``step`` (0-based).
    The scheduler's default cycle_momentum=True also anneals Adam's beta1 0.95 -> 0.85 -> 0.95."""
    max_lr = base_lr * 10.0
    init_lr = max_lr / 10.0
    min_lr = init_lr / 100.0
    up_end = float(0.1 * total) - 1.0
    down_end = float(total) - 1.0

    def cos_anneal(a, b, pct):
        return b + (a - b) / 2.0 * (math.cos(math.pi * pct) + 1.0)

    if step <= up_end:
        pct = step / up_end if up_end > 0 else 1.0
        return cos_anneal(init_lr, max_lr, pct), cos_anneal(0.95, 0.85, pct)
    pct = (step - up_end) / (down_end - up_end)
    return cos_anneal(max_lr, min_lr, pct), cos_anneal(0.85, 0.95, pct)


def onecycle_lr(step: int, total: int, base_lr: float) -> float:
    return onecycle(step, total, base_lr)[0]


def adamw_step(p, g, m, v, step, lr, b1=0.9, b2=0.999, eps=1e-8, wd=1e-6):
    """torch.optim.AdamW single-tensor update (step is 1-based)."""
    p = p * (1.0 - lr * wd)
    m = b1 * m + (1.0 - b1) * g
    v = b2 * v + (1.0 - b2) * g * g
    bc1 = 1.0 - b1 ** step
    bc2 = 1.0 - b2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)) + eps
    p = p - (lr / bc1) * m / denom
    return p, m, v


# ----------------------------------------------------------------------------------------------
# (f4) ray generation + ray / cube intersection -- scripts/precompute_rays.py
# (numpy like the reference: fp64 rays, fp32 intersection)
# ----------------------------------------------------------------------------------------------
def camera_rays(H: int, W: int, K, poses, colmap: bool = True, normalize: bool = True):
    """All pixel rays of M cameras, fp64, shape [M,H,W,3] each.  precompute_rays.py:79-80 (axis flip, inverse
    intrinsics), :81-87 (pixel grid, x fastest), :89-92 (back-projection, flip), :94-98 (unit length, norm clipped
    at 1e-8), :103-119 (rotate into the world, origin = camera position)."""
    import numpy as np
    K = np.asarray(K, dtype=np.float64)
    poses = np.asarray(poses, dtype=np.float64)
    flip = np.array([1.0, -1.0, -1.0]) if colmap else np.ones(3)
    rows, cols = np.divmod(np.arange(H * W), W)
    pix = np.stack([cols, rows, np.ones(H * W, dtype=np.int64)])          # [3, H*W]
    cam = flip[:, None] * (np.linalg.inv(K) @ pix)
    if normalize:
        cam = cam / np.maximum(np.sqrt((cam * cam).sum(axis=0, keepdims=True)), 1e-8)
    d = (poses[:, :3, :3] @ cam).transpose(0, 2, 1).reshape(-1, H, W, 3)
    o = np.broadcast_to(poses[:, None, None, :3, 3], d.shape)
    return o, d


def ray_cube_hits(rays_o, rays_d, cube_min: float = -1.0, cube_max: float = 1.0):
    """Slab test in fp32.  precompute_rays.py:28-29 (fp32 cast), :32-37 (|d| < 1e-8 -> sign(d) * 1e-8), :40-52
    (per-axis entry / exit, sorted), :55-56 (interval), :59 (hit flag, BEFORE the clamp), :62 (t_near >= 0).
    numpy's minimum / maximum propagate NaN (0/0 when an origin lies on a slab plane of a zero direction component)."""
    import numpy as np
    o = np.asarray(rays_o, dtype=np.float32)
    d = np.asarray(rays_d, dtype=np.float32)
    eps = np.float32(1e-8)
    d = np.where(np.abs(d) < eps, np.sign(d) * eps, d).astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        ta = (np.float32(cube_min) - o) / d
        tb = (np.float32(cube_max) - o) / d
    lo, hi = np.minimum(ta, tb), np.maximum(ta, tb)
    t_near = lo.max(axis=1)   # ndarray.max propagates NaN like np.maximum
    t_far = hi.min(axis=1)
    with np.errstate(invalid="ignore"):
        hits = (t_near <= t_far) & (t_far >= 0)
    return hits, np.maximum(t_near, np.float32(0.0)), t_far


def precompute_rays(H: int, W: int, K, poses, images_u8, cube_min: float = -1.0, cube_max: float = 1.0):
    """precompute_rays.py:206-228: rays of every pixel, the hits kept in order; colours are u8 / 255 in fp32
    (:131-132).  Returns the dict the reference writes to the .npz (rays fp64 there; fp32 after dataset.py:52-53)."""
    import numpy as np
    o, d = camera_rays(H, W, K, poses)
    o, d = o.reshape(-1, 3), d.reshape(-1, 3)
    hits, t_near, t_far = ray_cube_hits(o, d, cube_min, cube_max)
    rgb = np.asarray(images_u8).reshape(-1, 3).astype(np.float32) / np.float32(255.0)
    return {"rays_o": o[hits], "rays_d": d[hits], "rgbs": rgb[hits], "t_near": t_near[hits], "t_far": t_far[hits]}


# ----------------------------------------------------------------------------------------------
# (f5) occupied voxels as a point set -- scripts/post_process_vol.py
# ----------------------------------------------------------------------------------------------
def volume_points(occupancy, rgb_volume, aabb_min=None, aabb_max=None):
    """post_process_vol.py:82 (colours = u8 volume / its maximum, fp32), :89 (argwhere in C order), :90 (gather),
    :94-95 (index / (R-1) * 2 - 1, fp64), :37-50 (closed-box slice).  occupancy [R,R,R] u8, rgb_volume [R,R,R,3] u8."""
    import numpy as np
    occ = np.asarray(occupancy)
    rgb = np.asarray(rgb_volume)
    R = occ.shape[0]
    idx = np.stack(np.nonzero(occ), axis=1)                       # C order == argwhere
    cols = rgb[idx[:, 0], idx[:, 1], idx[:, 2]].astype(np.float32) / np.float32(rgb.max())
    pts = (idx / (R - 1)) * 2 - 1
    if aabb_min is not None:
        lo, hi = np.asarray(aabb_min, dtype=np.float64), np.asarray(aabb_max, dtype=np.float64)
        keep = np.all((pts >= lo) & (pts <= hi), axis=1)
        pts, cols = pts[keep], cols[keep]
    return pts, cols
