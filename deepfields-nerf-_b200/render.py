"""Inference render of a ray batch / an image (BASELINE config 5): the reference's forward_pass (scripts/train.py:129-186)
without the loss, with optional early ray termination in the fine pass.

    render_rays(model_coarse, model_fine, rays_o, rays_d, near, far, N_samples=64, N_importance=128, ...)
        -> {"rgb_map_coarse", "rgb_map", "acc", "depth", "weights", "z_vals", "evaluated_fraction"}
    render_image(model_coarse, model_fine, H, W, K, pose, ...) -> rgb [H,W,3] (rays from precompute_rays.get_rays, cube hits only)
    row_range(H, rank, world)   the rows a rank renders when an image is sharded over GPUs

Early ray termination: the fine network is queried front to back in `ert_segments` depth segments; after each one the
rays whose transmittance fell below `ert_eps` leave the active list (csrc/ert.cu) and their remaining samples are never
evaluated (they keep sigma = 0, i.e. add nothing in the final compositing pass).  With ert_eps = 0 the result is the
plain coarse + fine render; with ert_eps > 0 it differs from it by at most ert_eps per channel.
"""
from __future__ import annotations

import torch

from . import ops
from .sampling import StratifiedSampler


def row_range(H: int, rank: int, world: int):
    """Contiguous image rows of rank `rank` (the last ranks may get one row less / none)."""
    per = (H + world - 1) // world
    return min(rank * per, H), min((rank + 1) * per, H)


@torch.no_grad()
def render_rays(model_coarse, model_fine, rays_o, rays_d, near, far, N_samples=64, N_importance=128, white_bg=False,
                t_rand=None, u=None, ert_eps=0.0, ert_segments=4, dist_last=1e10):
    dev = rays_o.device
    if dev.type != "cuda":
        raise RuntimeError("deepfields_nerf_b200: render_rays needs CUDA tensors (there is no CPU fallback)")
    model_coarse._sync_packed()
    model_fine._sync_packed()
    B, S, I = rays_o.shape[0], int(N_samples), int(N_importance)
    T = S + I
    near, far = near.reshape(B, 1).float(), far.reshape(B, 1).float()
    if t_rand is None:
        t_rand = torch.rand(B, S, device=dev)   # the reference sampler always jitters (sampling.py:20-21)
    if u is None:
        u = torch.rand(B, I, device=dev)
    sampler = StratifiedSampler(S, dev)
    z_c = ops.stratified_sample(near, far, sampler.t_vals, t_rand)
    rgb_c, sig_c = ops.mlp_forward(model_coarse._packed, rays=(rays_o, rays_d, z_c))
    rgb_map_c, w_c = ops.composite_forward(rgb_c, sig_c, z_c, white_bg, dist_last)
    z_all = ops.sample_merge(z_c, w_c, u)
    evaluated = B * T
    if ert_eps <= 0.0 or ert_segments <= 1:
        rgb_f, sig_f = ops.mlp_forward(model_fine._packed, rays=(rays_o, rays_d, z_all))
    else:
        rgb_f = torch.zeros(B, T, 3, device=dev)
        sig_f = torch.zeros(B, T, device=dev)
        T_run = torch.ones(B, device=dev)
        active = None                                   # None = every ray, in order
        seg = (T + ert_segments - 1) // ert_segments
        evaluated = 0
        for s0 in range(0, T, seg):
            C = min(seg, T - s0)
            n = B if active is None else active.numel()
            if n == 0:
                break
            if active is None:
                o_a, d_a, z_a = rays_o, rays_d, z_all[:, s0:s0 + C].contiguous()
            else:
                ai = active.long()
                o_a, d_a, z_a = rays_o[ai], rays_d[ai], z_all[ai, s0:s0 + C].contiguous()
            seg_rgb, seg_sig = ops.mlp_forward(model_fine._packed, rays=(o_a, d_a, z_a))
            evaluated += n * C
            alive = ops.ert_update(active, seg_rgb, seg_sig.reshape(n, C), z_all, s0, rgb_f, sig_f, T_run, ert_eps, dist_last)
            if s0 + C < T:
                active = ops.compact_indices(alive, active)
    rgb_map, weights, acc, depth = ops.composite_forward(rgb_f, sig_f, z_all, white_bg, dist_last, want_extras=True)
    return {"rgb_map_coarse": rgb_map_c, "rgb_map": rgb_map, "acc": acc, "depth": depth, "weights": weights, "z_vals": z_all,
            "evaluated_fraction": evaluated / float(B * T)}


@torch.no_grad()
def render_image(model_coarse, model_fine, H, W, K, pose, rows=None, background=0.0, draws=None, **kw):
    """Render rows [rows[0], rows[1]) of one camera: pixels whose ray misses the [-1,1]^3 cube get `background`.
    `draws` = (t_rand [H*W,S], u [H*W,I]) optionally fixes the jitter / inverse-CDF draws PER PIXEL, so that a row-sharded
    render reproduces the unsharded image bit for bit."""
    from .precompute_rays import get_rays, ray_cube_intersection
    dev = next(model_fine.parameters()).device
    r0, r1 = rows if rows is not None else (0, H)
    pose = torch.as_tensor(pose, dtype=torch.float64).reshape(1, 4, 4)
    o, d = get_rays(H, W, K, pose, device=dev)
    o, d = o[0, r0:r1].reshape(-1, 3), d[0, r0:r1].reshape(-1, 3)
    hit, t_near, t_far = ray_cube_intersection(o, d)
    idx = ops.compact_indices(hit.to(torch.uint8)).long()
    img = torch.full(((r1 - r0) * W, 3), float(background), device=dev)
    info = {"evaluated_fraction": 0.0, "rays": int(idx.numel())}
    if idx.numel() > 0:
        if draws is not None:
            pix = idx + r0 * W
            kw = dict(kw, t_rand=draws[0][pix].contiguous(), u=draws[1][pix].contiguous())
        out = render_rays(model_coarse, model_fine, o[idx], d[idx], t_near[idx], t_far[idx], **kw)
        img[idx] = out["rgb_map"]
        info["evaluated_fraction"] = out["evaluated_fraction"]
    return img.reshape(r1 - r0, W, 3), info


def gather_rows(part: torch.Tensor, H: int, rank: int, world: int, group=None, dst: int = 0):
    """Assembles the row tiles of a row-sharded render (gather of rgb tiles): rank r's [rows_r, W, 3] tile lands
    at rows row_range(H, r, world) of the [H, W, 3] image.  Returns the image on `dst`, None elsewhere.  One all-gather on
    the ranks' communicator (tiles padded to the tallest one when H % world != 0), no point-to-point traffic."""
    if world == 1:
        return part
    dist = torch.distributed
    W = part.shape[1]
    rows = [r1 - r0 for r0, r1 in (row_range(H, r, world) for r in range(world))]
    if part.shape[0] != rows[rank]:
        raise ValueError("gather_rows: rank %d holds %d rows, its tile has %d" % (rank, part.shape[0], rows[rank]))
    h_max = max(rows)
    part = part.contiguous()
    if min(rows) == h_max:
        img = torch.empty(H, W, 3, dtype=part.dtype, device=part.device)
        dist.all_gather_into_tensor(img.view(-1), part.view(-1), group=group)
        return img if rank == dst else None
    mine = torch.zeros(h_max, W, 3, dtype=part.dtype, device=part.device)
    mine[:rows[rank]].copy_(part)
    every = torch.empty(world, h_max, W, 3, dtype=part.dtype, device=part.device)
    dist.all_gather_into_tensor(every.view(-1), mine.view(-1), group=group)
    if rank != dst:
        return None
    return torch.cat([every[r, :rows[r]] for r in range(world)])
