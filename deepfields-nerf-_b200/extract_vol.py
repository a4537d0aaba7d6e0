"""Drop-in for the reference's scripts/extract_vol.py: dense density/colour query of an AABB grid, Otsu
threshold, `volume.pth` in the reference layout.  The grid is never materialised on the fused path
(coordinates are generated in the MLP kernel from the flat index); slabs of x-planes shard across ranks."""
from __future__ import annotations

import argparse
import datetime
import os
import time
from pathlib import Path
from typing import Optional, Tuple

import torch

from . import ops
from .config import load_config
from .model import InstantNGP, OptimizedVanillaNeRF

# a collective that cannot complete (a rank that died, a mismatched call) fails after this long instead of holding the GPUs
# for NCCL's default ten minutes
_PG_TIMEOUT = datetime.timedelta(seconds=180)


def generate_3d_grid(bounds: Tuple[float, float], resolution: int, device="cuda") -> torch.Tensor:
    """scripts/extract_vol.py:57-82: [R^3,3], ij order (x slowest, z fastest).  Built on the device from the
    same torch.linspace values the reference meshgrids."""
    lin = torch.linspace(bounds[0], bounds[1], resolution, device=device)
    return ops.grid_points(lin, 0, resolution ** 3)


def volume_extraction(model, points, canonical_view, batch_size: int = 8192, device: str = "cuda",
                      show_progress: bool = True):
    """scripts/extract_vol.py:85-120 -> (densities [N,1], rgbs [N,3]).  One persistent launch replaces the
    reference's N/batch_size (H2D, model, 2x D2H) iterations; outputs stay on the device."""
    model.eval()
    with torch.no_grad():
        pts = points.to(device)
        view = canonical_view.to(device).float().reshape(1, 3).expand(pts.shape[0], 3).contiguous()
        rgb, sigma = model(pts, view)
    return sigma, rgb


def otsu_from_histogram(hist: torch.Tensor, lo: float, hi: float) -> float:
    """skimage.filters.threshold_otsu on a 256-bin histogram over [lo,hi] (bin centres, maximise the
    between-class variance w0*w1*(m0-m1)^2).  scikit-image is unpinned and absent: parity unpinned."""
    h = hist.double().cpu()
    n = h.numel()
    edges = torch.linspace(lo, hi, n + 1, dtype=torch.float64)
    centres = 0.5 * (edges[:-1] + edges[1:])
    w0 = torch.cumsum(h, 0)
    w1 = torch.cumsum(h.flip(0), 0).flip(0)
    m0 = torch.cumsum(h * centres, 0) / w0.clamp(min=1e-300)
    m1 = (torch.cumsum((h * centres).flip(0), 0) / w1.flip(0).clamp(min=1e-300)).flip(0)
    var12 = w0[:-1] * w1[1:] * (m0[:-1] - m1[1:]) ** 2
    return float(centres[int(torch.argmax(var12))])


def apply_otsu_threshold(densities: torch.Tensor, fallback_percentile: float = 50.0, group=None):
    """scripts/extract_vol.py:123-150 -> (binary occupancy [N] bool, threshold).  Histogram on the GPU; with a
    process group the min/max and the histogram are all-reduced so every rank gets the global threshold."""
    d = densities.reshape(-1).float()
    if d.numel():
        lo, hi = d.min(), d.max()
    else:   # a rank whose slab is empty (more ranks than x-planes) still takes part in the collectives
        lo, hi = torch.full((), float("inf"), device=d.device), torch.full((), float("-inf"), device=d.device)
    if group is not None:
        lohi = torch.stack([-lo, hi])
        torch.distributed.all_reduce(lohi, op=torch.distributed.ReduceOp.MAX, group=group)   # one collective for min and max
        lo, hi = -lohi[0], lohi[1]
    lo, hi = float(lo), float(hi)
    if hi <= lo:
        thr = lo
    else:
        hist = ops.histogram(d, lo, hi, 256) if d.numel() else torch.zeros(256, dtype=torch.int64, device=d.device)
        if group is not None:
            torch.distributed.all_reduce(hist, group=group)
        thr = otsu_from_histogram(hist, lo, hi)
    return d > thr, thr


def save_volume(occupancy, rgb_values, output_path, resolution, bounds, threshold, metadata: Optional[dict] = None):
    """scripts/extract_vol.py:153-206: {"occupancy_volume": u8 [R,R,R], "rgb_volume": u8 [R,R,R,3], "metadata"}."""
    output_path = Path(output_path)
    if output_path.suffix.lower() != ".pth":
        raise ValueError(f"Unsupported output format: {output_path.suffix}")
    R = resolution
    occ = occupancy.reshape(-1)
    meta = {
        "resolution": int(R), "bounds": [float(bounds[0]), float(bounds[1])], "threshold": float(threshold),
        "occupied_ratio": float(occ.float().mean().item()), "total_voxels": int(occ.numel()),
        "occupied_voxels": int(occ.sum().item()), "timestamp": time.strftime("%Y-%m-%d %H:%M:%S"),
    }
    meta.update(metadata or {})
    if rgb_values.dtype == torch.uint8:
        col = rgb_values
    else:
        col = (rgb_values.float() * 255).to(torch.uint8)  # truncation, extract_vol.py:198
    output_path.parent.mkdir(parents=True, exist_ok=True)
    torch.save({"occupancy_volume": occ.to(torch.uint8).reshape(R, R, R).cpu(),
                "rgb_volume": col.reshape(R, R, R, 3).cpu(), "metadata": meta}, output_path)


def slab_range(resolution: int, rank: int, world: int):
    """x-planes [x0, x1) owned by `rank`: x is the slowest index of the ij meshgrid
    (extract_vol.py:74-77), so a slab is the contiguous flat range [x0*R*R, x1*R*R)."""
    planes = (resolution + world - 1) // world
    return min(rank * planes, resolution), min((rank + 1) * planes, resolution)


def global_otsu(local_hist: torch.Tensor, lo: float, hi: float, group=None) -> float:
    """Threshold every rank agrees on: the per-slab 256-bin histograms (all over the same global [lo,hi]) are summed
    with one all-reduce and Otsu runs on the sum."""
    hist = local_hist.clone()
    if group is not None or (torch.distributed.is_available() and torch.distributed.is_initialized()):
        torch.distributed.all_reduce(hist, group=group)
    return otsu_from_histogram(hist, lo, hi)


def default_group(world: int, group=None):
    """The process group the slab collectives run on: the caller's, else the default group when `world` > 1 ranks are
    initialised (so that a sharded call can never silently threshold every slab on its own)."""
    if group is not None or world <= 1:
        return group
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        return torch.distributed.group.WORLD
    raise RuntimeError("extract_vol: world > 1 needs an initialised torch.distributed process group (one global Otsu "
                       "threshold, extract_vol.py:123-150); pass group=False to keep per-slab statistics on purpose")


def gather_slabs(occ: torch.Tensor, col: torch.Tensor, resolution: int, rank: int, world: int, group=None, dst: int = 0):
    """Stitches the ranks' u8 slabs into ONE volume: x-slabs are contiguous ranges of the flat grid in rank
    order, so the all-gather of equal slabs IS the volume; ragged slabs (R % world != 0) are padded to the largest one and
    cut to their ranges afterwards.  Returns (occupancy [R^3] u8, rgb [R^3,3] u8) on `dst`, (None, None) elsewhere.
    One NCCL all-gather per array on the ranks' communicator: no point-to-point traffic (unbatched send / recv pairs on an
    8-rank communicator did not complete on the 8-GPU box), 4 * R^3 bytes received per rank."""
    R = resolution
    occ, col = occ.reshape(-1).contiguous(), col.reshape(-1, 3).contiguous()
    if world == 1:
        return occ, col
    dist = torch.distributed
    sizes = [(x1 - x0) * R * R for x0, x1 in (slab_range(R, r, world) for r in range(world))]
    if occ.numel() != sizes[rank]:
        raise ValueError("gather_slabs: rank %d holds %d voxels, its slab has %d" % (rank, occ.numel(), sizes[rank]))
    n_max = max(sizes)
    dev = occ.device
    if min(sizes) == n_max:
        full_o = torch.empty(R * R * R, dtype=torch.uint8, device=dev)
        full_c = torch.empty(R * R * R, 3, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(full_o, occ, group=group)
        dist.all_gather_into_tensor(full_c.view(-1), col.view(-1), group=group)
    else:
        n = sizes[rank]
        mine = torch.zeros(4 * n_max, dtype=torch.uint8, device=dev)     # [occ | rgb] of this slab, padded
        mine[:n].copy_(occ)
        mine[n_max:n_max + 3 * n].copy_(col.view(-1))
        every = torch.empty(world, 4 * n_max, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(every.view(-1), mine, group=group)
        if rank != dst:
            return None, None
        full_o = torch.cat([every[r, :sizes[r]] for r in range(world)])
        full_c = torch.cat([every[r, n_max:n_max + 3 * sizes[r]] for r in range(world)]).view(-1, 3)
    return (full_o, full_c) if rank == dst else (None, None)


def global_stats(t: torch.Tensor, group=None) -> dict:
    """min / max / mean / std (unbiased, like Tensor.std) of a tensor that is sharded over the ranks of `group`
    (the density_stats / rgb_stats metadata of extract_vol.py:323-340)."""
    x = t.reshape(-1).double()
    n = torch.tensor([float(x.numel())], dtype=torch.float64, device=t.device)
    acc = torch.stack([x.sum(), (x * x).sum()]) if x.numel() else torch.zeros(2, dtype=torch.float64, device=t.device)
    mm = torch.stack([-x.min(), x.max()]) if x.numel() else torch.full((2,), float("-inf"), dtype=torch.float64, device=t.device)
    if group is not None:
        acc = torch.cat([acc, n])
        torch.distributed.all_reduce(acc, group=group)
        torch.distributed.all_reduce(mm, op=torch.distributed.ReduceOp.MAX, group=group)
        n = acc[2:3]
    cnt, mean = float(n), float(acc[0] / n)
    var = max(float(acc[1] / n) - mean * mean, 0.0) * (cnt / max(cnt - 1.0, 1.0))
    return {"min": float(-mm[0]), "max": float(mm[1]), "mean": mean, "std": var ** 0.5}


def extract_volume_fused(model: OptimizedVanillaNeRF, resolution: int, bounds=(-1.0, 1.0), view_direction=(0.0, 0.0, -1.0),
                         rank: int = 0, world: int = 1, group=None, out=None):
    """Slab-sharded dense query: rank r owns x-planes [r*R/G,(r+1)*R/G) == a contiguous range of
    the flat grid.  Returns (sigma [n], rgb [n,3], start, thr, occupancy u8 [n], rgb u8 [n,3]) for this rank's slab.
    With world > 1 the min / max and the 256-bin histogram are all-reduced over `group` (default: the initialised
    default group), so `thr` is the ONE global Otsu threshold of extract_vol.py:123-150 on every rank."""
    group = None if group is False else default_group(world, group)
    dev = next(model.parameters()).device
    model._sync_packed()
    R = resolution
    x0, x1 = slab_range(R, rank, world)
    start, count = x0 * R * R, (x1 - x0) * R * R
    lin = torch.linspace(bounds[0], bounds[1], R, device=dev)
    view = torch.tensor(view_direction, dtype=torch.float32, device=dev)
    view = view / torch.norm(view)
    if count:
        rgb, sigma = ops.mlp_forward(model._packed, grid=(lin, view, start, count), out=out)
    else:
        rgb, sigma = torch.empty(0, 3, device=dev), torch.empty(0, device=dev)
    occ_b, thr = apply_otsu_threshold(sigma, group=group)
    if count:
        occ, col = ops.pack_volume(sigma, rgb, thr)
    else:
        occ, col = torch.empty(0, dtype=torch.uint8, device=dev), torch.empty(0, 3, dtype=torch.uint8, device=dev)
    return sigma, rgb, start, thr, occ, col


def load_vanilla_model(checkpoint_path: str, cfg, device="cuda") -> OptimizedVanillaNeRF:
    """extract_vol.py:20-54 for the vanilla network (dispatch on cfg.ngp)."""
    if not os.path.exists(checkpoint_path):
        raise FileNotFoundError(f"Checkpoint not found: {checkpoint_path}")
    ckpt = torch.load(checkpoint_path, map_location=device, weights_only=False)
    if "nerf_fine_state_dict" not in ckpt:
        raise KeyError("nerf_fine_state_dict")
    model = OptimizedVanillaNeRF(cfg.pos_L, cfg.dir_L, cfg.hidden_dim).to(device)
    model.load_state_dict(ckpt["nerf_fine_state_dict"])
    return model.eval()


def load_instant_ngp_model(checkpoint_path: str, nerf_type: str = "large", device: str = "cuda") -> InstantNGP:
    """extract_vol.py:20-54: same errors (FileNotFoundError :33-34, KeyError :51), model in eval mode on `device`."""
    if not Path(checkpoint_path).exists():
        raise FileNotFoundError(f"Checkpoint not found: {checkpoint_path}")
    ckpt = torch.load(checkpoint_path, map_location=device, weights_only=False)
    model = InstantNGP(nerf_type)
    if "nerf_fine_state_dict" not in ckpt:
        raise KeyError("Fine model state dict not found in checkpoint")
    model.load_state_dict(ckpt["nerf_fine_state_dict"])
    return model.to(device).eval()


def extract_volume_ngp(model: InstantNGP, resolution: int, bounds=(-1.0, 1.0), view_direction=(0.0, 0.0, -1.0),
                       rank: int = 0, world: int = 1, group=None, slab_points: int = 1 << 25):
    """The slab-sharded dense query of extract_volume_fused for the Instant-NGP variant: grid points are generated on the
    device in pieces of `slab_points` (never the 1.6 GB grid), each piece is ONE fused launch (extract_vol.py:85-120 with
    the canonical view as a single direction, :100).  Same return tuple as extract_volume_fused."""
    group = None if group is False else default_group(world, group)
    dev = next(model.parameters()).device
    R = resolution
    x0, x1 = slab_range(R, rank, world)
    start, count = x0 * R * R, (x1 - x0) * R * R
    lin = torch.linspace(bounds[0], bounds[1], R, device=dev)
    view = torch.tensor(view_direction, dtype=torch.float32, device=dev)
    view = view / torch.norm(view)
    sigma = torch.empty(count, 1, device=dev, dtype=torch.float32)
    rgb = torch.empty(count, 3, device=dev, dtype=torch.float32)
    with torch.no_grad():
        for s in range(0, count, slab_points):
            n = min(slab_points, count - s)
            c, d = model(ops.grid_points(lin, start + s, n), view)
            sigma[s:s + n], rgb[s:s + n] = d.float(), c.float()
    occ_b, thr = apply_otsu_threshold(sigma, group=group)
    if count:
        occ, col = ops.pack_volume(sigma, rgb, thr)
    else:
        occ, col = torch.empty(0, dtype=torch.uint8, device=dev), torch.empty(0, 3, dtype=torch.uint8, device=dev)
    return sigma, rgb, start, thr, occ, col


def get_args(argv=None):
    p = argparse.ArgumentParser(description="Dense occupancy / colour volume from a trained NeRF")
    p.add_argument("--cfg_path", type=str, required=True)
    p.add_argument("--resolution", type=int, default=512)  # ignored in favour of cfg.volume_resolution (:283-284)
    p.add_argument("--bounds", type=float, nargs=2, default=[-1.0, 1.0])
    p.add_argument("--batch-size", type=int, default=8192)
    p.add_argument("--percentile-threshold", type=float, default=50.0)
    p.add_argument("--view-direction", type=float, nargs=3, default=[0.0, 0.0, -1.0])
    return p.parse_args(argv)


def main(argv=None):
    """scripts/extract_vol.py:264-357.  Under torchrun (RANK / WORLD_SIZE / LOCAL_RANK) the grid is sharded in x-slabs
    over the ranks: NCCL all-reduces make the Otsu threshold and the metadata statistics global, the u8 slabs are gathered
    to rank 0, and rank 0 alone writes ONE volume.pth."""
    args = get_args(argv)
    if args.resolution <= 0:
        raise ValueError("Resolution must be positive")
    if args.bounds[0] >= args.bounds[1]:
        raise ValueError("Invalid bounds: min must be less than max")
    cfg = load_config(args.cfg_path)
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)   # the ops launch on the current device's stream
    own_pg = False
    if world > 1 and not torch.distributed.is_initialized():
        torch.distributed.init_process_group("nccl", device_id=dev, timeout=_PG_TIMEOUT)
        own_pg = True
    group = torch.distributed.group.WORLD if world > 1 else None
    R = int(cfg.volume_resolution)   # --resolution is ignored in favour of the cfg (:283-284, :346-347)
    with torch.no_grad():
        if getattr(cfg, "ngp", False):
            model = load_instant_ngp_model(str(cfg.checkpoint), getattr(cfg, "nerf_type", "large"), dev)
            sigma, rgb, _, thr, occ, col = extract_volume_ngp(model, R, tuple(args.bounds), tuple(args.view_direction),
                                                              rank=rank, world=world, group=group)
        else:
            model = load_vanilla_model(str(cfg.checkpoint), cfg, dev)
            sigma, rgb, _, thr, occ, col = extract_volume_fused(model, R, tuple(args.bounds), tuple(args.view_direction),
                                                                rank=rank, world=world, group=group)
        meta = {"checkpoint_path": str(cfg.checkpoint), "device_used": str(dev), "batch_size": int(args.batch_size),
                "density_stats": global_stats(sigma, group), "view_direction": list(args.view_direction),
                "rgb_stats": global_stats(rgb, group)}
        occ, col = gather_slabs(occ, col, R, rank, world, group)
    if rank == 0:
        save_volume(occ, col, str(cfg.volume_output_path), R, tuple(args.bounds), thr, meta)
        print(f"Saved colored occupancy volume with metadata: {cfg.volume_output_path}")
    if own_pg:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
