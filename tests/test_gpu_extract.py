"""GPU parity tests of what bench.py times on the extract_vol side (BASELINE config 4) and of the two on-disk layouts:
extract_volume_fused / volume_extraction against goldens the reference's own dense query produced
(tests/golden/make_golden_extract.py: extract_occupancy_volume.query_density_and_rgb_batch, fp32), save_volume against
the dict the reference's save_volume wrote, extract_vol.main and train(cfg) end to end on temporary files.

Bars: sigma / rgb <= 2e-2 (north_star's bf16 MLP tolerance; sigma relative to max(1, |sigma|)); occupancy and u8 colours
bit-exact given the same sigma / rgb / threshold; slabs concatenate to the single-rank volume bit for bit."""
import json
import os

import numpy as np
import pytest
import torch

from common import make_rays, make_state_dict, peaked

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def D():
    import deepfields_nerf_b200 as d
    assert torch.cuda.is_available()
    return d


@pytest.fixture(scope="module")
def G():
    return dict(np.load(os.path.join(HERE, "golden", "extract_golden.npz")))


@pytest.fixture(scope="module")
def SPEC():
    with open(os.path.join(HERE, "golden", "extract_golden.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="module")
def model(D):
    m = D.OptimizedVanillaNeRF().cuda()
    m.load_state_dict(peaked(make_state_dict(32, 256), 3.0))
    return m.eval()


def close_sigma(got, want):
    got, want = got.float().cpu().reshape(-1), torch.from_numpy(want).reshape(-1)
    return float(((got - want).abs() / want.abs().clamp(min=1.0)).max())


def test_extract_volume_fused_r64_vs_reference(D, G, model):
    R = 64
    with torch.no_grad():
        sigma, rgb, start, thr, occ, col = D.extract_volume_fused(model, R)
    assert start == 0 and sigma.shape == (R ** 3,) and rgb.shape == (R ** 3, 3)
    assert occ.dtype == torch.uint8 and col.dtype == torch.uint8 and col.shape == (R ** 3, 3)
    assert close_sigma(sigma, G["r64_sigma"]) <= 2e-2
    assert float((rgb.cpu() - torch.from_numpy(G["r64_rgb_f16"]).float()).abs().max()) <= 2e-2 + 5e-4   # + fp16 storage of the golden
    # packing is exact arithmetic on the kernel's own sigma / rgb: strict >, truncating * 255 (extract_vol.py:142,198)
    assert torch.equal(occ, (sigma > thr).to(torch.uint8))
    assert torch.equal(col, (rgb * 255).to(torch.uint8))
    # ... and lands within the MLP tolerance of the u8 volume the reference packs from its fp32 query
    du8 = (col.cpu().int() - torch.from_numpy(G["r64_rgb_u8"]).int()).abs().max()
    assert int(du8) <= int(2e-2 * 255) + 1
    # the threshold: Otsu of the kernel's sigma == the oracle's Otsu restatement on the same values (within one bin)
    from oracle import nerf_oracle as O
    lo, hi = float(sigma.min()), float(sigma.max())
    assert abs(thr - O.otsu_threshold(sigma.cpu())) <= (hi - lo) / 256 + 1e-6
    # occupancy against the reference's sigma at the same threshold: only voxels within the MLP tolerance of it may differ
    ref_sigma = torch.from_numpy(G["r64_sigma"])
    differ = occ.cpu().bool() != (ref_sigma > thr)
    assert bool(((ref_sigma - thr).abs()[differ] <= 2e-2 * max(1.0, abs(thr))).all())


@pytest.mark.parametrize("world", [2, 3, 8])
def test_extract_volume_fused_slabs_concatenate(D, model, world):
    R = 64
    with torch.no_grad():
        full = D.extract_volume_fused(model, R)
        parts = [D.extract_volume_fused(model, R, rank=r, world=world, group=False) for r in range(world)]   # per-slab statistics on purpose
    assert [p[2] for p in parts] == [D.slab_range(R, r, world)[0] * R * R for r in range(world)]
    assert torch.equal(torch.cat([p[0] for p in parts]), full[0]) and torch.equal(torch.cat([p[1] for p in parts]), full[1])
    # with the ONE global threshold the packed slabs are the packed volume
    occ = torch.cat([D.ops.pack_volume(p[0], p[1], full[3])[0] for p in parts])
    col = torch.cat([D.ops.pack_volume(p[0], p[1], full[3])[1] for p in parts])
    assert torch.equal(occ, full[4]) and torch.equal(col, full[5])


def test_sharded_extract_without_group_raises(D, model):
    with pytest.raises(RuntimeError, match="process group"):
        D.extract_volume_fused(model, 16, rank=0, world=2)


def test_extract_volume_fused_r128_slab_vs_reference(D, G, model):
    R, world, rank = 128, 8, 3
    with torch.no_grad():
        sigma, rgb, start, thr, occ, col = D.extract_volume_fused(model, R, rank=rank, world=world, group=False)
    a, b = (int(v) for v in G["r128_slab_range"])
    assert start == a and sigma.numel() == b - a
    assert close_sigma(sigma, G["r128_slab_sigma"]) <= 2e-2
    assert float((rgb.cpu() - torch.from_numpy(G["r128_slab_rgb_f16"]).float()).abs().max()) <= 2e-2 + 5e-4


def test_volume_extraction_points_api_vs_reference(D, G, model):
    """extract_vol.volume_extraction(model, points, canonical_view) -- the reference signature -- on the 64^3 grid."""
    R = 64
    pts = D.generate_3d_grid((-1.0, 1.0), R)
    dens, cols = D.volume_extraction(model, pts, torch.tensor([0.0, 0.0, -1.0]), batch_size=8192, device="cuda", show_progress=False)
    assert dens.shape == (R ** 3, 1) and cols.shape == (R ** 3, 3)
    assert close_sigma(dens, G["r64_sigma"]) <= 2e-2
    assert float((cols.cpu() - torch.from_numpy(G["r64_rgb_f16"]).float()).abs().max()) <= 2e-2 + 5e-4
    with torch.no_grad():
        sigma, rgb, *_ = D.extract_volume_fused(model, R)
    assert torch.equal(dens.reshape(-1), sigma) and torch.equal(cols, rgb)     # in-kernel grid == explicit points


def test_save_volume_matches_reference_file(D, G, SPEC, tmp_path):
    sv = SPEC["save_volume"]
    occ = torch.from_numpy(G["sv_occ_in"]).cuda()
    rgb = torch.from_numpy(G["sv_rgb_in"]).cuda()
    path = tmp_path / "volume.pth"
    D.save_volume(occ, rgb, str(path), sv["resolution"], tuple(sv["bounds"]), sv["threshold"], dict(sv["metadata_in"]))
    vol = torch.load(str(path), map_location="cpu", weights_only=False)
    assert list(vol.keys()) == sv["top_keys"]
    for k in ("occupancy_volume", "rgb_volume"):
        assert str(vol[k].dtype) == sv[k]["dtype"] and list(vol[k].shape) == sv[k]["shape"], k
        assert vol[k].device.type == "cpu"
        assert np.array_equal(vol[k].numpy(), G["sv_" + k]), k
    assert list(vol["metadata"].keys()) == sv["metadata_keys"]
    got = {k: v for k, v in vol["metadata"].items() if k != "timestamp"}
    want = dict(sv["metadata"])
    assert got.pop("occupied_ratio") == pytest.approx(want.pop("occupied_ratio"), rel=1e-6)
    assert got == want
    assert isinstance(vol["metadata"]["timestamp"], str) and len(vol["metadata"]["timestamp"]) == 19
    with pytest.raises(ValueError, match="Unsupported output format"):
        D.save_volume(occ, rgb, str(tmp_path / "volume.npy"), sv["resolution"], tuple(sv["bounds"]), sv["threshold"], {})
    assert sv["bad_suffix_raises"].startswith("ValueError: Unsupported output format")


def _write_cfg(tmp_path, n_rays, batch, epochs, R):
    o, d, near, far, rgb = make_rays(3, n_rays)
    rays = tmp_path / "scene_ray_data.npz"
    np.savez(rays, rays_o=o.numpy().astype(np.float64), rays_d=d.numpy().astype(np.float64), rgbs=rgb.numpy(),
             t_near=near.numpy().reshape(-1), t_far=far.numpy().reshape(-1))
    cfg = tmp_path / "scene.yml"
    cfg.write_text(f"""
scene_name: "scene"
output_dir: "{tmp_path}/outputs/${{scene_name}}"
checkpoint_dir: "${{output_dir}}/checkpoints"
rays_file: "{rays}"
volume_resolution: {R}
checkpoint: "${{checkpoint_dir}}/nerf_final.pth"
volume_output_path: "${{output_dir}}/volume.pth"
batch_size: {batch}
num_epochs: {epochs}
lr: 0.0005
ngp: false
nerf_type: large
hidden_dim: 256
pos_L: 10
dir_L: 4
N_samples: 64
N_importance: 128
white_bg: false
chunk_size: 32768
use_memmap: False
""")
    return cfg


def test_train_cfg_then_extract_main(D, SPEC, tmp_path):
    """scripts/train.py --cfg_path ... followed by scripts/extract_vol.py --cfg_path ... on temporary files: the two on-disk
    layouts (nerf_final.pth: train.py:343-364; volume.pth: extract_vol.py:181-202,323-340) and the TensorBoard tags."""
    n_rays, batch, epochs, R = 700, 256, 2, 24
    cfg_path = _write_cfg(tmp_path, n_rays, batch, epochs, R)
    cfg = D.load_config(str(cfg_path))
    tr = D.train(cfg)
    steps = epochs * ((n_rays + batch - 1) // batch)                      # len(dataloader) = ceil(n / batch): drop_last=False
    assert int(tr.step_dev.item()) == steps == tr.total_steps
    ckpt_dir = tmp_path / "outputs" / "scene" / "checkpoints"
    assert (ckpt_dir / "nerf_final.pth").exists() and (ckpt_dir / "nerf_epoch_000002.pth").exists()
    ck = torch.load(str(ckpt_dir / "nerf_final.pth"), map_location="cpu", weights_only=False)
    assert list(ck.keys()) == ["iteration", "nerf_coarse_state_dict", "nerf_fine_state_dict", "optimizer_state_dict"]
    assert ck["iteration"] == epochs                                         # "iteration" holds the EPOCH (train.py:345)
    # state dicts load into the reference's module: same keys in the same order, shapes, fp32 (spec generated from model.py)
    for name in ("nerf_coarse_state_dict", "nerf_fine_state_dict"):
        got = [[k, list(v.shape), str(v.dtype)] for k, v in ck[name].items()]
        assert got == SPEC["state_dict"], name
        assert all(torch.isfinite(v).all() for v in ck[name].values())
    init = D.OptimizedVanillaNeRF()
    assert not torch.equal(ck["nerf_fine_state_dict"]["feature_layer.weight"], init.state_dict()["feature_layer.weight"])
    # the optimiser state resumes in a stock AdamW + OneCycleLR built like train.py:244-259
    params = [torch.nn.Parameter(v.clone()) for nm in ("nerf_coarse_state_dict", "nerf_fine_state_dict")
              for k, v in ck[nm].items() if "freq_bands" not in k]
    opt = torch.optim.AdamW(params, lr=5e-4, weight_decay=1e-6)
    opt.load_state_dict(ck["optimizer_state_dict"])
    assert len(opt.state_dict()["state"]) == 48 and float(opt.state_dict()["state"][0]["step"]) == steps
    torch.optim.lr_scheduler.OneCycleLR(opt, max_lr=5e-3, total_steps=steps + 5, pct_start=0.1, anneal_strategy="cos",
                                        div_factor=10, final_div_factor=100, last_epoch=steps - 1)
    # TensorBoard: the reference's five tags at step = epoch * steps_per_epoch + idx (train.py:333-338)
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
    acc = EventAccumulator(str(tmp_path / "outputs" / "scene" / "logs"))
    acc.Reload()
    assert set(acc.Tags()["scalars"]) == {"Loss/Coarse", "Loss/Fine", "Loss/Total", "PSNR", "Scheduler Step"}
    spe = steps // epochs
    assert [e.step for e in acc.Scalars("Loss/Total")] == [1 * spe, 2 * spe]
    tot, lc, lf = (acc.Scalars(t)[0].value for t in ("Loss/Total", "Loss/Coarse", "Loss/Fine"))
    assert tot == pytest.approx(lc + lf, rel=1e-5) and acc.Scalars("PSNR")[0].value == pytest.approx(-10 * np.log10(lf), rel=1e-4)

    # ---- extract_vol.main on the checkpoint just written ----
    from deepfields_nerf_b200 import extract_vol as X
    X.main(["--cfg_path", str(cfg_path), "--resolution", "999", "--batch-size", "4096"])     # --resolution is ignored (:283-284)
    vol = torch.load(str(tmp_path / "outputs" / "scene" / "volume.pth"), map_location="cpu", weights_only=False)
    sv = SPEC["save_volume"]
    assert list(vol.keys()) == sv["top_keys"]
    assert vol["occupancy_volume"].dtype == torch.uint8 and tuple(vol["occupancy_volume"].shape) == (R, R, R)
    assert vol["rgb_volume"].dtype == torch.uint8 and tuple(vol["rgb_volume"].shape) == (R, R, R, 3)
    md = vol["metadata"]
    assert list(md.keys()) == sv["metadata_keys"]
    assert md["resolution"] == R and md["total_voxels"] == R ** 3 and md["bounds"] == [-1.0, 1.0] and md["batch_size"] == 4096
    assert md["occupied_voxels"] == int(vol["occupancy_volume"].sum()) and md["occupied_ratio"] == pytest.approx(md["occupied_voxels"] / R ** 3)
    assert set(md["density_stats"]) == {"min", "max", "mean", "std"} == set(md["rgb_stats"])
    # the volume is the fused query of the loaded fine network
    m = X.load_vanilla_model(str(ckpt_dir / "nerf_final.pth"), cfg, "cuda")
    with torch.no_grad():
        sigma, rgb, _, thr, occ, col = D.extract_volume_fused(m, R)
    assert md["threshold"] == pytest.approx(thr) and torch.equal(vol["occupancy_volume"].reshape(-1), occ.cpu())
    assert torch.equal(vol["rgb_volume"].reshape(-1, 3), col.cpu())
    assert md["density_stats"]["max"] == pytest.approx(float(sigma.max())) and md["density_stats"]["std"] == pytest.approx(float(sigma.std()), rel=1e-4)
    with pytest.raises(FileNotFoundError):
        X.load_vanilla_model(str(tmp_path / "missing.pth"), cfg, "cuda")


def test_ray_store_batches(D, tmp_path):
    """The GPU-resident replacement of Dataset + DataLoader (dataset.py:32-105, train.py:209-217): every ray once per
    epoch, last short batch kept (drop_last=False), ranks take disjoint shards of the SAME permutation."""
    n, batch = 1000, 256
    cfg_path = _write_cfg(tmp_path, n, batch, 1, 8)
    cfg = D.load_config(str(cfg_path))
    o, d, near, far, rgb = make_rays(3, n)
    st = D.RayStore(str(cfg.rays_file), torch.device("cuda"))
    assert len(st) == n and st.steps_per_epoch(batch) == 4
    g = torch.Generator(device="cuda").manual_seed(0)
    got = list(st.batches(batch, g))
    assert [b[0].shape[0] for b in got] == [256, 256, 256, 232]
    all_rgb = torch.cat([b[4] for b in got]).cpu()
    assert torch.equal(all_rgb.sort(dim=0).values, rgb.sort(dim=0).values)            # a permutation of the data set
    assert got[0][2].shape == (256,) and got[0][0].dtype == torch.float32
    # rows stay together: the gathered near / far belong to the gathered origin
    key = {tuple(np.round(r, 6)): i for i, r in enumerate(o.numpy())}
    idx = [key[tuple(np.round(r, 6))] for r in got[1][0].cpu().numpy()]
    assert torch.equal(got[1][2].cpu(), near.reshape(-1)[idx]) and torch.equal(got[1][4].cpu(), rgb[idx])
    # two ranks: disjoint halves of each global batch of the same shuffle
    parts = []
    for r in range(2):
        sr = D.RayStore(str(cfg.rays_file), torch.device("cuda"), rank=r, world=2)
        parts.append(list(sr.batches(batch, torch.Generator(device="cuda").manual_seed(0))))
    assert [b[0].shape[0] for b in parts[0]] == [128, 128, 128, 116]
    for b0, b1, bg in zip(parts[0], parts[1], got):
        assert torch.equal(torch.cat([b0[4], b1[4]]), bg[4])
