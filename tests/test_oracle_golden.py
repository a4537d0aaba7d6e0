"""Pins oracle/nerf_oracle.py against the golden vectors produced by the reference itself
(tests/golden/make_golden.py).  CPU only."""
import numpy as np
import torch

from common import make_draws, make_noise, make_rays, make_state_dict, peaked
from oracle import nerf_oracle as O

B, S, I = 48, 64, 128
T = lambda a: torch.from_numpy(np.asarray(a))


def test_stratified_and_points_bit_exact(golden):
    o, d, near, far, _ = make_rays(11, B)
    t_rand, _ = make_draws(12, B, S, I)
    z = O.stratified_z(near, far, t_rand)
    assert np.array_equal(z.numpy(), golden["sampler_z"])
    pts, dirs = O.points_on_rays(o, d, z)
    assert np.array_equal(pts.numpy(), golden["sampler_pts"])
    assert np.array_equal(dirs.numpy(), golden["sampler_dirs"])


def test_sample_pdf_indices_and_samples_bit_exact(golden):
    z = T(golden["sampler_z"])
    mids = 0.5 * (z[:, :-1] + z[:, 1:])
    samples, inds, _ = O.sample_pdf(mids, T(golden["pdf_weights"]), T(golden["pdf_u"]))
    assert np.array_equal(inds.numpy(), golden["pdf_inds"])
    assert np.array_equal(samples.numpy(), golden["pdf_samples"])
    assert inds.min() >= 1 and inds.max() <= 63


def test_merge_bit_exact(golden):
    zz = O.merge_depths(T(golden["sampler_z"]), T(golden["pdf_samples"]))
    assert np.array_equal(zz.numpy(), golden["fine_z"])


def test_posenc_bit_exact(golden):
    x = T(golden["pe_in"])
    assert np.array_equal(O.posenc(x, 10).numpy(), golden["pe_L10"])
    assert np.array_equal(O.posenc(x, 4).numpy(), golden["pe_L4"])


def test_mlp_forward(golden):
    sd = make_state_dict(21, 256)
    pos, view = T(golden["mlp_pos"]), T(golden["mlp_view"])
    rgb, sig = O.mlp_forward(sd, pos, view)
    np.testing.assert_allclose(rgb.numpy(), golden["mlp_fp32_rgb"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(sig.numpy(), golden["mlp_fp32_sigma"], rtol=1e-5, atol=1e-6)
    rgb, sig = O.mlp_forward(sd, pos, view, noise=make_noise(16, 512))
    np.testing.assert_allclose(sig.numpy(), golden["mlp_fp32_train_sigma"], rtol=1e-5, atol=1e-6)
    rgb, sig = O.mlp_forward(peaked(sd), pos, view)
    np.testing.assert_allclose(sig.numpy(), golden["mlp_peaked_sigma"], rtol=1e-4, atol=1e-5)
    # autocast(bf16) restatement: same rounding points -> agreement far inside the 2e-2 budget
    rgb, sig = O.mlp_forward(sd, pos, view, low=torch.bfloat16)
    np.testing.assert_allclose(rgb.float().numpy(), golden["mlp_bf16_rgb"], atol=8e-3)
    np.testing.assert_allclose(sig.float().numpy(), golden["mlp_bf16_sigma"], atol=8e-3)


def test_composite(golden):
    for tag, zkey in (("c", "sampler_z"), ("f", "fine_z")):
        rgb, sig, z = T(golden[f"comp_{tag}_rgb"]), T(golden[f"comp_{tag}_sigma"]), T(golden[zkey])
        for wb in (False, True):
            out = O.composite(rgb, sig, z, wb)
            assert np.array_equal(out["rgb_map"].numpy(), golden[f"comp_{tag}_map_wb{int(wb)}"])
            assert np.array_equal(out["weights"].numpy(), golden[f"comp_{tag}_w_wb{int(wb)}"])
        out = O.composite(rgb, sig, z, False, dist_last=1e6)
        assert np.array_equal(out["weights"].numpy(), golden[f"comp_{tag}_w_1e6"])
        r2, s2 = rgb.clone().requires_grad_(), sig.clone().requires_grad_()
        (O.composite(r2, s2, z)["rgb_map"] * T(golden[f"comp_{tag}_G"])).sum().backward()
        np.testing.assert_allclose(r2.grad.numpy(), golden[f"comp_{tag}_drgb"], rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(s2.grad.numpy(), golden[f"comp_{tag}_dsigma"], rtol=1e-5, atol=1e-9)


def test_render_losses_and_grads_fp32(golden):
    o, d, near, far, tgt = make_rays(11, B)
    t_rand, u = make_draws(12, B, S, I)
    for train_mode in (False, True):
        sdc = {k: v.clone().requires_grad_(v.dim() > 0 and "freq" not in k) for k, v in make_state_dict(31, 64).items()}
        sdf = {k: v.clone().requires_grad_(v.dim() > 0 and "freq" not in k) for k, v in make_state_dict(32, 64).items()}
        nc = make_noise(41, B * S) if train_mode else None
        nf = make_noise(42, B * (S + I)) if train_mode else None
        out = O.render_losses(sdc, sdf, o, d, near, far, tgt, t_rand, u, nc, nf, dist_last=1e6)
        out["loss"].backward()
        tag = f"fp_fp32_{'train' if train_mode else 'eval'}"
        g = golden[tag + "_losses"]
        np.testing.assert_allclose([out["loss"].item(), out["loss_coarse"].item(), out["loss_fine"].item()], g, rtol=2e-5)
        assert np.array_equal(out["inds"].numpy(), golden[tag + "_inds"])
        for nm, sd in (("c", sdc), ("f", sdf)):
            for k, v in sd.items():
                if v.requires_grad:
                    ref = golden[f"{tag}_grad_{nm}_{k}"]
                    np.testing.assert_allclose(v.grad.numpy(), ref, rtol=2e-3, atol=2e-6 + 1e-4 * np.abs(ref).max())


def test_render_losses_bf16(golden):
    o, d, near, far, tgt = make_rays(11, B)
    t_rand, u = make_draws(12, B, S, I)
    sdc, sdf = make_state_dict(31, 64), make_state_dict(32, 64)
    out = O.render_losses(sdc, sdf, o, d, near, far, tgt, t_rand, u, low=torch.bfloat16)
    g = golden["fp_bf16_eval_losses"]
    np.testing.assert_allclose([out["loss"].item(), out["loss_coarse"].item(), out["loss_fine"].item()], g, rtol=2e-2)


def test_grid_and_dense_query(golden):
    g = O.grid_points((-1.0, 1.0), 12)
    assert np.array_equal(g.numpy(), golden["grid_R12"])
    sd = make_state_dict(21, 256)
    sig, rgb = O.dense_query(sd, g, torch.tensor([0.0, 0.0, -1.0]))
    np.testing.assert_allclose(sig.numpy(), golden["grid_sigma"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(rgb.numpy(), golden["grid_rgb"], rtol=1e-5, atol=1e-6)


def test_adamw_onecycle(golden):
    p = T(golden["adamw_p0"]).clone()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    total = 50
    for s in range(total - 1):
        lr, b1 = O.onecycle(s, total, 5e-4)
        assert abs(lr - golden["adamw_lrs"][s]) <= 1e-12 + 1e-9 * lr, (s, lr, golden["adamw_lrs"][s])
        p, m, v = O.adamw_step(p, T(golden["adamw_grads"][s]), m, v, s + 1, lr, b1=b1)
        np.testing.assert_allclose(p.numpy(), golden["adamw_params"][s], rtol=2e-6, atol=1e-7)


# ---- ray generation / cube intersection (scripts/precompute_rays.py) ---------------------------------
def _rays_golden():
    import os
    return dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "rays_golden.npz")))


def test_camera_rays_match_reference():
    from common import make_cameras
    g = _rays_golden()
    H, W, K, poses, _ = make_cameras(7, M=3, H=12, W=16)
    o, d = O.camera_rays(H, W, K, poses)
    assert np.array_equal(o, g["rays_o"])
    # fp64 sums in a different association than numpy's einsum: a few ulp of fp64, nothing after the fp32 cast
    assert np.allclose(d, g["rays_d"], rtol=0, atol=4e-16)
    assert np.array_equal(d.astype(np.float32), g["rays_d"].astype(np.float32))


def test_ray_cube_hits_bit_exact():
    from common import edge_case_rays
    g = _rays_golden()
    hits, t_near, t_far = O.ray_cube_hits(g["rays_o"].reshape(-1, 3), g["rays_d"].reshape(-1, 3))
    assert np.array_equal(hits, g["hits"])
    assert np.array_equal(t_near, g["t_near"], equal_nan=True) and np.array_equal(t_far, g["t_far"], equal_nan=True)
    eo, ed = edge_case_rays()
    hits, t_near, t_far = O.ray_cube_hits(eo, ed, -0.5, 0.75)
    assert np.array_equal(hits, g["edge_hits"])
    assert np.array_equal(t_near, g["edge_near"], equal_nan=True) and np.array_equal(t_far, g["edge_far"], equal_nan=True)
    assert np.isnan(g["edge_near"]).any() and g["edge_hits"].any() and not g["edge_hits"].all()   # the fixture covers the cases


def test_precompute_rays_keeps_hits_in_order():
    from common import make_cameras
    g = _rays_golden()
    H, W, K, poses, images = make_cameras(7, M=3, H=12, W=16)
    out = O.precompute_rays(H, W, K, poses, images)
    assert np.array_equal(out["rays_o"], g["kept_o"]) and np.array_equal(out["rgbs"], g["kept_rgb"])
    assert np.array_equal(out["rays_d"].astype(np.float32), g["kept_d"].astype(np.float32))
    assert np.array_equal(out["t_near"], g["kept_near"]) and np.array_equal(out["t_far"], g["kept_far"])


# ---- occupied voxels as points (scripts/post_process_vol.py) -----------------------------------------
VOL_AABB = {"aabb_min": [-0.55, -0.4, -0.7], "aabb_max": [0.6, 1.0, 0.22580645161290325]}


def test_volume_points_bit_exact():
    import os
    from common import make_volume
    g = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "volume_points_golden.npz")))
    occ, rgb = make_volume(5, 32)
    pts, cols = O.volume_points(occ, rgb)
    assert pts.dtype == np.float64 and cols.dtype == np.float32
    assert np.array_equal(pts, g["points"]) and np.array_equal(cols, g["rgbs"])
    pts, cols = O.volume_points(occ, rgb, VOL_AABB["aabb_min"], VOL_AABB["aabb_max"])
    assert np.array_equal(pts, g["sliced_points"]) and np.array_equal(cols, g["sliced_rgbs"])
    assert 0 < len(g["sliced_points"]) < len(g["points"])
    assert (g["sliced_points"][:, 2] == VOL_AABB["aabb_max"][2]).any()   # the closed upper bound is hit exactly


def test_dense_query_64_and_packing_vs_reference():
    """The oracle's dense query (extract_vol.py:85-120) and u8 packing (:142,197-198) against the 64^3 goldens the reference's
    query_density_and_rgb_batch produced (tests/golden/make_golden_extract.py) -- a strided sample keeps the CPU suite fast."""
    import os
    from common import peaked
    G = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "extract_golden.npz")))
    sd = peaked(make_state_dict(32, 256), 3.0)
    pts = O.grid_points((-1.0, 1.0), 64)
    idx = torch.arange(0, 64 ** 3, 37)
    with torch.no_grad():
        dens, cols = O.dense_query(sd, pts[idx], torch.tensor([0.0, 0.0, -1.0]))
    assert float((dens.reshape(-1) - torch.from_numpy(G["r64_sigma"])[idx]).abs().max()) <= 1e-5
    assert float((cols - torch.from_numpy(G["r64_rgb_f16"]).float()[idx]).abs().max()) <= 5e-4      # the golden rgb is stored in fp16
    occ, col = O.pack_volume(torch.from_numpy(G["r64_sigma"]), torch.from_numpy(G["r64_rgb_f16"]).float(), 0.25, 64)
    assert occ.dtype == torch.uint8 and int(occ.sum()) == int((G["r64_sigma"] > 0.25).sum())
    # (rgb * 255).byte() truncates: recomputed from the fp16-stored rgb it is within one level of the reference's u8
    assert int((col.reshape(-1, 3).int() - torch.from_numpy(G["r64_rgb_u8"]).int()).abs().max()) <= 1


def test_save_volume_spec_is_the_reference_layout():
    import json, os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "extract_golden.json")) as fh:
        spec = json.load(fh)
    sv = spec["save_volume"]
    assert sv["top_keys"] == ["occupancy_volume", "rgb_volume", "metadata"]
    assert sv["metadata_keys"][:7] == ["resolution", "bounds", "threshold", "occupied_ratio", "total_voxels", "occupied_voxels", "timestamp"]
    # the state-dict spec of model.py:66-137 is what common.make_state_dict builds (key order and shapes)
    sd = make_state_dict(1, 256)
    assert [[k, list(v.shape), str(v.dtype)] for k, v in sd.items()] == spec["state_dict"]
