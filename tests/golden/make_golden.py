"""Generates tests/golden/*.npz by RUNNING THE REFERENCE ITSELF (imported read-only from
/root/reference/scripts) on the deterministic inputs of tests/golden/common.py.

Run once in the build container:  PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py
The reference cannot travel to the GPU box, so the outputs are committed as small fixtures.

How the reference is imported:
  * model.py, sampling.py, extract_occupancy_volume.py import unmodified;
  * train.py / train_no_amp.py need two absent third-party modules at import time only
    (omegaconf, tinycudann) -> empty stub modules in sys.modules; forward_pass reads the module
    global ``cfg`` -> set to a SimpleNamespace;
  * the reference draws its randoms internally -> torch.rand / torch.randn_like are wrapped to pop
    pre-generated tensors (draw order: sampler rand, coarse-chunk randn, sample_pdf rand,
    fine-chunk randn).
"""
from __future__ import annotations

import os
import sys
import types
from contextlib import contextmanager
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.dont_write_bytecode = True
REF = "/root/reference/scripts"

from common import make_draws, make_noise, make_rays, make_state_dict, peaked  # noqa: E402


def import_reference():
    om = types.ModuleType("omegaconf")
    om.OmegaConf = type("OmegaConf", (), {})
    sys.modules.setdefault("omegaconf", om)
    sys.modules.setdefault("tinycudann", types.ModuleType("tinycudann"))
    sys.path.insert(0, REF)
    import model as ref_model
    import sampling as ref_sampling
    import train as ref_train
    import train_no_amp as ref_train_fp32
    import extract_occupancy_volume as ref_eov

    return ref_model, ref_sampling, ref_train, ref_train_fp32, ref_eov


@contextmanager
def injected_draws(rands, randns):
    """Replace torch.rand / torch.randn_like by queues of pre-drawn tensors."""
    rands, randns = list(rands), list(randns)
    real_rand, real_randn_like = torch.rand, torch.randn_like
    log = {"searchsorted": []}
    real_ss = torch.searchsorted

    def fake_rand(*a, **k):
        t = rands.pop(0)
        shape = a[0] if len(a) == 1 and not isinstance(a[0], int) else a
        assert tuple(shape) == tuple(t.shape), (shape, t.shape)
        return t.clone()

    def fake_randn_like(x, *a, **k):
        t = randns.pop(0)
        assert t.shape == x.shape, (t.shape, x.shape)
        return t.to(x.dtype)

    def spy_ss(*a, **k):
        r = real_ss(*a, **k)
        log["searchsorted"].append(r.clone())
        return r

    torch.rand, torch.randn_like, torch.searchsorted = fake_rand, fake_randn_like, spy_ss
    try:
        yield log
    finally:
        torch.rand, torch.randn_like, torch.searchsorted = real_rand, real_randn_like, real_ss
        assert not rands and not randns, "unused injected draws"


def np_(t):
    return t.detach().float().cpu().numpy().copy()


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    ref_model, ref_sampling, ref_train, ref_fp32, ref_eov = import_reference()
    out = {}
    B, S, I = 48, 64, 128
    CH = 32768
    cfg = SimpleNamespace(N_samples=S, N_importance=I, chunk_size=CH)
    ref_train.cfg = cfg
    ref_fp32.cfg = cfg
    dev = torch.device("cpu")

    rays_o, rays_d, near, far, target = make_rays(11, B)
    t_rand, u = make_draws(12, B, S, I)

    # ---- sampler (sampling.py:10-32) ---------------------------------------------------------
    sampler = ref_sampling.StratifiedSampler(S, dev)
    with injected_draws([t_rand], []):
        pts, dirs, z = sampler(rays_o, rays_d, near, far)
    out["sampler_z"], out["sampler_pts"], out["sampler_dirs"] = np_(z), np_(pts), np_(dirs)

    # ---- sample_pdf (train.py:31-75) on synthetic peaked + flat weights ----------------------
    rs = np.random.RandomState(13)
    w = rs.uniform(0, 1, (B, S - 2)).astype(np.float32) ** 8
    w[: B // 4] = 0.0                                    # all-zero rows -> uniform pdf
    w[B // 4 : B // 2, 5:] = 0.0                         # mass in the first few bins only
    w = torch.from_numpy(w)
    mids = 0.5 * (z[:, :-1] + z[:, 1:])
    u_edge = u.clone()
    u_edge[:, 0] = 0.0                                   # u == cdf[0]
    u_edge[:, 1] = 1.0 - 2.0 ** -24                      # largest fp32 below 1
    with injected_draws([u_edge], []) as log:
        zs = ref_train.sample_pdf(mids, w, I, dev)
    out["pdf_weights"], out["pdf_u"] = np_(w), np_(u_edge)
    out["pdf_samples"] = np_(zs)
    out["pdf_inds"] = log["searchsorted"][0].numpy().astype(np.int64)

    # ---- fine_points (sampling.py:34-49) -----------------------------------------------------
    pts_f, dirs_f, z_all = sampler.fine_points(rays_o, rays_d, zs, z, near, far, S, I)
    out["fine_z"], out["fine_pts"] = np_(z_all), np_(pts_f)

    # ---- positional encoding (model.py:6-53) -------------------------------------------------
    x = torch.from_numpy(np.random.RandomState(14).uniform(-1, 1, (256, 3)).astype(np.float32))
    out["pe_in"] = np_(x)
    out["pe_L10"] = np_(ref_model.OptimizedPositionalEncoding(10)(x))
    out["pe_L4"] = np_(ref_model.OptimizedPositionalEncoding(4)(x))

    # ---- MLP forward, hidden 256 (model.py:139-187) -------------------------------------------
    sd = make_state_dict(21, 256)
    net = ref_model.OptimizedVanillaNeRF(10, 4, 256)
    net.load_state_dict(sd)
    P = 512
    rsx = np.random.RandomState(15)
    pos = torch.from_numpy(rsx.uniform(-1, 1, (P, 3)).astype(np.float32))
    vd = rsx.normal(size=(P, 3))
    vd = torch.from_numpy((vd / np.linalg.norm(vd, axis=1, keepdims=True)).astype(np.float32))
    out["mlp_pos"], out["mlp_view"] = np_(pos), np_(vd)
    net.eval()
    with torch.no_grad():
        rgb, sig = net(pos, vd)
        out["mlp_fp32_rgb"], out["mlp_fp32_sigma"] = np_(rgb), np_(sig)
        sgm, feat = net.query_density(pos)
        out["mlp_fp32_feat"] = np_(feat)
        with torch.autocast("cpu", dtype=torch.bfloat16):
            rgb, sig = net(pos, vd)
        out["mlp_bf16_rgb"], out["mlp_bf16_sigma"] = np_(rgb), np_(sig)
    noise = make_noise(16, P)
    net.train()
    with torch.no_grad(), injected_draws([], [noise]):
        rgb, sig = net(pos, vd)
    out["mlp_fp32_train_rgb"], out["mlp_fp32_train_sigma"] = np_(rgb), np_(sig)
    # peaked-sigma weights
    netp = ref_model.OptimizedVanillaNeRF(10, 4, 256)
    netp.load_state_dict(peaked(sd))
    netp.eval()
    with torch.no_grad():
        rgb, sig = netp(pos, vd)
    out["mlp_peaked_rgb"], out["mlp_peaked_sigma"] = np_(rgb), np_(sig)

    # ---- raw2outputs (train.py:78-106 ; train_no_amp.py:78-105) --------------------------------
    rsr = np.random.RandomState(17)
    for tag, T, zz in (("c", S, z), ("f", S + I, z_all)):
        rgb_in = torch.from_numpy(rsr.uniform(0, 1, (B, T, 3)).astype(np.float32))
        sig_in = torch.from_numpy((rsr.uniform(0, 1, (B, T, 1)) ** 6 * 60).astype(np.float32))
        sig_in[:4] = 0.0
        sig_in[4:8, T // 2] = 1e4                       # alpha == 1 exactly -> T = 1e-10 path
        out[f"comp_{tag}_rgb"], out[f"comp_{tag}_sigma"] = np_(rgb_in), np_(sig_in)
        for wb in (False, True):
            rm, ww = ref_train.raw2outputs(rgb_in, sig_in, zz, wb)
            out[f"comp_{tag}_map_wb{int(wb)}"], out[f"comp_{tag}_w_wb{int(wb)}"] = np_(rm), np_(ww)
        rm, ww = ref_fp32.raw2outputs(rgb_in, sig_in, zz, False)
        out[f"comp_{tag}_map_1e6"], out[f"comp_{tag}_w_1e6"] = np_(rm), np_(ww)
        # backward of sum(rgb_map * G) wrt rgb and sigma
        G = torch.from_numpy(rsr.normal(size=(B, 3)).astype(np.float32))
        r2, s2 = rgb_in.clone().requires_grad_(), sig_in.clone().requires_grad_()
        rm, _ = ref_train.raw2outputs(r2, s2, zz, False)
        (rm * G).sum().backward()
        out[f"comp_{tag}_G"], out[f"comp_{tag}_drgb"], out[f"comp_{tag}_dsigma"] = np_(G), np_(r2.grad), np_(s2.grad)

    # ---- forward_pass + backward (train.py:129-186), hidden 64 so grads stay small ------------
    H = 64
    sdc, sdf = make_state_dict(31, H), make_state_dict(32, H)
    for mode, mod in (("fp32", ref_fp32), ("bf16", ref_train)):
        mc, mf = ref_model.OptimizedVanillaNeRF(10, 4, H), ref_model.OptimizedVanillaNeRF(10, 4, H)
        mc.load_state_dict(sdc)
        mf.load_state_dict(sdf)
        for train_mode in (False, True):
            mc.train(train_mode)
            mf.train(train_mode)
            mc.zero_grad()
            mf.zero_grad()
            nz_c = [make_noise(41, B * S)] if train_mode else []
            nz_f = [make_noise(42, B * (S + I))] if train_mode else []
            crit = torch.nn.MSELoss(reduction="mean")
            with injected_draws([t_rand, u], nz_c + nz_f) as log:
                if mode == "bf16":
                    with torch.amp.autocast("cpu"):
                        tot, lc, lf = mod.forward_pass(mc, mf, rays_o, rays_d, near, far, target,
                                                       sampler, crit, dev, False)
                else:
                    tot, lc, lf = mod.forward_pass(mc, mf, rays_o, rays_d, near, far, target,
                                                   sampler, crit, dev, False)
            tot.backward()
            tag = f"fp_{mode}_{'train' if train_mode else 'eval'}"
            out[tag + "_losses"] = np.array([tot.item(), lc.item(), lf.item()], np.float64)
            out[tag + "_inds"] = log["searchsorted"][0].numpy().astype(np.int64)
            for nm, m in (("c", mc), ("f", mf)):
                for k, p in m.named_parameters():
                    out[f"{tag}_grad_{nm}_{k}"] = np_(p.grad)

    # ---- dense grid + query (extract_occupancy_volume.py:90-172 == extract_vol.py:57-120) -----
    R = 12
    import io
    from contextlib import redirect_stdout
    with redirect_stdout(io.StringIO()):
        grid = ref_eov.generate_3d_grid((-1.0, 1.0), R)
    out["grid_R12"] = np_(grid)
    out["linspace_512"] = np_(torch.linspace(-1.0, 1.0, 512))
    net.eval()
    with redirect_stdout(io.StringIO()):
        dens, cols = ref_eov.query_density_and_rgb_batch(net, grid, batch_size=500, device="cpu",
                                                         show_progress=False)
    out["grid_sigma"], out["grid_rgb"] = np_(dens), np_(cols)

    # ---- AdamW + OneCycleLR (train.py:244-259, :321-322) ---------------------------------------
    p = torch.nn.Parameter(torch.from_numpy(np.random.RandomState(51).normal(size=(257,)).astype(np.float32)))
    opt = torch.optim.AdamW([p], lr=5e-4, weight_decay=1e-6)
    total = 50
    sch = torch.optim.lr_scheduler.OneCycleLR(opt, max_lr=5e-3, total_steps=total, pct_start=0.1,
                                              anneal_strategy="cos", div_factor=10, final_div_factor=100)
    lrs, ps = [], []
    gr = np.random.RandomState(52)
    gs = []
    for s in range(total - 1):
        g = torch.from_numpy(gr.normal(size=(257,)).astype(np.float32))
        gs.append(np_(g))
        p.grad = g
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        sch.step()
        ps.append(np_(p))
    out["adamw_p0"] = np.random.RandomState(51).normal(size=(257,)).astype(np.float32)
    out["adamw_grads"], out["adamw_lrs"], out["adamw_params"] = np.stack(gs), np.array(lrs, np.float64), np.stack(ps)

    path = os.path.join(HERE, "vanilla_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, f"{os.path.getsize(path)/1e6:.2f} MB,", len(out), "arrays")


if __name__ == "__main__":
    main()
