"""GPU parity tests (run on the B200 box: pytest -m gpu).  Every CUDA op is called through the C ABI
(deepfields_nerf_b200.ops -> ctypes -> libdeepfields_b200.so) and compared with oracle/nerf_oracle.py on the
same seeded inputs / injected random draws, and with the committed reference-generated golden vectors."""
import numpy as np
import pytest
import torch

from common import make_draws, make_noise, make_rays, make_state_dict, peaked
from oracle import nerf_oracle as O

pytestmark = pytest.mark.gpu
T = lambda a: torch.from_numpy(np.asarray(a))  # noqa: E731


@pytest.fixture(scope="module")
def D():
    import deepfields_nerf_b200 as d
    assert torch.cuda.is_available()
    return d


def chunks(t):
    """[rows, cols] -> UMMA chunk image [cols/8][rows][8]."""
    r, c = t.shape
    return t.reshape(r, c // 8, 8).permute(1, 0, 2).contiguous()


def flat_params(sd):
    keys = [k for k in sd if "freq_bands" not in k]
    return torch.cat([sd[k].reshape(-1) for k in keys])


# ---- tcgen05 descriptor conventions ------------------------------------------------------------------
@pytest.mark.parametrize("N,K", [(256, 64), (256, 256), (128, 32), (16, 128)])
def test_umma_kmajor(D, N, K):
    g = torch.Generator().manual_seed(N + K)
    A = torch.randn(128, K, generator=g).bfloat16().cuda()
    B = torch.randn(N, K, generator=g).bfloat16().cuda()
    d = D.ops.umma_selftest(chunks(A), chunks(B), N, K, 0, 0, 128 * 16, 128, N * 16, 128, 2 * 128 * 16, 2 * N * 16)
    ref = A.float() @ B.float().t()
    torch.testing.assert_close(d, ref, rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("N,K", [(256, 128), (64, 128), (16, 128)])
def test_umma_mnmajor(D, N, K):
    """Both operands read 'transposed' from [points x features] tiles: D = At^T * Bt (wgrad form)."""
    g = torch.Generator().manual_seed(7 * N + K)
    At = torch.randn(K, 128, generator=g).bfloat16().cuda()   # [reduction rows, M]
    Bt = torch.randn(K, N, generator=g).bfloat16().cuda()     # [reduction rows, N]
    d = D.ops.umma_selftest(chunks(At), chunks(Bt), N, K, 1, 1, 128, K * 16, 128, K * 16, 256, 256)
    ref = At.float().t() @ Bt.float()
    torch.testing.assert_close(d, ref, rtol=1e-3, atol=1e-2)


# ---- sampling -------------------------------------------------------------------------------------------
def test_stratified_and_points_bit_exact(D, golden):
    B, S, I = 48, 64, 128
    o, d, near, far, _ = make_rays(11, B)
    t_rand, _ = make_draws(12, B, S, I)
    smp = D.StratifiedSampler(S, "cuda")
    pts, dirs, z = smp(o.cuda(), d.cuda(), near.cuda(), far.cuda(), t_rand.cuda())
    # torch.linspace on CUDA may round differently from the CPU one the golden used: compare through the
    # oracle fed with the sampler's own t grid as well as the golden
    assert np.array_equal(z.cpu().numpy(), golden["sampler_z"]) or torch.equal(
        z.cpu(), O.stratified_z(near, far, t_rand))
    z_ref = T(golden["sampler_z"]).cuda()
    pts2, dirs2 = D.ops.points_on_rays(o.cuda(), d.cuda(), z_ref)
    assert np.array_equal(pts2.cpu().numpy(), golden["sampler_pts"])
    assert np.array_equal(dirs2.cpu().numpy(), golden["sampler_dirs"])


def test_stratified_given_tvals_bit_exact(D):
    for B, S in ((1, 2), (5, 64), (333, 256), (4096, 64)):
        o, d, near, far, _ = make_rays(B + S, B)
        t_rand, _ = make_draws(B, B, S, 1)
        ref = O.stratified_z(near, far, t_rand)
        tv = torch.linspace(0.0, 1.0, S)
        z = D.ops.stratified_sample(near.cuda(), far.cuda(), tv.cuda(), t_rand.cuda())
        assert torch.equal(z.cpu(), ref), (B, S)


def test_sample_pdf_shared_cdf_bit_exact(D, golden):
    z = T(golden["sampler_z"])
    mids = 0.5 * (z[:, :-1] + z[:, 1:])
    w, u = T(golden["pdf_weights"]), T(golden["pdf_u"])
    cdf = O.pdf_to_cdf(w)
    s, inds = D.ops.sample_pdf_with_cdf(mids.cuda(), cdf.cuda(), u.cuda())
    assert np.array_equal(inds.cpu().numpy(), golden["pdf_inds"])          # tier A: bit-exact indices
    assert np.array_equal(s.cpu().numpy(), golden["pdf_samples"])


def test_sample_pdf_fused(D, golden):
    z = T(golden["sampler_z"])
    mids = 0.5 * (z[:, :-1] + z[:, 1:])
    w, u = T(golden["pdf_weights"]), T(golden["pdf_u"])
    s, inds, cdf = D.ops.sample_pdf(mids.cuda(), w.cuda(), u.cuda(), return_inds=True, return_cdf=True)
    s2 = D.ops.sample_pdf(z.cuda(), w.cuda(), u.cuda(), bins_mid_of_z=True)
    assert torch.equal(s, s2)
    ref_cdf = O.pdf_to_cdf(w)
    assert (cdf.cpu() - ref_cdf).abs().max() <= 6e-7   # a few ulp(1): fp32 sum order differs from torch's
    # tier A on the kernel's own cdf: searchsorted indices and samples are BIT-EXACT with torch (train.py:56-73)
    s_chk, i_chk = O.invert_cdf(mids, cdf.cpu(), u)
    assert torch.equal(inds.cpu(), i_chk)
    assert torch.equal(s.cpu(), s_chk)
    ref_inds = T(golden["pdf_inds"])
    bad = (inds.cpu() != ref_inds)
    # tier B: the fused cdf may differ from torch's CPU cdf by an ulp (sum order); every index flip against the
    # reference-generated golden must be an edge case (u within an ulp of a cdf edge)
    if bad.any():
        uu, cc = u[bad], torch.gather(ref_cdf, 1, ref_inds.clamp(max=62))[bad]
        lo = torch.gather(ref_cdf, 1, (ref_inds - 1).clamp(min=0))[bad]
        assert torch.minimum((uu - cc).abs(), (uu - lo).abs()).max() <= 7e-7
        # the golden puts u = 1 - 2^-24 on the top edge of EVERY row (cdf[-1] = 1 +- 1 ulp), hence up to B flips
        assert bad.sum() <= 2 * u.shape[0]
    # where the index agrees, the sample may still move with the cdf's last bit: t = (u - c0) / (c1 - c0) amplifies an
    # ulp of the cdf by 1 / (c1 - c0) (>= 1e-5 by train.py:71), so the bound is per sample
    ok = ~bad
    c1 = torch.gather(ref_cdf, 1, ref_inds.clamp(max=62))
    c0 = torch.gather(ref_cdf, 1, (ref_inds - 1).clamp(min=0))
    b1 = torch.gather(mids, 1, ref_inds.clamp(max=62))
    b0 = torch.gather(mids, 1, (ref_inds - 1).clamp(min=0))
    denom = (c1 - c0)
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    ref_s = T(golden["pdf_samples"])
    tol = 1e-5 * ref_s.abs() + 2e-6 + 2.5e-7 * (b1 - b0).abs() / denom
    assert bool(((s.cpu() - ref_s).abs() <= tol)[ok].all())
    # a strided weights view (train.py:158 passes weights[..., 1:-1])
    wfull = torch.zeros(w.shape[0], 64)
    wfull[:, 1:-1] = w
    s3 = D.ops.sample_pdf(mids.cuda(), wfull.cuda()[:, 1:-1], u.cuda())
    assert torch.equal(s, s3)


def test_sample_pdf_unnormalised_weights(D):
    """dfn_sample_pdf is a public entry point and takes any non-negative weights, not only compositing weights (row sum
    <= 1, the case whose cdf prefixes are exact in double).  With rows that sum to ~1e6 and entries spread over 12 decades
    the warp scan's partial sums round, so the prefixes are no longer guaranteed bit-identical to a sequential double cumsum -
    they must stay within an fp32 ulp of it, monotone, end at 1, and the inversion on the kernel's own cdf stays bit-exact."""
    g = torch.Generator().manual_seed(5)
    B, NB, NS = 300, 63, 128
    z = torch.sort(torch.rand(B, NB + 1, generator=g) * 4 + 2, dim=1).values
    mids = 0.5 * (z[:, :-1] + z[:, 1:])
    w = 10.0 ** (torch.rand(B, NB - 1, generator=g) * 12 - 6)          # 1e-6 .. 1e6
    w[::7, 5] = 0.0
    u = torch.rand(B, NS, generator=g)
    s, inds, cdf = D.ops.sample_pdf(mids.cuda(), w.cuda(), u.cuda(), return_inds=True, return_cdf=True)
    cdf = cdf.cpu()
    pdf = (w + 1e-5) / (w + 1e-5).sum(-1, keepdim=True)                # train.py:42-43 in fp32
    ref = torch.cat([torch.zeros(B, 1), torch.cumsum(pdf.double(), -1).float()], -1)
    assert float((cdf - ref).abs().max()) <= 6e-7                      # same bar as the normalised case: a few ulp(1)
    assert bool((cdf[:, 1:] >= cdf[:, :-1]).all()) and bool((cdf[:, 0] == 0).all())
    assert float((cdf[:, -1] - 1).abs().max()) <= 6e-7
    s_chk, i_chk = O.invert_cdf(mids, cdf, u)
    assert torch.equal(inds.cpu(), i_chk) and torch.equal(s.cpu(), s_chk)
    assert bool(torch.isfinite(s).all())


def test_merge_sort_bit_exact(D, golden):
    zz = D.ops.merge_sort_z(T(golden["sampler_z"]).cuda(), T(golden["pdf_samples"]).cuda())
    assert np.array_equal(zz.cpu().numpy(), golden["fine_z"])
    g = torch.Generator().manual_seed(3)
    for B, NA, NB in ((1, 1, 1), (7, 64, 192), (100, 0, 5), (33, 17, 0)):
        a = torch.rand(B, NA, generator=g).sort(-1).values
        b = torch.rand(B, NB, generator=g)
        out = D.ops.merge_sort_z(a.cuda(), b.cuda())
        assert torch.equal(out.cpu(), torch.sort(torch.cat([a, b], -1), -1).values)


def test_merge_sort_unsorted_and_long_first_list(D):
    """the rank merge needs an ascending first list; anything else must take the generic network - same values either way"""
    g = torch.Generator().manual_seed(4)
    for B, NA, NB in ((50, 64, 128), (9, 200, 40), (5, 128, 128), (3, 100, 156)):
        a = torch.rand(B, NA, generator=g)                        # NOT sorted
        a[0] = a[0].sort().values                                # one ascending row, with duplicates across the lists
        b = torch.rand(B, NB, generator=g)
        b[0, : min(NA, NB)] = a[0, : min(NA, NB)]
        out = D.ops.merge_sort_z(a.cuda(), b.cuda())
        assert torch.equal(out.cpu(), torch.sort(torch.cat([a, b], -1), -1).values)


@pytest.mark.parametrize("B,NA,NS", [(48, 64, 128), (1000, 64, 128), (37, 64, 192), (5, 3, 1), (21, 128, 256), (64, 33, 77)])
def test_sample_merge_equals_sample_pdf_then_merge(D, B, NA, NS):
    """dfn_sample_merge (one launch, the train step / render path) is bit-identical to dfn_sample_pdf on the mid-points and
    weights[:, 1:-1] followed by dfn_merge_sort_z (scripts/train.py:157-158 + scripts/sampling.py:40)."""
    g = torch.Generator().manual_seed(B + NS)
    z = (torch.rand(B, NA, generator=g) * 4 + 2).sort(-1).values.cuda()
    w = (torch.rand(B, NA, generator=g) ** 6).cuda()
    w[: B // 4] = 0.0                                              # all-zero rows -> uniform pdf
    u = torch.rand(B, NS, generator=g).cuda()
    u[:, 0] = 0.0
    z_new_ref = D.ops.sample_pdf(z, w[:, 1:-1], u, bins_mid_of_z=True)
    z_all_ref = (D.ops.merge_sort_z(z, z_new_ref) if NA + NS <= 256                 # (dfn_merge_sort_z takes at most 256 keys)
                 else torch.sort(torch.cat([z, z_new_ref], -1), -1).values)
    z_all, z_new = D.ops.sample_merge(z, w, u, want_new=True)
    assert torch.equal(z_new, z_new_ref) and torch.equal(z_all, z_all_ref)
    assert torch.equal(D.ops.sample_merge(z, w, u), z_all_ref)
    assert torch.equal(z_all_ref, torch.sort(torch.cat([z, z_new_ref], -1), -1).values)
    # coarse depths that are not ascending (a ray with far < near): the generic network gives the same sorted values
    zr = z.flip(-1).contiguous()
    z_all_r, z_new_r = D.ops.sample_merge(zr, w, u, want_new=True)
    assert torch.equal(z_new_r, D.ops.sample_pdf(zr, w[:, 1:-1], u, bins_mid_of_z=True))
    assert torch.equal(z_all_r, torch.sort(torch.cat([zr, z_new_r], -1), -1).values)


def test_posenc(D, golden):
    x = T(golden["pe_in"]).cuda()
    for L, key in ((10, "pe_L10"), (4, "pe_L4")):
        enc = D.OptimizedPositionalEncoding(L).cuda()
        out = enc(x)
        assert out.shape[-1] == enc.output_dim
        np.testing.assert_allclose(out.cpu().numpy(), golden[key], atol=2e-6, rtol=0)


# ---- compositing --------------------------------------------------------------------------------------------
@pytest.mark.parametrize("tag,zkey", [("c", "sampler_z"), ("f", "fine_z")])
def test_composite_forward_backward(D, golden, tag, zkey):
    rgb, sig, z = T(golden[f"comp_{tag}_rgb"]), T(golden[f"comp_{tag}_sigma"]), T(golden[zkey])
    for wb in (False, True):
        rm, w = D.raw2outputs(rgb.cuda(), sig.cuda(), z.cuda(), wb)
        np.testing.assert_allclose(rm.cpu().numpy(), golden[f"comp_{tag}_map_wb{int(wb)}"], rtol=1e-5, atol=1e-7)
        # alpha = 1 - exp(-x) carries an absolute error of one ulp(1) = 6e-8 from exp's last bit (CPU vs GPU libm)
        np.testing.assert_allclose(w.cpu().numpy(), golden[f"comp_{tag}_w_wb{int(wb)}"], rtol=1e-5, atol=1.2e-7)
    rm, w = D.raw2outputs(rgb.cuda(), sig.cuda(), z.cuda(), False, 1e6)
    np.testing.assert_allclose(w.cpu().numpy(), golden[f"comp_{tag}_w_1e6"], rtol=1e-5, atol=1.2e-7)
    r2, s2 = rgb.cuda().requires_grad_(), sig.cuda().requires_grad_()
    rm, _ = D.raw2outputs(r2, s2, z.cuda())
    (rm * T(golden[f"comp_{tag}_G"]).cuda()).sum().backward()
    # d_rgb = w * G inherits the one-ulp(1) absolute error of alpha (see above), |G| ~ N(0,1)
    np.testing.assert_allclose(r2.grad.cpu().numpy(), golden[f"comp_{tag}_drgb"], rtol=1e-5, atol=5e-7)
    ref = golden[f"comp_{tag}_dsigma"]
    np.testing.assert_allclose(s2.grad.cpu().numpy(), ref, rtol=1e-4, atol=1e-6 * np.abs(ref).max())
    # extras + white background + weight gradients against oracle autograd
    r3, s3 = rgb.clone().requires_grad_(), sig.clone().requires_grad_()
    out = O.composite(r3, s3, z, True)
    gw = torch.randn(out["weights"].shape, generator=torch.Generator().manual_seed(1))
    gm = torch.randn(out["rgb_map"].shape, generator=torch.Generator().manual_seed(2))
    ((out["rgb_map"] * gm).sum() + (out["weights"] * gw).sum()).backward()
    d_rgb, d_sig = D.ops.composite_backward(rgb.cuda(), sig.cuda(), z.cuda(), gm.cuda(), gw.cuda(), True)
    torch.testing.assert_close(d_rgb.cpu(), r3.grad, rtol=1e-5, atol=5e-7)
    torch.testing.assert_close(d_sig.cpu().reshape(s3.shape), s3.grad, rtol=1e-4, atol=1e-6 * float(s3.grad.abs().max()))
    rm, w, acc, dep = D.ops.composite_forward(rgb.cuda(), sig.cuda(), z.cuda(), want_extras=True)
    o2 = O.composite(rgb, sig, z)
    torch.testing.assert_close(acc.cpu(), o2["acc"], rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(dep.cpu(), o2["depth"], rtol=1e-5, atol=1e-6)


def test_render_loss_fused(D, golden):
    rgb, sig, z = T(golden["comp_f_rgb"]), T(golden["comp_f_sigma"]), T(golden["fine_z"])
    B = z.shape[0]
    tgt = torch.rand(B, 3, generator=torch.Generator().manual_seed(5))
    r, s = rgb.clone().requires_grad_(), sig.clone().requires_grad_()
    out = O.composite(r, s, z)
    loss = torch.nn.functional.mse_loss(out["rgb_map"], tgt)
    loss.backward()
    ls = torch.zeros(1, device="cuda")
    rm, w, d_rgb, d_sig = D.ops.render_loss(rgb.cuda(), sig.cuda(), z.cuda(), tgt.cuda(), ls)
    assert abs(ls.item() / (3 * B) - loss.item()) <= 1e-5 * loss.item()
    torch.testing.assert_close(d_rgb.cpu(), r.grad, rtol=1e-4, atol=1e-8)
    torch.testing.assert_close(d_sig.cpu().reshape(s.shape), s.grad, rtol=1e-4, atol=1e-6 * float(s.grad.abs().max()))
    # bf16 inputs are accepted as-is (MLP outputs)
    rm2, w2 = D.ops.composite_forward(rgb.cuda().bfloat16(), sig.cuda().bfloat16(), z.cuda())
    o2 = O.composite(rgb.bfloat16(), sig.bfloat16(), z)
    torch.testing.assert_close(rm2.cpu(), o2["rgb_map"], rtol=1e-5, atol=1e-7)


# ---- MLP forward ------------------------------------------------------------------------------------------------
def _packed(D, sd):
    pk = D.ops.PackedMLP("cuda")
    pk.pack(flat_params(sd).cuda())
    return pk


def test_mlp_forward_golden(D, golden):
    sd = make_state_dict(21, 256)
    pos, view = T(golden["mlp_pos"]).cuda(), T(golden["mlp_view"]).cuda()
    rgb, sig = D.ops.mlp_forward(_packed(D, sd), pos=pos, view=view)
    # north_star tolerance for the bf16 MLP: 2e-2 against the reference (fp32 and bf16-autocast runs)
    for tag in ("fp32", "bf16"):
        np.testing.assert_allclose(rgb.cpu().numpy(), golden[f"mlp_{tag}_rgb"], atol=2e-2, rtol=0)
        np.testing.assert_allclose(sig.cpu().numpy(), golden[f"mlp_{tag}_sigma"][:, 0], atol=2e-2, rtol=2e-2)
    # same rounding points as the bf16 oracle -> much tighter
    r_o, s_o = O.mlp_forward(sd, pos.cpu(), view.cpu(), low=torch.bfloat16)
    assert (rgb.cpu() - r_o.float()).abs().max() < 8e-3
    assert (sig.cpu() - s_o.float()[:, 0]).abs().max() < 8e-3
    # training-mode noise and the peaked-sigma weight set
    noise = make_noise(16, 512)
    rgb, sig = D.ops.mlp_forward(_packed(D, sd), pos=pos, view=view, noise=noise.cuda())
    np.testing.assert_allclose(sig.cpu().numpy(), golden["mlp_fp32_train_sigma"][:, 0], atol=2e-2, rtol=2e-2)
    rgb, sig = D.ops.mlp_forward(_packed(D, peaked(sd)), pos=pos, view=view)
    np.testing.assert_allclose(sig.cpu().numpy(), golden["mlp_peaked_sigma"][:, 0], atol=0.3, rtol=2e-2)


@pytest.mark.parametrize("P", [1, 127, 128, 129, 1000, 148 * 256 + 77])
def test_mlp_forward_ragged_sizes(D, P):
    sd = make_state_dict(22, 256)
    g = torch.Generator().manual_seed(P)
    pos = torch.rand(P, 3, generator=g) * 2 - 1
    view = torch.nn.functional.normalize(torch.randn(P, 3, generator=g), dim=-1)
    rgb, sig = D.ops.mlp_forward(_packed(D, sd), pos=pos.cuda(), view=view.cuda())
    n = min(P, 4096)
    idx = torch.randperm(P, generator=g)[:n]
    r_o, s_o = O.mlp_forward(sd, pos[idx], view[idx], low=torch.bfloat16)
    assert (rgb.cpu()[idx] - r_o.float()).abs().max() < 1e-2
    assert (sig.cpu()[idx] - s_o.float()[:, 0]).abs().max() < 1e-2


def test_mlp_forward_sources_and_module(D, golden):
    sd = make_state_dict(21, 256)
    B, S = 37, 64
    o, d, near, far, _ = make_rays(3, B)
    t_rand, _ = make_draws(4, B, S, 1)
    z = O.stratified_z(near, far, t_rand)
    pts, dirs = O.points_on_rays(o, d, z)
    pk = _packed(D, sd)
    a = D.ops.mlp_forward(pk, pos=pts.cuda(), view=dirs.cuda())
    b = D.ops.mlp_forward(pk, rays=(o.cuda(), d.cuda(), z.cuda()))
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    # dense grid source == explicit grid points (extract_vol.py:57-120)
    R = 12
    lin = torch.linspace(-1.0, 1.0, R, device="cuda")
    view = torch.tensor([0.0, 0.0, -1.0], device="cuda")
    g1 = D.ops.mlp_forward(pk, grid=(lin, view, 0, R ** 3))
    gp = D.generate_3d_grid((-1.0, 1.0), R)
    assert np.array_equal(gp.cpu().numpy(), golden["grid_R12"]) or torch.equal(gp.cpu(), O.grid_points((-1.0, 1.0), R))
    g2 = D.ops.mlp_forward(pk, pos=gp, view=view.expand(R ** 3, 3).contiguous())
    assert torch.equal(g1[0], g2[0]) and torch.equal(g1[1], g2[1])
    np.testing.assert_allclose(g1[1].cpu().numpy(), golden["grid_sigma"][:, 0], atol=2e-2, rtol=2e-2)
    np.testing.assert_allclose(g1[0].cpu().numpy(), golden["grid_rgb"], atol=2e-2, rtol=0)
    # slab offset
    g3 = D.ops.mlp_forward(pk, grid=(lin, view, 5 * R * R, 2 * R * R))
    assert torch.equal(g3[1], g1[1][5 * R * R: 7 * R * R])
    # nn.Module surface: state-dict round trip with the reference key set
    m = D.OptimizedVanillaNeRF().cuda()
    assert list(m.state_dict().keys()) == list(sd.keys())
    m.load_state_dict(sd)
    m.eval()
    with torch.no_grad():
        rgb, sig = m(pts.cuda(), dirs.cuda())
        sg, feat = m.query_density(pts.cuda())
    assert rgb.shape == (B * S, 3) and sig.shape == (B * S, 1) and feat.shape == (B * S, 256)
    assert torch.equal(rgb, a[0]) and torch.equal(sig[:, 0], a[1]) and torch.equal(sg, sig)
    np.testing.assert_allclose(feat.float().cpu().numpy()[:512], O_feat(sd, pts[:512]), atol=2e-2, rtol=2e-2)


def O_feat(sd, pos):
    pe = O.posenc(pos, 10)
    h = pe
    for i in (0, 2, 4, 6):
        h = torch.relu(O._lin(h, sd[f"pos_layers.{i}.weight"], sd[f"pos_layers.{i}.bias"], None))
    h = torch.cat([h, pe], -1)
    for i in (0, 2, 4, 6):
        h = torch.relu(O._lin(h, sd[f"skip_layers.{i}.weight"], sd[f"skip_layers.{i}.bias"], None))
    return O._lin(h, sd["feature_layer.weight"], sd["feature_layer.bias"], None).numpy()


# ---- optimiser / volume -----------------------------------------------------------------------------------------------
def test_adamw_onecycle(D, golden):
    total = 50
    lrs, b1s = D.onecycle_tables(total, 5e-4)
    np.testing.assert_allclose(lrs[: total - 1], golden["adamw_lrs"], rtol=1e-9)
    p = T(golden["adamw_p0"]).clone().cuda()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    step = torch.zeros(1, dtype=torch.int32, device="cuda")
    lt, bt = torch.tensor(lrs, device="cuda", dtype=torch.float32), torch.tensor(b1s, device="cuda", dtype=torch.float32)
    for s in range(total - 1):
        D.ops.adamw_step(p, T(golden["adamw_grads"][s]).cuda(), m, v, lt, bt, step)
        np.testing.assert_allclose(p.cpu().numpy(), golden["adamw_params"][s], rtol=1e-5, atol=2e-6)
    assert step.item() == total - 1


def test_histogram_otsu_pack(D):
    g = torch.Generator().manual_seed(9)
    v = torch.cat([torch.rand(50000, generator=g) * 0.2, 5 + torch.randn(30000, generator=g)])
    rgb = torch.rand(v.numel(), 3, generator=g)
    occ, thr = D.apply_otsu_threshold(v.cuda())
    ref_thr = O.otsu_threshold(v)
    assert abs(thr - ref_thr) <= (float(v.max()) - float(v.min())) / 256 + 1e-6
    o8, c8 = D.ops.pack_volume(v.cuda(), rgb.cuda(), thr)
    assert torch.equal(o8.cpu(), (v > thr).to(torch.uint8))
    assert torch.equal(c8.cpu(), (rgb * 255).to(torch.uint8))
    h = D.ops.histogram(v.cuda(), float(v.min()), float(v.max()), 256)
    assert int(h.sum()) == v.numel()


def test_errors_are_loud(D):
    with pytest.raises(RuntimeError):
        D.ops.stratified_sample(torch.zeros(4), torch.ones(4), torch.linspace(0, 1, 8), torch.rand(4, 8))  # CPU tensors
    with pytest.raises(RuntimeError):
        D.ops.merge_sort_z(torch.zeros(2, 200, device="cuda"), torch.zeros(2, 100, device="cuda"))
    with pytest.raises(NotImplementedError):
        m = D.OptimizedVanillaNeRF(hidden_dim=64).cuda()
        m(torch.zeros(4, 3, device="cuda"), torch.zeros(4, 3, device="cuda"))


# ---- ray generation / cube intersection (scripts/precompute_rays.py) ---------------------------------
def _rays_golden():
    import os
    return dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "rays_golden.npz")))


def test_generate_rays_golden(D):
    from common import make_cameras
    g = _rays_golden()
    H, W, K, poses, _ = make_cameras(7, M=3, H=12, W=16)
    o, d = D.get_rays(H, W, K, poses)
    assert o.shape == (3, H, W, 3) and o.dtype == torch.float32
    assert np.array_equal(o.cpu().numpy(), g["rays_o"].astype(np.float32))
    want = g["rays_d"].astype(np.float32)
    got = d.cpu().numpy()
    # fp64 arithmetic in another association than numpy's BLAS: equal after the fp32 cast up to a rare 1-ulp tie
    assert np.abs(got - want).max() <= 6e-8 and (got == want).mean() > 0.99
    o2, d2 = O.camera_rays(H, W, K, poses, colmap=False, normalize=False)
    o3, d3 = D.get_rays(H, W, K, poses, colmap=False, normalize_directions=False)
    assert np.allclose(d3.cpu().numpy(), d2.astype(np.float32), rtol=2e-7, atol=0)


def test_ray_cube_intersection_bit_exact(D):
    from common import edge_case_rays
    g = _rays_golden()
    o32, d32 = g["rays_o"].reshape(-1, 3).astype(np.float32), g["rays_d"].reshape(-1, 3).astype(np.float32)
    hits, t_near, t_far = D.ray_cube_intersection_numpy(T(o32).cuda(), T(d32).cuda())
    assert np.array_equal(hits.cpu().numpy(), g["hits"])
    assert np.array_equal(t_near.cpu().numpy(), g["t_near"], equal_nan=True)
    assert np.array_equal(t_far.cpu().numpy(), g["t_far"], equal_nan=True)
    eo, ed = edge_case_rays()
    hits, t_near, t_far = D.ray_cube_intersection(T(eo).cuda(), T(ed).cuda(), -0.5, 0.75)
    assert np.array_equal(hits.cpu().numpy(), g["edge_hits"])
    assert np.array_equal(t_near.cpu().numpy(), g["edge_near"], equal_nan=True)
    assert np.array_equal(t_far.cpu().numpy(), g["edge_far"], equal_nan=True)
    # random rays against the oracle, bit for bit
    rng = np.random.default_rng(5)
    o = rng.uniform(-3, 3, size=(100003, 3)).astype(np.float32)
    dd = rng.normal(size=(100003, 3)).astype(np.float32)
    dd[::7, 1] = 0.0
    dd[::11, 0] = 1e-9
    oh, on, of = O.ray_cube_hits(o, dd)
    hits, t_near, t_far = D.ray_cube_intersection(T(o).cuda(), T(dd).cuda())
    assert np.array_equal(hits.cpu().numpy(), oh)
    assert np.array_equal(t_near.cpu().numpy(), on, equal_nan=True) and np.array_equal(t_far.cpu().numpy(), of, equal_nan=True)


def test_precompute_rays_fused(D):
    from common import make_cameras
    g = _rays_golden()
    H, W, K, poses, images = make_cameras(7, M=3, H=12, W=16)
    out = D.precompute_rays_on_device(H, W, K, poses, T(images).cuda())
    k = int(g["hits"].sum())
    assert out["rays_o"].shape == (k, 3) and out["t_near"].shape == (k,)
    assert np.array_equal(out["rays_o"].cpu().numpy(), g["kept_o"].astype(np.float32))
    assert np.abs(out["rays_d"].cpu().numpy() - g["kept_d"].astype(np.float32)).max() <= 6e-8
    assert np.array_equal(out["rgbs"].cpu().numpy(), g["kept_rgb"])
    assert np.allclose(out["t_near"].cpu().numpy(), g["kept_near"], rtol=1e-5, atol=1e-6)
    assert np.allclose(out["t_far"].cpu().numpy(), g["kept_far"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("M,H,W", [(1, 1, 1), (2, 31, 33), (3, 1080, 1920)])
def test_precompute_rays_equals_unfused_path(D, M, H, W):
    """Size-independent property at full size: the fused generate + intersect + compaction equals generating every ray,
    intersecting it and boolean-masking, bit for bit and in order (ragged chunk tails, multi-tile scan of the counts)."""
    from common import make_cameras
    _, _, K, poses, images = make_cameras(3 + M, M=M, H=H, W=W, focal=0.45 * W)
    img = T(images).cuda()
    o, d = D.get_rays(H, W, K, poses)
    hits, t_near, t_far = D.ray_cube_intersection(o, d)
    out = D.precompute_rays_on_device(H, W, K, poses, img)
    k = int(hits.sum())
    assert out["rays_o"].shape[0] == k
    if M > 1:
        assert 0 < k < M * H * W
    assert torch.equal(out["rays_o"], o.reshape(-1, 3)[hits]) and torch.equal(out["rays_d"], d.reshape(-1, 3)[hits])
    assert torch.equal(out["t_near"], t_near[hits]) and torch.equal(out["t_far"], t_far[hits])
    want_rgb = (images.reshape(-1, 3).astype(np.float32) / np.float32(255.0))[hits.cpu().numpy()]   # IEEE division, as numpy does
    assert np.array_equal(out["rgbs"].cpu().numpy(), want_rgb)
    assert bool((out["t_near"] >= 0).all()) and bool((out["t_near"] <= out["t_far"]).all())


def test_precompute_rays_no_hits_and_store(D):
    from common import make_cameras
    H, W, K, poses, images = make_cameras(9, M=2, H=8, W=8)
    away = poses.copy()
    away[:, :3, 2] *= -1.0   # look the other way: nothing hits the cube
    away[:, :3, 0] *= -1.0
    out = D.precompute_rays_on_device(H, W, K, away, T(images).cuda())
    assert out["rays_o"].shape == (0, 3) and out["rgbs"].shape == (0, 3)
    out = D.precompute_rays_on_device(H, W, K, poses, T(images).cuda())
    store = D.RayStore.from_tensors(out)
    assert len(store) == out["rays_o"].shape[0] > 0
    with pytest.raises(ValueError):
        D.precompute_rays_on_device(H, W, K, poses, T(images).cuda().float())
    with pytest.raises(RuntimeError):
        D.get_rays(H, W, K, poses, device="cpu")


# ---- occupied voxels as points (scripts/post_process_vol.py) -----------------------------------------
VOL_AABB = {"aabb_min": [-0.55, -0.4, -0.7], "aabb_max": [0.6, 1.0, 0.22580645161290325]}


def test_volume_points_golden_bit_exact(D):
    import os
    from common import make_volume
    g = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "volume_points_golden.npz")))
    occ, rgb = make_volume(5, 32)
    pts, cols = D.volume_points(T(occ).cuda(), T(rgb).cuda())
    assert pts.dtype == torch.float64 and cols.dtype == torch.float32
    assert np.array_equal(pts.cpu().numpy(), g["points"]) and np.array_equal(cols.cpu().numpy(), g["rgbs"])
    pts, cols = D.volume_points(T(occ).cuda(), T(rgb).cuda(), VOL_AABB)
    assert np.array_equal(pts.cpu().numpy(), g["sliced_points"]) and np.array_equal(cols.cpu().numpy(), g["sliced_rgbs"])
    p2, c2 = D.slice_aabb(T(g["points"]).cuda(), T(g["rgbs"]).cuda(), VOL_AABB)
    assert np.array_equal(p2.cpu().numpy(), g["sliced_points"]) and np.array_equal(c2.cpu().numpy(), g["sliced_rgbs"])


@pytest.mark.parametrize("R", [2, 33, 160])
def test_volume_points_vs_oracle_and_edges(D, R):
    from common import make_volume
    occ, rgb = make_volume(40 + R, R)
    box = ([-0.3, -1.0, -0.25], [0.9, 0.5, 1.0])
    for lo, hi in ((None, None), box):
        want_p, want_c = O.volume_points(occ, rgb, lo, hi)
        meta = None if lo is None else {"aabb_min": lo, "aabb_max": hi}
        pts, cols = D.volume_points(T(occ).cuda(), T(rgb).cuda(), meta)
        assert np.array_equal(pts.cpu().numpy(), want_p) and np.array_equal(cols.cpu().numpy(), want_c)
    empty = np.zeros_like(occ)
    pts, cols = D.volume_points(T(empty).cuda(), T(rgb).cuda())
    assert pts.shape == (0, 3) and cols.shape == (0, 3)
    full = np.ones_like(occ)
    pts, cols = D.volume_points(T(full).cuda(), T(rgb).cuda())
    assert pts.shape == (R ** 3, 3) and float(pts.min()) == -1.0 and float(pts.max()) == 1.0
    assert bool((pts[1:, 0] >= pts[:-1, 0]).all())   # C order: x slowest
    with pytest.raises(ValueError):
        D.volume_points(T(occ).cuda(), T(rgb).cuda().float())


def test_volume_points_full_size_properties(D):
    """512^3 (BASELINE config 4 size): count, order and values against torch.nonzero on the device; the sliced set is the
    masked raw set."""
    R = 512
    g = torch.Generator(device="cuda").manual_seed(3)
    occ = (torch.rand(R, R, R, device="cuda", generator=g) < 0.07).to(torch.uint8)
    rgb = torch.randint(0, 256, (R, R, R, 3), device="cuda", dtype=torch.uint8, generator=g)
    pts, cols = D.volume_points(occ, rgb)
    idx = torch.nonzero(occ)
    assert pts.shape[0] == idx.shape[0] == int(occ.sum())
    want = torch.div(idx.double(), torch.full_like(idx, R - 1, dtype=torch.float64)) * 2 - 1
    assert torch.equal(pts, want)
    want_c = torch.div(rgb[idx[:, 0], idx[:, 1], idx[:, 2]].float(), torch.full((1,), float(rgb.max()), device="cuda"))
    assert torch.equal(cols, want_c)
    meta = {"aabb_min": [-0.5, -0.25, -1.0], "aabb_max": [0.25, 0.75, 0.0]}
    sp, sc = D.volume_points(occ, rgb, meta)
    p2, c2 = D.slice_aabb(pts, cols, meta)
    assert torch.equal(sp, p2) and torch.equal(sc, c2) and 0 < sp.shape[0] < pts.shape[0]
