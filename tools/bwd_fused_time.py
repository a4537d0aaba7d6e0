"""Times the one-launch backward (csrc/mlp_bwd_fused.cu) against the two-launch path at the train step's sizes.
    DFN_BWD_SPLIT="pairs,big,pe,col" DFN_BWD_RING_SLOTS=48 python tools/bwd_fused_time.py [rays]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import deepfields_nerf_b200 as D  # noqa: E402
from common import make_rays, make_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
sd = make_state_dict(32, 256)
flat = torch.cat([sd[k].reshape(-1) for k in sd if "freq_bands" not in k]).cuda()
pk = D.ops.PackedMLP("cuda")
pk.pack(flat)
o, d, near, far, _ = make_rays(1, B)
g = torch.Generator(device="cuda").manual_seed(3)
out = {}
for tag, S in (("coarse", 64), ("fine", 192)):
    z = torch.sort(torch.rand(B, S, device="cuda", generator=g) * (far.cuda() - near.cuda()) + near.cuda(), dim=1).values.contiguous()
    rays = (o.cuda(), d.cuda(), z)
    P = B * S
    stash = torch.empty(D.ops.mlp_stash_bytes(P), dtype=torch.uint8, device="cuda")
    noise = torch.randn(P, device="cuda", generator=g)
    rgb, sig = D.ops.mlp_forward(pk, rays=rays, noise=noise, stash=stash)
    G_rgb = torch.randn(P, 3, device="cuda", generator=g) * 1e-3
    G_sig = torch.randn(P, device="cuda", generator=g) * 1e-3
    dz = torch.empty(D.ops.mlp_dz_bytes(P), dtype=torch.uint8, device="cuda")
    ws = torch.empty(D.ops.mlp_bwd_fused_ws_bytes(P), dtype=torch.uint8, device="cuda")
    g1 = torch.zeros(D.ops.MLP_NUM_PARAMS, device="cuda")
    g2 = torch.zeros(D.ops.MLP_NUM_PARAMS, device="cuda")

    def two():
        D.ops.mlp_backward(pk, rgb, sig, G_rgb, G_sig, stash, dz, g1, rays=rays)

    def one():
        D.ops.mlp_backward_fused(pk, rgb, sig, G_rgb, G_sig, stash, ws, g2, rays=rays)

    def timed(fn, reps=10):
        fn(); fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    def dgrad_only():
        D._lib.call("dfn_mlp_backward_dgrad", P, pk.bwd.data_ptr(), pk.consts.data_ptr(), rgb.data_ptr(), sig.data_ptr(),
                    G_rgb.data_ptr(), G_sig.data_ptr(), stash.data_ptr(), dz.data_ptr(), torch.cuda.current_stream().cuda_stream)

    td = timed(dgrad_only)
    print(f"{tag}: dgrad alone {td:.3f} ms (DFN_DGRAD_PAIRS={os.environ.get('DFN_DGRAD_PAIRS', 'all')})", flush=True)
    t2, t1 = timed(two), timed(one)
    err = float((g1 - g2).norm() / g1.norm())
    out[tag] = (t2, t1)
    print(f"{tag}: P={P} two-launch {t2:.3f} ms  fused {t1:.3f} ms  ratio {t2 / t1:.2f}  grad rel diff {err:.2e}", flush=True)
print(f"sum: two-launch {sum(v[0] for v in out.values()):.3f} ms fused {sum(v[1] for v in out.values()):.3f} ms "
      f"split={os.environ.get('DFN_BWD_SPLIT', 'default')} slots={os.environ.get('DFN_BWD_RING_SLOTS', 'default')}")
