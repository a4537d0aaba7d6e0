"""One fused query + one fused backward of the Instant-NGP variant at 2^22 points along rays (ncu target, tools/gpu_profile_r2.sh)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import deepfields_nerf_b200 as D  # noqa: E402
from common import make_rays  # noqa: E402

B, S = 1 << 16, 64
o, d, near, far, _ = (t.cuda() for t in make_rays(3, B))
torch.manual_seed(0)
net = D.InstantNGP("large").cuda()
with torch.no_grad():
    net.pos_encoding.params.uniform_(-1.0, 1.0)
tab, dw, cw = net._half_params()
z = D.ops.stratified_sample(near.reshape(-1), far.reshape(-1), torch.linspace(0, 1, S, device="cuda"), torch.rand(B, S, device="cuda"))
pts = D.ops.points_on_rays(o, d, z, want_dirs=False)
g_t = torch.zeros(2 * net.grid.n_entries, device="cuda")
g_d = torch.zeros(D.ops.NGP_DMLP_PARAMS, device="cuda")
g_c = torch.zeros(D.ops.NGP_CMLP_PARAMS, device="cuda")
G_rgb = torch.randn(B * S, 3, device="cuda") * 1e-3
G_sig = torch.randn(B * S, device="cuda") * 1e-3
for _ in range(3):
    rgb, sig = D.ops.ngp_forward_f32(pts, d, tab, net.grid, dw, cw, dirs_group=S)
    D.ops.ngp_backward(pts, d, tab, net.grid, dw, cw, G_rgb, G_sig, g_t, g_d, g_c, dirs_group=S)
torch.cuda.synchronize()
print("ok", float(rgb.float().mean()), float(g_t.abs().sum()))
