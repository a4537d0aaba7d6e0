"""A/B timing of the Instant-NGP kernels (CUDA events, 5 repetitions after a warm-up); bring-up tool, not the bench."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepfields_nerf_b200 as D  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(7)
net = D.InstantNGP("large").to(dev)
with torch.no_grad():
    net.pos_encoding.params.uniform_(-1.0, 1.0)
P = 1 << 24
pos, dirs = torch.rand(P, 3, device=dev), torch.rand(P, 3, device=dev)
table, dw, cw = net._half_params()


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


ms = timed(lambda: D.ops.ngp_forward(pos, dirs, table, net.grid, dw, cw))
print(f"ngp_forward random  {P} pts: {ms:.3f} ms  {P / ms / 1e6:.2f} G pts/s")
ms = timed(lambda: D.ops.ngp_forward(pos, None, table, net.grid, dw, cw, want_rgb=False))
print(f"ngp density only    {P} pts: {ms:.3f} ms  {P / ms / 1e6:.2f} G pts/s")
# 256^3 dense grid (extract_vol order): neighbouring points share cells on the coarse levels
lin = torch.linspace(0, 1, 256, device=dev)
grid = D.ops.grid_points(lin, 0, 256 ** 3)
view = torch.tensor([0.5, 0.5, 0.0], device=dev)
ms = timed(lambda: D.ops.ngp_forward(grid, view, table, net.grid, dw, cw))
print(f"ngp_forward 256^3 grid     : {ms:.3f} ms  {256 ** 3 / ms / 1e6:.2f} G pts/s")
Pe = 1 << 22
ms = timed(lambda: D.ops.hashgrid_forward(pos[:Pe], table, net.grid))
print(f"hashgrid_forward    {Pe} pts: {ms:.3f} ms  {Pe / ms / 1e6:.2f} G pts/s")
enc = D.ops.hashgrid_forward(pos[:Pe], table, net.grid)
dt = torch.zeros(2 * net.grid.n_entries, device=dev)
ms = timed(lambda: D.ops.hashgrid_backward(pos[:Pe], enc, net.grid, d_table=dt))
print(f"hashgrid_backward   {Pe} pts: {ms:.3f} ms  {Pe / ms / 1e6:.2f} G pts/s")
