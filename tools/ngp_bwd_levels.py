"""Which levels cost what in the hash-grid backward: time it with the gradient of a level range zeroed (those threads skip
their atomics).  Bring-up tool."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepfields_nerf_b200 as D  # noqa: E402

dev = torch.device("cuda:0")
g = D.ops.hashgrid_levels(19)
P = 1 << 22
torch.manual_seed(0)
pos = torch.rand(P, 3, device=dev)
d = (torch.rand(P, 32, device=dev) - 0.5).half()
dt = torch.zeros(2 * g.n_entries, device=dev)


def timed(x, reps=5):
    D.ops.hashgrid_backward(pos, x, g, d_table=dt)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        D.ops.hashgrid_backward(pos, x, g, d_table=dt)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


print(f"all 16 levels        : {timed(d):.3f} ms")
for lo, hi in ((0, 1), (0, 2), (0, 4), (4, 16), (0, 16)):
    x = d.clone()
    x[:, 2 * lo:2 * hi] = 0
    print(f"levels {lo:2d}..{hi - 1:2d} zeroed : {timed(x):.3f} ms")
