"""One launch each of the Instant-NGP kernels at the bench sizes, for ncu (profiles/r1_ngp_*).  Not a benchmark."""
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
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
pos, dirs = torch.rand(P, 3, device=dev), torch.rand(P, 3, device=dev)
table, dw, cw = net._half_params()
for _ in range(2):
    D.ops.ngp_forward(pos, dirs, table, net.grid, dw, cw)
    enc = D.ops.hashgrid_forward(pos, table, net.grid)
    D.ops.hashgrid_backward(pos, enc, net.grid)
torch.cuda.synchronize()
print("ok")
