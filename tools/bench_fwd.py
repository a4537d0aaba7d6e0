"""Quick device-side timing of the fused MLP forward (grid source). Usage: python tools/bench_fwd.py [log2_points]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
import torch
import deepfields_nerf_b200 as d
from common import make_state_dict

lg = int(sys.argv[1]) if len(sys.argv) > 1 else 24
R = 512
P = 1 << lg
sd = make_state_dict(1, 256)
pk = d.ops.PackedMLP("cuda")
pk.pack(torch.cat([v.reshape(-1) for k, v in sd.items() if "freq" not in k]).cuda())
lin = torch.linspace(-1, 1, R, device="cuda")
view = torch.tensor([0.0, 0.0, -1.0], device="cuda")
out = (torch.empty(P, 3, device="cuda"), torch.empty(P, device="cuda"))
for _ in range(3):
    d.ops.mlp_forward(pk, grid=(lin, view, 0, P), out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 5
e0.record()
for _ in range(n):
    d.ops.mlp_forward(pk, grid=(lin, view, 0, P), out=out)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
flop = P * 1186816
print(f"mlp_forward P=2^{lg}: {ms:.3f} ms  {P/ms/1e6:.3f} Gpts/s  {flop/ms/1e9:.1f} TFLOP/s (algorithmic)")
