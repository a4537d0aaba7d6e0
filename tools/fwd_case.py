"""Run one forward of P points and compare with the bf16 oracle (debug helper). Usage: python tools/fwd_case.py P [train]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch
import deepfields_nerf_b200 as D
from common import make_state_dict
from oracle import nerf_oracle as O
P = int(sys.argv[1]); train = len(sys.argv) > 2
sd = make_state_dict(22, 256)
g = torch.Generator().manual_seed(P)
pos = torch.rand(P, 3, generator=g) * 2 - 1
view = torch.nn.functional.normalize(torch.randn(P, 3, generator=g), dim=-1)
pk = D.ops.PackedMLP("cuda"); pk.pack(torch.cat([v.reshape(-1) for k, v in sd.items() if "freq" not in k]).cuda())
stash = torch.empty(D.ops.mlp_stash_bytes(P), dtype=torch.uint8, device="cuda") if train else None
rgb, sig = D.ops.mlp_forward(pk, pos=pos.cuda(), view=view.cuda(), stash=stash)
torch.cuda.synchronize()
idx = torch.randperm(P, generator=g)[:2048]
r_o, s_o = O.mlp_forward(sd, pos[idx], view[idx], low=torch.bfloat16)
print("max rgb err", float((rgb.cpu()[idx] - r_o.float()).abs().max()), "max sigma err", float((sig.cpu()[idx] - s_o.float()[:, 0]).abs().max()))
