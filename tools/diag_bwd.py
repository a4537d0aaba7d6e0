"""Diagnostic: per-tensor gradient error of the CUDA backward vs the fp32 oracle and vs the bf16-autocast oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import deepfields_nerf_b200 as D
from common import make_state_dict, make_noise
from oracle import nerf_oracle as O
from test_gpu_train import leaf_sd, flat_params, split_flat, rel_err, param_keys

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
sd = make_state_dict(40 + P, 256)
g = torch.Generator().manual_seed(P)
pos = torch.rand(P, 3, generator=g) * 2 - 1
view = torch.nn.functional.normalize(torch.randn(P, 3, generator=g), dim=-1)
G_rgb = torch.randn(P, 3, generator=g) * 1e-3
G_sig = torch.randn(P, generator=g) * 1e-3
refs = {}
for name, low in (("fp32", None), ("bf16", torch.bfloat16)):
    ref = leaf_sd(sd)
    rgb_o, sig_o = O.mlp_forward(ref, pos, view, None, low)
    ((rgb_o.float() * G_rgb).sum() + (sig_o.float()[:, 0] * G_sig).sum()).backward()
    refs[name] = ref
pk = D.ops.PackedMLP("cuda"); pk.pack(flat_params(sd).cuda())
stash = torch.empty(D.ops.mlp_stash_bytes(P), dtype=torch.uint8, device="cuda")
dz = torch.empty(D.ops.mlp_dz_bytes(P), dtype=torch.uint8, device="cuda")
rgb, sig = D.ops.mlp_forward(pk, pos=pos.cuda(), view=view.cuda(), stash=stash)
grads = torch.zeros(D.ops.MLP_NUM_PARAMS, device="cuda")
D.ops.mlp_backward(pk, rgb, sig, G_rgb.cuda(), G_sig.cuda(), stash, dz, grads, pos=pos.cuda(), view=view.cuda())
got = split_flat(grads.cpu(), sd)
print(f"{'tensor':28s} {'cuda-vs-fp32':>12s} {'cuda-vs-bf16':>12s} {'bf16-vs-fp32':>12s}  |ref|")
for k in param_keys(sd):
    a, b = refs["fp32"][k].grad, refs["bf16"][k].grad
    print(f"{k:28s} {rel_err(got[k], a):12.4f} {rel_err(got[k], b):12.4f} {rel_err(b, a):12.4f}  {float(a.norm()):.3e}")
w = got["pos_layers.0.weight"]; a = refs["fp32"]["pos_layers.0.weight"].grad
colerr = ((w - a).norm(dim=0) / a.norm(dim=0).clamp(min=1e-30))
print("pos_layers.0.weight per-input-column rel err:", [round(float(x), 3) for x in colerr])
w = got["skip_layers.0.weight"]; a = refs["fp32"]["skip_layers.0.weight"].grad
colerr = ((w - a).norm(dim=0) / a.norm(dim=0).clamp(min=1e-30))
print("skip_layers.0.weight PE columns:", [round(float(x), 3) for x in colerr[256:]])
print("skip_layers.0.weight hidden cols (first 8):", [round(float(x), 3) for x in colerr[:8]])
