"""The HBM-bound small kernels of the hot path at a SATURATING batch (2^20 rays): CUDA-event time, algorithmic bytes
(SURVEY 8d per-ray figures) and GB/s against the measured copy peak.  Run under
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none
for the DRAM traffic of the same launches (profiles/r2_small_kernels_*.csv).

    python tools/small_kernels_profile.py [log2_rays=20] [--once]      (--once: one launch per kernel, for ncu)
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import deepfields_nerf_b200 as D  # noqa: E402

LOG2 = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 20
ONCE = "--once" in sys.argv
B, S, I = 1 << LOG2, 64, 128
T = S + I
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
peak = 6545.0
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(pk):
    peak = json.load(open(pk))["hbm_gbs"]

near = torch.rand(B, device=dev, generator=g) + 2.0
far = near + 1.0 + torch.rand(B, device=dev, generator=g)
t_vals = torch.linspace(0, 1, S, device=dev)
t_rand = torch.rand(B, S, device=dev, generator=g)
u = torch.rand(B, I, device=dev, generator=g)
o = torch.randn(B, 3, device=dev, generator=g)
d = torch.nn.functional.normalize(torch.randn(B, 3, device=dev, generator=g), dim=-1)


def timed(fn, reps=5):
    if ONCE:
        fn()
        torch.cuda.synchronize()
        return float("nan")
    fn(); fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


rows = []


def report(name, ms, bytes_per_ray, note=""):
    gbs = B * bytes_per_ray / (ms * 1e-3) / 1e9 if ms == ms else float("nan")
    rows.append({"kernel": name, "rays": B, "ms": ms, "bytes_per_ray": bytes_per_ray, "gbs": gbs, "frac_of_hbm_peak": gbs / peak, "note": note})
    print(f"{name:34s} {ms:8.3f} ms  {bytes_per_ray:6d} B/ray  {gbs:8.0f} GB/s  {gbs / peak:5.2f} of {peak:.0f}  {note}", flush=True)


# stratified z (fused mode: z only): 8 B near/far + 256 B draws in, 256 B out
z = D.ops.stratified_sample(near, far, t_vals, t_rand)
report("stratified_kernel (z only)", timed(lambda: D.ops.stratified_sample(near, far, t_vals, t_rand)), 8 + 4 * S + 4 * S)
# points emission at the reference signature: + pts 768 + dirs 768 (SURVEY: 2080 B/ray with the z kernel)
report("points_kernel (pts + dirs)", timed(lambda: D.ops.points_on_rays(o, d, z)), 24 + 4 * S + 2 * 12 * S)
# sample_pdf fused (kernel builds the cdf): weights 256 + z 256 + u 512 in, samples 512 out
w = torch.rand(B, S, device=dev, generator=g) ** 4
report("sample_pdf_kernel (64 -> 128)", timed(lambda: D.ops.sample_pdf(z, w[:, 1:-1], u, bins_mid_of_z=True)), 4 * S + 4 * S + 4 * I + 4 * I)
z_new = D.ops.sample_pdf(z, w[:, 1:-1], u, bins_mid_of_z=True)
# merge sort: 768 in, 768 out
report("merge_sort_kernel (64 + 128)", timed(lambda: D.ops.merge_sort_z(z, z_new)), 4 * T + 4 * T)
z_all = D.ops.merge_sort_z(z, z_new)
# the two fused (train step / render): weights 256 + z 256 + u 512 in, z_all 768 out
report("sample_merge_kernel (64 -> 192)", timed(lambda: D.ops.sample_merge(z, w, u)), 4 * S + 4 * S + 4 * I + 4 * T)
del t_rand, u, w, z_new
# compositing forward, S = 192, fp32 inputs (what the train step feeds it): rgb 2304 + sigma 768 + z 768 in, weights 768 + 12 out
rgb = torch.rand(B, T, 3, device=dev, generator=g)
sig = torch.rand(B, T, device=dev, generator=g) * 3
report("composite_kernel fwd (S=192, fp32)", timed(lambda: D.ops.composite_forward(rgb, sig, z_all)), 12 * T + 4 * T + 4 * T + 4 * T + 12)
# fused compositing + MSE + backward: the same reads + target 12, writes rgb_map 12 + d_rgb 2304 + d_sigma 768 (no weights)
tgt = torch.rand(B, 3, device=dev, generator=g)
ls = torch.zeros(1, device=dev)
report("render_loss_kernel (S=192, fp32)", timed(lambda: D.ops.render_loss(rgb, sig, z_all, tgt, ls, want_weights=False)),
       12 * T + 4 * T + 4 * T + 12 + 12 + 12 * T + 4 * T)
# bf16 inputs at the boundary signature (SURVEY: 3084 B/ray)
rgb16, sig16 = rgb.bfloat16(), sig.bfloat16()
del rgb, sig
report("composite_kernel fwd (S=192, bf16)", timed(lambda: D.ops.composite_forward(rgb16, sig16, z_all)), 6 * T + 2 * T + 4 * T + 4 * T + 12)
del rgb16, sig16, z_all
# positional encoding at the boundary: 12 B in, 252 B out per POINT; 2^22 points
Pn = 1 << 22
x = torch.rand(Pn, 3, device=dev, generator=g) * 2 - 1
fb = (2.0 ** torch.arange(10, device=dev)) * torch.pi
ms = timed(lambda: D.ops.posenc(x, fb))
gbs = Pn * (12 + 252) / (ms * 1e-3) / 1e9 if ms == ms else float("nan")
rows.append({"kernel": "posenc_kernel (L=10)", "points": Pn, "ms": ms, "bytes_per_point": 264, "gbs": gbs, "frac_of_hbm_peak": gbs / peak})
print(f"{'posenc_kernel (L=10), 2^22 points':34s} {ms:8.3f} ms  {264:6d} B/pt   {gbs:8.0f} GB/s  {gbs / peak:5.2f} of {peak:.0f}", flush=True)
if not ONCE:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "small_kernels.json"), "w") as fh:
        json.dump({"hbm_peak_gbs": peak, "rows": rows}, fh, indent=1)
