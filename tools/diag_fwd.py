"""Forward kernel diagnostics: time train / no-train forward under the DFN_FWD_DEBUG switches.
Usage: python tools/diag_fwd.py [log2_points]"""
import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 2 and sys.argv[2] == "child":
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import torch
    import deepfields_nerf_b200 as d
    from common import make_state_dict
    lg = int(sys.argv[1]); train = sys.argv[3] == "1"
    P = 1 << lg; R = 512
    sd = make_state_dict(1, 256)
    pk = d.ops.PackedMLP("cuda")
    pk.pack(torch.cat([v.reshape(-1) for k, v in sd.items() if "freq" not in k]).cuda())
    lin = torch.linspace(-1, 1, R, device="cuda"); view = torch.tensor([0.0, 0.0, -1.0], device="cuda")
    out = (torch.empty(P, 3, device="cuda"), torch.empty(P, device="cuda"))
    stash = torch.empty(d.ops.mlp_stash_bytes(P), dtype=torch.uint8, device="cuda") if train else None
    for _ in range(2):
        d.ops.mlp_forward(pk, grid=(lin, view, 0, P), out=out, stash=stash)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    e0.record()
    for _ in range(n):
        d.ops.mlp_forward(pk, grid=(lin, view, 0, P), out=out, stash=stash)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"train={int(train)} debug={os.environ.get('DFN_FWD_DEBUG','0')}: {ms:.3f} ms {P*1186816/ms/1e9:.1f} TF/s", flush=True)
else:
    lg = sys.argv[1] if len(sys.argv) > 1 else "20"
    for train in ("0", "1"):
        for dbg in ("0",):
            env = dict(os.environ, DFN_FWD_DEBUG=dbg)
            subprocess.run([sys.executable, __file__, lg, "child", train], env=env)
    for train in ():
        for dbg in ("8", "24"):
            env = dict(os.environ, DFN_FWD_DEBUG=dbg, DFN_FWD_STATS="1")
            subprocess.run([sys.executable, __file__, lg, "child", train], env=env)
