#!/usr/bin/env python
"""Benchmark of the NeRF render/train hot path (BASELINE.json config 2: vanilla NeRF train step, 4096 rays per
GPU x (64 stratified + 128 hierarchical) samples, bf16 MLP, data-parallel rays).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One "step" = one full training iteration of scripts/train.py:282-323 on one 4096-ray batch per GPU: jittered
sampling, coarse MLP, compositing + loss, inverse-CDF resampling, merge, fine MLP, compositing + loss, backward of
both networks, gradient all-reduce (N > 1), AdamW + OneCycleLR, bf16 weight repack.  Synthetic rays, random-init
weights.  Rank 0 prints ONE JSON line.

`value`   rays/s over all ranks with the ray batches already resident in HBM.
`e2e`     the same step driven from pinned HOST buffers through FusedTrainer.step_from_host: H2D copy of the
          batch and D2H read of the losses inside the timed region.
`roofline` the dominant kernel (tensor-bound MLP kernel) timed live with CUDA events inside the timed region.
`cpu_baseline` the oracle port of the reference path timed on this box's host cores (bounded sample).
`--impl reference` times that CPU port alone (the reference is pure Python on torch; /root/reference does not
exist on the GPU box, so the port in oracle/ stands in for it -- kind "port").
"""
from __future__ import annotations

import argparse
import datetime
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)

B_RAYS, S_COARSE, N_IMPORTANCE = 4096, 64, 128
T_FINE = S_COARSE + N_IMPORTANCE
# algorithmic FLOP per MLP evaluation (unpadded layer dims, SURVEY 8d / DESIGN.md)
FLOP_FWD = 2 * 593408
FLOP_DGRAD = 2 * 557696
FLOP_WGRAD = 2 * 593408
FLOP_PER_RAY_TRAIN = (S_COARSE + T_FINE) * (FLOP_FWD + FLOP_DGRAD + FLOP_WGRAD)
STAGE_FLOP = {"mlp_fwd": FLOP_FWD, "dgrad": FLOP_DGRAD, "wgrad": FLOP_WGRAD}
# algorithmic HBM bytes per MLP evaluation: stash = h1..h8 + features + colour hidden (bf16) + 8 bit
# masks; dz = dZ1..dZ8 + dZ_feat + dZ_c0 + 16 B of head gradients.  fwd writes the stash (+16 B outputs); dgrad reads the masks,
# the colour tile and 32 B of outputs/gradients and writes dz; wgrad reads stash tiles and dz once each.
STASH_B, MASK_B, DZ_B = 4864, 256, 4880
STAGE_BYTES = {"mlp_fwd": STASH_B + MASK_B + 16, "dgrad": MASK_B + 256 + 32 + DZ_B, "wgrad": STASH_B + DZ_B}
STAGE_BOUND = {"mlp_fwd": "tensor", "dgrad": "tensor", "wgrad": "hbm"}   # what limits the kernel by design
METRIC, UNIT = "train_rays_per_s", "rays/s"
WORKLOAD = ("vanilla NeRF train step (model.py, PE L=10/4, hidden 256), 4096 rays/GPU x (64 stratified + 128 "
            "hierarchical) samples, coarse+fine nets, bf16 MLP, fp32 sampling/compositing/AdamW")


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        return {"hbm": d["hbm_gbs"], "tensor_burst": d["bf16_tflops"], "tensor_sustained": d["bf16_tflops_sustained"],
                "src": "measured (MEASURED_PEAKS.json)"}
    return {"hbm": 6650.0, "tensor_burst": 1590.0, "tensor_sustained": 1400.0, "src": "fallback (B200_PROFILING.md)"}


class AuxDeadline:
    """Bounds everything after the headline measurement.  The auxiliary sections (multi-GPU consistency checks, extract_vol,
    render, NGP, comparators) run AFTER the contract's numbers are in `line`; if one of them does not finish within `seconds`
    (a collective that never completes, a slow host leg) rank 0 prints the line as it stands, naming the section it was in,
    and every rank exits 0 - the headline must not be lost to an auxiliary measurement."""

    def __init__(self, seconds: float, rank: int, line: dict):
        self.seconds, self.rank, self.line = seconds, rank, line
        self.section = "start"
        self.lock = threading.Lock()
        self.printed = False
        self.timer = threading.Timer(seconds, self._fire)
        self.timer.daemon = True
        self.timer.start()

    def enter(self, section: str):
        self.section = section

    def _fire(self):
        with self.lock:
            if self.printed:
                return
            self.printed = True
            if self.rank == 0:
                self.line["aux_timeout"] = {"section": self.section, "seconds": self.seconds}
                try:
                    text = json.dumps(with_summary(self.line))
                except Exception:          # the main thread was mutating a section's dict: drop the auxiliary sections
                    text = json.dumps({k: v for k, v in self.line.items() if k in HEADLINE_KEYS or k == "aux_timeout"})
                print(text, flush=True)
            sys.stdout.flush()
            os._exit(0)

    def finish(self) -> bool:
        """True if the caller is the one that prints."""
        with self.lock:
            if self.printed:
                return False
            self.printed = True
        self.timer.cancel()
        return True


def with_summary(line: dict) -> dict:
    """Puts a compact `summary` object LAST in the line: the second half of BASELINE.json's metric (extract_vol points/s), the
    inference render, strong scaling and the N > 1 equality flags in < 1 200 characters, so that they survive a reader that
    keeps only the tail of the line.  Values are copied from the sections above it; absent sections give null."""
    def g(d, *ks):
        for k in ks:
            d = d.get(k) if isinstance(d, dict) else None
        return d

    def r(x, n=4):
        return round(x, n) if isinstance(x, float) else x

    ex, mg, rn, ng = line.get("extract_vol"), line.get("multi_gpu_checks"), line.get("render"), line.get("ngp")
    pmm = g(mg, "param_max_minus_min_over_ranks_after_steps")
    summ = {
        "n_gpus": line.get("n_gpus"),
        "extract_vol_gpts": r(g(ex, "value") / 1e9) if g(ex, "value") else None,
        "extract_vol_ms": r(g(ex, "ms"), 2),
        "extract_vol_frac_tensor": r(g(ex, "frac_of_sustained_tensor_peak")),
        "extract_includes_allgather": bool(ex and line.get("n_gpus", 1) > 1) if ex else None,
        "extract_cpu_pts_per_s": r(g(ex, "cpu_baseline", "value"), 1),
        "render_full_ms": r(g(rn, "ms_full"), 2),
        "render_ert_ms": r(g(rn, "ms_ert"), 2),
        "strong_ms_per_step": r(g(mg, "strong_scaling", "ms_per_step")),
        "strong_rays_per_s": r(g(mg, "strong_scaling", "rays_per_s"), 0),
        "strong_rays_per_gpu": g(mg, "strong_scaling", "rays_per_gpu"),
        "params_identical": (pmm == 0.0) if pmm is not None else None,
        "extract_sharded_eq_single": g(mg, "extract_64_sharded_equals_single_gpu"),
        "render_sharded_eq_single": g(mg, "render_64x48_sharded_equals_single_gpu"),
        "threshold_identical": g(mg, "extract_threshold_identical_on_all_ranks"),
        "ngp_train_ms": r(g(ng, "train_step", "ms"), 2),
        "ngp_train_rays_per_s": r(g(ng, "train_step", "rays_per_s"), 0),
        "ngp_dp_ms": r(g(ng, "data_parallel", "ms"), 2),
        "ngp_dp_rays_per_s": r(g(ng, "data_parallel", "rays_per_s"), 0),
        "ngp_dp_allreduce_mb": r(g(ng, "data_parallel", "allreduce_mb"), 1),
        "aux_timeout": g(line, "aux_timeout", "section"),
    }
    line.pop("summary", None)
    line["summary"] = summ
    return line


HEADLINE_KEYS = ("metric", "value", "unit", "n_gpus", "steps", "warmup", "repeats", "timed_region_s", "ms_per_step",
                 "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config", "roofline", "e2e", "gpu_launches",
                 "clocks", "step_tflops", "step_frac_of_sustained_tensor_peak", "stages")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            self.thr.join(timeout=2)

    def summary(self):
        sm, mx, reasons, pw = [], [], set(), []
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# ------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's train step (fp32, as scripts/train_no_amp.py computes on CPU)
# ------------------------------------------------------------------------------------------------------------
def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm is meant to use the host cores the box has."""
    import torch
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        pass
    torch.set_num_threads(max(n, 1))
    return torch.get_num_threads()


class CpuStep:
    """One training step of the CPU port: render_losses + backward + AdamW (scripts/train.py:282-323).
    autocast=False: fp32, what scripts/train_no_amp.py computes; autocast=True: under torch.autocast("cpu", bfloat16),
    what scripts/train.py:120,262 does when no GPU is present."""

    def __init__(self, n_rays: int, autocast: bool = False):
        import torch
        from common import make_state_dict
        from oracle import nerf_oracle as O
        self.torch, self.O, self.n, self.autocast = torch, O, n_rays, autocast
        self.sd = [make_state_dict(31, 256), make_state_dict(32, 256)]
        self.leaves = []
        for sd in self.sd:
            for k in sd:
                if "freq_bands" not in k:
                    sd[k] = sd[k].clone().requires_grad_(True)
                    self.leaves.append(sd[k])
        self.opt = torch.optim.AdamW(self.leaves, lr=5e-4, weight_decay=1e-6)  # scripts/train.py:244-248

    def __call__(self, seed: int, n_rays=None):
        from common import make_rays
        torch, O = self.torch, self.O
        n = n_rays or self.n
        o, d, near, far, tgt = make_rays(seed, n)
        t_rand, u = torch.rand(n, S_COARSE), torch.rand(n, N_IMPORTANCE)
        nc, nf = torch.randn(n * S_COARSE, 1), torch.randn(n * T_FINE, 1)
        t0 = time.perf_counter()
        self.opt.zero_grad(set_to_none=True)
        with torch.autocast("cpu", dtype=torch.bfloat16, enabled=self.autocast):
            out = O.render_losses(self.sd[0], self.sd[1], o, d, near, far, tgt, t_rand, u, nc, nf)
        out["loss"].backward()
        self.opt.step()
        return time.perf_counter() - t0


class RefStep:
    """One training step of the REFERENCE ITSELF on the host cores: the unmodified scripts/train.py `forward_pass` (staged under
    oracle/_ref by oracle/make_ref.py) with the reference's models, sampler and MSELoss, under torch.autocast("cpu") as
    train.py:294-295 does on a machine without CUDA, then backward and AdamW (:244-248, :310-321).  The GradScaler of :262 is
    left out: on CPU it is disabled by torch and only forwards the calls.  Rays are processed in slices of `n` so that a step
    of 4096 rays can be sampled; weights = the bench's synthetic state dicts."""

    def __init__(self):
        import torch
        from types import SimpleNamespace
        from common import make_state_dict
        from oracle import make_ref
        self.torch = torch
        self.ref_train, ref_model, ref_sampling = make_ref.load()
        self.ref_train.cfg = SimpleNamespace(N_samples=S_COARSE, N_importance=N_IMPORTANCE, chunk_size=32768)   # cfg/base_cfg.yml:64-67
        self.dev = torch.device("cpu")
        self.mc, self.mf = ref_model.get_nerf_models(10, 4, 256, device=self.dev)
        self.mc.load_state_dict(make_state_dict(31, 256))
        self.mf.load_state_dict(make_state_dict(32, 256))
        self.mc.train(); self.mf.train()
        self.sampler = ref_sampling.StratifiedSampler(S_COARSE, self.dev)
        self.crit = torch.nn.MSELoss(reduction="mean")
        self.opt = torch.optim.AdamW(list(self.mc.parameters()) + list(self.mf.parameters()), lr=5e-4, weight_decay=1e-6)

    def __call__(self, seed: int, n_rays: int):
        from common import make_rays
        torch = self.torch
        o, d, near, far, tgt = make_rays(seed, n_rays)
        t0 = time.perf_counter()
        self.opt.zero_grad()
        with torch.amp.autocast("cpu"):
            total, _, _ = self.ref_train.forward_pass(self.mc, self.mf, o, d, near, far, tgt, self.sampler, self.crit, self.dev, False)
        total.backward()
        self.opt.step()
        return time.perf_counter() - t0


def cpu_rate(budget_s: float):
    """rays/s of the CPU port on a bounded sample (>= 1 warm-up, median of >= 5 repetitions), fp32 and
    bf16 CPU-autocast; the fp32 figure is the headline `value` (it is the faster of the two on hosts without bf16 GEMMs)."""
    cores = use_all_host_threads()
    out = {}
    for tag, ac in (("fp32", False), ("bf16_autocast", True)):
        step = CpuStep(64, ac)
        step(0, 64)                      # warm-up (first-call allocator / thread-pool effects)
        t = step(1, 64)
        # 1 warm-up + 5 timed repetitions of n rays inside budget_s / 2 per variant
        n = int(max(64, min(1024, (budget_s / 2 / 6 / max(t / 64, 1e-9)) // 64 * 64)))
        step(2, n)
        ts = sorted(step(3 + i, n) for i in range(5))
        out[tag] = {"rays_per_s": n / ts[2], "rays": n, "median_s": ts[2], "min_s": ts[0], "max_s": ts[-1]}
    best = max(out, key=lambda k: out[k]["rays_per_s"])
    sample = (f"median of 5 steps of {out[best]['rays']} rays x (64+192) samples after a warm-up step, torch CPU, {cores} threads; "
              f"value = the faster variant ({best}); fp32 = train_no_amp.py semantics, bf16_autocast = train.py:120,262 on CPU")
    return out[best]["rays_per_s"], sample, cores, out


def cpu_extract_rate(R: int = 64):
    """points/s of the CPU port of extract_vol (the dense query on a 64^3 sub-grid, per-point rate comparable with
    the 512^3 figure since the work per point does not depend on R): grid, eval-mode query, Otsu, u8 pack; fp32, all host threads,
    one warm-up batch, median of 3."""
    import torch
    from common import make_state_dict
    from oracle import nerf_oracle as O
    cores = use_all_host_threads()
    sd = make_state_dict(32, 256)
    view = torch.tensor([0.0, 0.0, 1.0])
    with torch.no_grad():
        O.dense_query(sd, O.grid_points((-1.0, 1.0), 16), view)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            pts = O.grid_points((-1.0, 1.0), R)
            sig, rgb = O.dense_query(sd, pts, view)
            thr = O.otsu_threshold(sig)
            O.pack_volume(sig, rgb, thr, R)
            ts.append(time.perf_counter() - t0)
    ts.sort()
    return {"value": R ** 3 / ts[1], "unit": "points/s", "cores": cores, "kind": "port",
            "sample": f"median of 3 passes over a {R}^3 grid (grid + query + Otsu + u8 pack), oracle port, torch CPU fp32"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = use_all_host_threads()
    total = max(args.steps + args.warmup, 1)
    kind, why_port = "port", None
    try:
        from oracle import make_ref
        if not make_ref.available():
            raise RuntimeError("oracle/_ref not staged (python oracle/make_ref.py where /root/reference exists)")
        step = RefStep()
        kind, variant = "reference", "bf16_autocast"
        step(0, 64)
        t64 = step(1, 64)
    except Exception as e:   # the staged reference is missing or does not import here: fall back to the pinned port
        why_port = f"{type(e).__name__}: {e}"[:200]
        # the port's CPU path is train.py under torch.autocast("cpu") (bf16) or train_no_amp.py (fp32): time the faster
        probes = {}
        for tag, ac in (("fp32", False), ("bf16_autocast", True)):
            pr = CpuStep(64, ac)
            pr(0, 64)
            probes[tag] = (pr(1, 64), pr)
        variant = min(probes, key=lambda k: probes[k][0])
        t64, step = probes[variant]
    # the configuration's 4096-ray step when steps+warmup of it fit ~150 s, else the largest slice of it that does
    n = int(max(32, min(B_RAYS, (150.0 / total) / max(t64 / 64, 1e-9) // 32 * 32)))
    for i in range(args.warmup):
        step(10 + i, n)
    times = [step(100 + i, n) for i in range(args.steps)]
    elapsed = sum(times)
    value = n * args.steps / elapsed
    what = ("the reference's own scripts/train.py forward_pass + backward + AdamW under torch.autocast('cpu') (oracle/_ref)"
            if kind == "reference" else f"CPU port oracle/nerf_oracle.py, {variant} (the faster of fp32 / bf16 autocast on this host)")
    sample = (f"{args.steps} steps of {n} rays x (64+192) samples ("
              + ("the whole 4096-ray batch" if n == B_RAYS else "bounded sample of the 4096-ray batch") + f"), {what}, {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * elapsed / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16" if variant != "fp32" else "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "rays_per_step": n, "variant": variant, "same_config": n == B_RAYS,
                   "note": what + ("" if why_port is None else "; staged reference unavailable: " + why_port)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from common import make_rays, make_state_dict
    import deepfields_nerf_b200 as D

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    pk = peaks()

    K, W = args.steps, max(args.warmup, 3)
    mc, mf = D.get_nerf_models(device=dev)
    mc.load_state_dict(make_state_dict(31, 256))
    mf.load_state_dict(make_state_dict(32, 256))
    tr = D.FusedTrainer(mc, mf, S_COARSE, N_IMPORTANCE, lr=5e-4, total_steps=10 * (K + W) + 64, train_noise=True)

    n_batches = 8
    host, resident = [], []   # pinned flat batches (the .npz columns of dataset.py:32-60) and their HBM copies
    for i in range(n_batches):
        o, d, near, far, rgb = make_rays(1000 + 17 * rank + i, B_RAYS)
        host.append(D.FusedTrainer.pack_host_batch(o, d, near, far, rgb))
        resident.append(tuple(t.reshape(-1).to(dev) if t.shape[-1] == 1 else t.to(dev) for t in (o, d, near, far, rgb)))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident timed region -------------------------------------------------------------------
    # The K steps the driver asks for are repeated REP times back to back inside ONE timed region so that it lasts
    # >= --min-seconds: a 0.1 s burst runs at boost clocks, the sustained bf16 peak it is compared with does not.
    w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(W):
        if i == W - 1:
            w0.record()
        tr.step(*resident[i % n_batches])
    w1.record()
    barrier()
    est_ms = max(reduce_max(w0.elapsed_time(w1)), 1e-3)
    REP = max(1, int(math.ceil(args.min_seconds * 1e3 / (est_ms * K))))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tr.profile = True
    with ClockSampler(local) as clk:
        barrier()
        e0.record()
        for i in range(K * REP):
            tr.step(*resident[(W + i) % n_batches])
        e1.record()
        barrier()
    tr.profile = False
    ms_total = reduce_max(e0.elapsed_time(e1)) / REP          # per K steps
    stages = tr.stage_times_ms()
    launches = tr.launches_per_step * K
    value = world * B_RAYS * K / (ms_total * 1e-3)

    # ---- end-to-end timed region: pinned host batches in, losses out --------------------------------------
    for i in range(2):
        tr.step_from_host(host[i % n_batches])
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for i in range(K * REP):
        losses_host = tr.step_from_host(host[(2 + i) % n_batches])
    e3.record()
    barrier()
    ms_e2e = reduce_max(e2.elapsed_time(e3)) / REP
    e2e_value = world * B_RAYS * K / (ms_e2e * 1e-3)
    assert all(map(lambda v: v == v, losses_host.tolist())), "non-finite loss"

    # ---- roofline of the dominant kernel ------------------------------------------------------------------
    pts = {"coarse": B_RAYS * S_COARSE, "fine": B_RAYS * T_FINE}
    mean = {k: sum(v) / len(v) for k, v in stages.items()}
    tot = sum(sum(v) for v in stages.values())
    mlp = {k: v for k, v in mean.items() if k.split("_")[-1] in pts and k.rsplit("_", 1)[0] in STAGE_FLOP}
    dom = max(mlp, key=lambda k: mlp[k] * len(stages[k]))
    kind, net = dom.rsplit("_", 1)
    flop = STAGE_FLOP[kind] * pts[net]
    nbytes = STAGE_BYTES[kind] * pts[net]
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as fh:
            traffic = json.load(fh).get(dom)
    # SURVEY 8d counts this kernel's work in FLOP on the tensor pipe; that fraction comes first.  The HBM fraction of the
    # same launch (the design's own stash / dz bytes) is reported beside it: it says whether the kernel could go faster
    # WITH THESE BYTES, not whether the bytes are necessary.
    tf_ach = flop / (mlp[dom] * 1e-3) / 1e12
    hbm_ach = nbytes / (mlp[dom] * 1e-3) / 1e9
    roofline = {"bound": "tensor", "kernel": dom, "achieved": tf_ach, "peak": pk["tensor_sustained"], "unit": "TFLOP/s",
                "frac": tf_ach / pk["tensor_sustained"], "frac_of_burst_peak": tf_ach / pk["tensor_burst"], "traffic": traffic,
                "peak_source": pk["src"] + ", sustained bf16 (timed region %.1f s)" % (ms_total * REP * 1e-3),
                "flop_per_launch": flop, "ms_per_launch": mlp[dom],
                "hbm": {"achieved": hbm_ach, "peak": pk["hbm"], "unit": "GB/s", "frac": hbm_ach / pk["hbm"],
                        "bytes_per_launch": nbytes, "design_bound": STAGE_BOUND[kind]}}
    step_tflops = world * FLOP_PER_RAY_TRAIN * B_RAYS * K / (ms_total * 1e-3) / 1e12
    stage_report = {k: {"ms": round(mean[k], 4), "share": round(sum(stages[k]) / tot, 4)} for k in stages}
    for k in mlp:
        kk, nn = k.rsplit("_", 1)
        stage_report[k]["tflops"] = round(STAGE_FLOP[kk] * pts[nn] / (mean[k] * 1e-3) / 1e12, 1)
        stage_report[k]["hbm_gbs"] = round(STAGE_BYTES[kk] * pts[nn] / (mean[k] * 1e-3) / 1e9, 0)
        stage_report[k]["bound"] = STAGE_BOUND[kk]

    # ---- the contract's line is complete here; what follows only adds auxiliary sections to it, under a deadline ----------
    stash_gb = (D.ops.mlp_stash_bytes(pts["coarse"]) + D.ops.mlp_stash_bytes(pts["fine"]) + D.ops.mlp_dz_bytes(pts["fine"])) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "repeats": REP, "timed_region_s": ms_total * REP * 1e-3,
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "rays_per_gpu": B_RAYS, "global_rays": B_RAYS * world, "N_samples": S_COARSE,
                   "N_importance": N_IMPORTANCE, "parallelism": f"dp{world} (rays sharded, one NCCL all-reduce of the flat gradient)",
                   "l2": f"no explicit flush: each step streams {stash_gb:.2f} GB of activation/gradient scratch through HBM "
                         "(>> 126 MB L2) and rotates over 8 distinct ray batches"},
        "roofline": roofline,
        "cpu_baseline": None,
        "gpu_eager_baseline": None,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": host[0].numel() * 4 * world,
                "d2h_bytes_per_step": 8 * world, "ms_per_step": ms_e2e / K},
        "gpu_launches": launches,
        "clocks": clk.summary(),
        "step_tflops": step_tflops, "step_frac_of_sustained_tensor_peak": step_tflops / (pk["tensor_sustained"] * world),
        "stages": stage_report,
        "multi_gpu_checks": None, "extract_vol": None, "precompute_rays": None, "volume_points": None, "render": None,
        "ngp": None,
    }
    aux_s = args.aux_timeout if args.aux_timeout > 0 else (420.0 if world == 1 else 150.0)
    guard = AuxDeadline(aux_s, rank, line)

    # ---- N > 1: on-hardware consistency of the data-parallel state + strong scaling of the same 4096 global rays --------
    multi = None
    guard.enter("multi_gpu_checks: parameters + strong scaling")
    if world > 1:
        pmax, pmin = tr.flat.clone(), tr.flat.clone()
        dist.all_reduce(pmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(pmin, op=dist.ReduceOp.MIN)
        multi = {"param_max_minus_min_over_ranks_after_steps": float((pmax - pmin).abs().max()),
                 "steps_checked": int(tr.step_dev.item())}
        line["multi_gpu_checks"] = multi
        del pmax, pmin
        # strong scaling: ONE 4096-ray global batch, rank r owns rays [r*B/G, (r+1)*B/G)
        from deepfields_nerf_b200.train import shard_slice
        sl = shard_slice(B_RAYS, rank, world)
        strong = [tuple(t[sl].contiguous() for t in b_) for b_ in resident]
        for i in range(W):
            tr.step(*strong[i % n_batches])
        barrier()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        Rs = max(1, int(math.ceil(1000.0 / (est_ms * K))))            # ~1 s at the weak-scaling step time, shorter here
        s0.record()
        for i in range(K * Rs):
            tr.step(*strong[(W + i) % n_batches])
        s1.record()
        barrier()
        ms_s = reduce_max(s0.elapsed_time(s1)) / Rs
        multi["strong_scaling"] = {"global_rays": B_RAYS, "rays_per_gpu": sl.stop - sl.start, "ms_per_step": ms_s / K,
                                   "rays_per_s": B_RAYS * K / (ms_s * 1e-3), "steps": K, "repeats": Rs}
        del strong

    # ---- extract_vol 512^3 (BASELINE.json config 4), slab-sharded, reported beside the headline ---------------
    # N > 1: min/max + 256-bin histogram all-reduces (ONE global Otsu threshold) and the all-gather of the u8 slabs into the volume
    # are inside the timed region.
    extract = None
    guard.enter("extract_vol")
    if not args.no_extract:
        R = args.extract_res
        mf.eval()
        group = dist.group.WORLD if world > 1 else None
        x0s, x1s = D.slab_range(R, rank, world)
        n_pts = (x1s - x0s) * R * R
        out = (torch.empty(n_pts, 3, device=dev), torch.empty(n_pts, device=dev))

        def run_extract():
            sig_x, rgb_x, _, thr_x, occ_x, col_x = D.extract_volume_fused(mf, R, rank=rank, world=world, group=group, out=out)
            occ_g, col_g = D.gather_slabs(occ_x, col_x, R, rank, world, group)
            return thr_x, occ_g, col_g

        with torch.no_grad():
            run_extract()
            barrier()
            x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            x0.record()
            thr_x, occ_g, col_g = run_extract()
            x1.record()
            barrier()
        ms_x = reduce_max(x0.elapsed_time(x1))
        tf = R ** 3 * FLOP_FWD / (ms_x * 1e-3) / 1e12
        extract = {"metric": "extract_vol_points_per_s", "value": R ** 3 / (ms_x * 1e-3), "unit": "points/s", "resolution": R,
                   "ms": ms_x, "threshold": thr_x,
                   "includes": "MLP query + min/max + histogram + Otsu + u8 pack" + (
                       " + min/max and histogram all-reduces + all-gather of the u8 slabs into one volume (%d MB)" % (4 * R ** 3 // 2 ** 20)
                       if world > 1 else " (outputs stay in HBM)"),
                   "tflops": tf, "frac_of_sustained_tensor_peak": tf / (pk["tensor_sustained"] * world)}
        if rank == 0:
            extract["gathered_voxels"] = int(occ_g.numel())
        line["extract_vol"] = extract
        if rank == 0 and world == 1 and not args.no_cpu:
            try:
                extract["cpu_baseline"] = cpu_extract_rate(64)
            except Exception as e:  # an auxiliary measurement must not take the headline line down with it
                extract["cpu_baseline"] = {"error": f"{type(e).__name__}: {e}"[:300]}
        del out, occ_g, col_g
        if world > 1:
            # the sharded volume IS the single-GPU volume: same threshold on every rank, and on a 64^3 grid every rank's
            # slab equals the same range of the volume one GPU computes alone (bit for bit), threshold included
            thr_mm = torch.tensor([float(thr_x), -float(thr_x)], dtype=torch.float64, device=dev)
            dist.all_reduce(thr_mm, op=dist.ReduceOp.MAX)          # max and -min over the ranks
            Rc = 64
            with torch.no_grad():
                s_sig, s_rgb, s_start, s_thr, s_occ, s_col = D.extract_volume_fused(mf, Rc, rank=rank, world=world, group=group)
                f_sig, f_rgb, _, f_thr, f_occ, f_col = D.extract_volume_fused(mf, Rc, rank=0, world=1, group=False)
            sl_c = slice(s_start, s_start + s_occ.numel())
            same = bool(s_thr == f_thr and torch.equal(s_occ, f_occ[sl_c]) and torch.equal(s_col, f_col[sl_c])
                        and torch.equal(s_sig, f_sig[sl_c]))
            flag = torch.tensor([1.0 if same else 0.0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            multi["extract_threshold_identical_on_all_ranks"] = bool(float(thr_mm[0]) == -float(thr_mm[1]))
            multi["extract_64_sharded_equals_single_gpu"] = bool(flag.item() == 1.0)
            multi["extract_64_threshold"] = s_thr

    # ---- ray generation + cube intersection + compaction (scripts/precompute_rays.py), the step in front of the path ----
    raygen = None
    guard.enter("precompute_rays")
    if rank == 0 and not args.no_extract:
        import numpy as np
        sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
        from common import make_cameras
        Mc, Hc, Wc = 8, 1080, 1920
        _, _, Kc, poses_c, _ = make_cameras(21, M=Mc, H=2, W=2, focal=0.45 * Wc)
        Kc[0, 2], Kc[1, 2] = Wc / 2.0, Hc / 2.0
        img = torch.randint(0, 256, (Mc, Hc, Wc, 3), device=dev, dtype=torch.uint8)
        res = D.precompute_rays_on_device(Hc, Wc, Kc, poses_c, img)
        torch.cuda.synchronize()
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0.record()
        for _ in range(3):
            res = D.precompute_rays_on_device(Hc, Wc, Kc, poses_c, img)
        r1.record()
        torch.cuda.synchronize()
        ms_r = r0.elapsed_time(r1) / 3
        n_all, n_kept = Mc * Hc * Wc, res["rays_o"].shape[0]
        algo_bytes = 3 * n_all + 56 * n_kept     # u8 image in; o, d, rgb (12 B each) + t_near, t_far out per kept ray
        raygen = {"metric": "precompute_rays_per_s", "value": n_all / (ms_r * 1e-3), "unit": "rays/s", "cameras": Mc, "H": Hc, "W": Wc,
                  "kept_fraction": n_kept / n_all, "ms": ms_r, "hbm_gbs": algo_bytes / (ms_r * 1e-3) / 1e9,
                  "frac_of_hbm_peak": algo_bytes / (ms_r * 1e-3) / 1e9 / pk["hbm"],
                  "includes": "generate + intersect + ordered compaction + 8-byte count read-back, through the public API"}
        if not args.no_cpu:
            from oracle import nerf_oracle as O
            t0 = time.perf_counter()
            O.precompute_rays(Hc, Wc, Kc, poses_c[:1], np.zeros((1, Hc, Wc, 3), dtype=np.uint8))
            raygen["cpu_port_rays_per_s"] = Hc * Wc / (time.perf_counter() - t0)
            raygen["cpu_sample"] = "1 camera of 1080x1920, numpy oracle port, 1 thread"
        line["precompute_rays"] = raygen
        del img, res

    # ---- occupied voxels -> coloured point set (scripts/post_process_vol.py), the step after extract_vol ----
    volpts = None
    guard.enter("volume_points")
    if rank == 0 and not args.no_extract:
        Rv = args.extract_res
        gen = torch.Generator(device=dev).manual_seed(5)
        occ = (torch.rand(Rv, Rv, Rv, device=dev, generator=gen) < 0.1).to(torch.uint8)
        rgbv = torch.randint(0, 256, (Rv, Rv, Rv, 3), device=dev, dtype=torch.uint8, generator=gen)
        meta = {"aabb_min": [-0.8, -0.8, -0.8], "aabb_max": [0.8, 0.8, 0.8]}
        pv, cv = D.volume_points(occ, rgbv, meta)
        torch.cuda.synchronize()
        v0, v1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        v0.record()
        for _ in range(3):
            pv, cv = D.volume_points(occ, rgbv, meta)
        v1.record()
        torch.cuda.synchronize()
        ms_v = v0.elapsed_time(v1) / 3
        algo = 4 * Rv ** 3 + (36 + 3) * pv.shape[0]   # occupancy + colours in (max), colours again + 24 B xyz + 12 B rgb per kept voxel
        volpts = {"metric": "volume_points_voxels_per_s", "value": Rv ** 3 / (ms_v * 1e-3), "unit": "voxels/s", "resolution": Rv,
                  "kept_fraction": pv.shape[0] / Rv ** 3, "ms": ms_v, "hbm_gbs": algo / (ms_v * 1e-3) / 1e9,
                  "frac_of_hbm_peak": algo / (ms_v * 1e-3) / 1e9 / pk["hbm"],
                  "includes": "argwhere + colour max + gather + normalise + AABB slice + 8-byte count read-back, through the public API"}
        line["volume_points"] = volpts
        del occ, rgbv, pv, cv

    # ---- inference render of a 1920x1080 image, 256 samples/ray, with and without early ray termination (config 5) ----
    render = None
    guard.enter("render")
    if not args.no_extract:
        from common import make_cameras, peaked
        Hr, Wr = 1080, 1920
        _, _, Kr, poses_r, _ = make_cameras(33, M=1, H=2, W=2, focal=0.9 * Wr)
        Kr[0, 2], Kr[1, 2] = Wr / 2.0, Hr / 2.0
        sd_r = peaked(make_state_dict(32, 256))
        sd_r["density_layer.bias"] = sd_r["density_layer.bias"] + 60.0    # an opaque medium, so that rays do terminate
        mr_c, mr_f = D.get_nerf_models(device=dev)
        mr_c.load_state_dict(peaked(make_state_dict(31, 256)))
        mr_f.load_state_dict(sd_r)
        rows = D.row_range(Hr, rank, world)
        kw = dict(rows=rows, N_samples=64, N_importance=192)
        res_r = {}
        for tag, eps in (("full", 0.0), ("ert", 1e-4)):   # ERT epsilon = 1e-4
            D.render_image(mr_c, mr_f, Hr, Wr, Kr, poses_r[0], ert_eps=eps, ert_segments=8, **kw)
            barrier()
            q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            q0.record()
            part_r, info_r = D.render_image(mr_c, mr_f, Hr, Wr, Kr, poses_r[0], ert_eps=eps, ert_segments=8, **kw)
            img_r = D.gather_rows(part_r, Hr, rank, world, dist.group.WORLD if world > 1 else None)   # tiles -> rank 0
            q1.record()
            barrier()
            res_r[tag] = (reduce_max(q0.elapsed_time(q1)), info_r)
        n_rays = res_r["full"][1]["rays"]
        render = {"metric": "render_pixels_per_s", "image": [Hr, Wr], "samples_per_ray": 256, "rows_per_rank": rows[1] - rows[0],
                  "rays_hitting_cube_rank0": n_rays, "ms_full": res_r["full"][0], "ms_ert": res_r["ert"][0],
                  "ert_eps": 1e-4, "ert_segments": 8, "ert_evaluated_fraction_rank0": res_r["ert"][1]["evaluated_fraction"],
                  "value": Hr * Wr / (res_r["ert"][0] * 1e-3), "unit": "pixels/s",
                  "includes": "ray generation, cube test, coarse + fine MLP, resampling, compositing" + (
                      ", all-gather of the row tiles into one image" if world > 1 else "") + "; synthetic opaque medium"}
        line["render"] = render
        if world > 1:
            # row-sharded render == unsharded render (per-pixel draws fixed), 64 x 48 image
            Hs, Ws = 48, 64
            _, _, Ks, poses_s, _ = make_cameras(34, M=1, H=2, W=2, focal=0.9 * Ws)
            Ks[0, 2], Ks[1, 2] = Ws / 2.0, Hs / 2.0
            gd = torch.Generator(device=dev).manual_seed(11)
            draws = (torch.rand(Hs * Ws, 64, device=dev, generator=gd), torch.rand(Hs * Ws, 128, device=dev, generator=gd))
            kw_s = dict(N_samples=64, N_importance=128, draws=draws)
            part_s, _ = D.render_image(mr_c, mr_f, Hs, Ws, Ks, poses_s[0], rows=D.row_range(Hs, rank, world), **kw_s)
            img_s = D.gather_rows(part_s, Hs, rank, world, dist.group.WORLD)
            full_s, _ = D.render_image(mr_c, mr_f, Hs, Ws, Ks, poses_s[0], **kw_s)
            if rank == 0:
                multi["render_64x48_sharded_equals_single_gpu"] = bool(torch.equal(img_s, full_s))
                multi["render_64x48_max_abs_diff"] = float((img_s - full_s).abs().max())
        del mr_c, mr_f

    # ---- Instant-NGP variant (config 3, scripts/instantngp.py): fused inference query, hash-grid encode and its backward ----
    ngp = None
    guard.enter("ngp")
    if rank == 0 and not args.no_extract:
        torch.manual_seed(7)
        net = D.InstantNGP("large").to(dev)
        with torch.no_grad():
            net.pos_encoding.params.uniform_(-1.0, 1.0)
        gen = torch.Generator(device=dev).manual_seed(8)
        Pn = 1 << 24                                   # 2^18 rays x 64 samples
        pos_n = torch.rand(Pn, 3, device=dev, generator=gen)
        dir_n = torch.rand(Pn, 3, device=dev, generator=gen)
        table16, dw16, cw16 = net._half_params()

        def timed(fn, reps=3):
            fn()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b) / reps

        ms_q = timed(lambda: D.ops.ngp_forward(pos_n, dir_n, table16, net.grid, dw16, cw16))
        Pe = 1 << 22
        enc_holder = {}
        ms_e = timed(lambda: enc_holder.__setitem__("e", D.ops.hashgrid_forward(pos_n[:Pe], table16, net.grid)))
        d_enc = (torch.rand(Pe, 32, device=dev, generator=gen) - 0.5).half()
        d_tab = torch.zeros(2 * net.grid.n_entries, device=dev)
        ms_b = timed(lambda: D.ops.hashgrid_backward(pos_n[:Pe], d_enc, net.grid, d_table=d_tab))
        # one training step of the cfg.ngp path at config 3's batch: 2^18 rays x (64 + 192) points through NGPTrainer (ONE InstantNGP
        # as coarse and fine model, train.py:219-220): launches of this library only - fused query, compositing + loss, fused
        # backward (recompute + mma.sync dgrad / wgrad + table scatter-add), fused AdamW with fp16 shadow; in-kernel draws.
        # "occupancy-guided sampling" of config 3's wording has no counterpart in the reference (train.py:219-220 feeds the same
        # stratified + hierarchical sampler to the NGP model): it is the reference's sampler that runs here.
        ngp_train = None
        try:
            from common import make_rays
            Bn = 1 << 18
            o_n, d_n, near_n, far_n, tgt_n = (t.to(dev) for t in make_rays(9, Bn))
            near_n, far_n = near_n.reshape(-1), far_n.reshape(-1)
            trn = D.NGPTrainer(net, S_COARSE, N_IMPORTANCE, lr=5e-4, total_steps=64,
                               process_group=False)     # rank 0 alone runs this section: no gradient exchange
            for _ in range(2):
                trn.step(o_n, d_n, near_n, far_n, tgt_n)
            torch.cuda.synchronize()
            trn.profile = True
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record()
            for _ in range(3):
                trn.step(o_n, d_n, near_n, far_n, tgt_n)
            b_.record()
            torch.cuda.synchronize()
            ms_t = a_.elapsed_time(b_) / 3
            st_n = trn.stage_times_ms()
            trn.profile = False
            pts_n = Bn * (S_COARSE + T_FINE)
            ngp_train = {"rays": Bn, "points": pts_n, "ms": ms_t, "rays_per_s": Bn / (ms_t * 1e-3), "points_per_s": pts_n / (ms_t * 1e-3),
                         "stages_ms": {k: round(sum(v) / len(v), 3) for k, v in st_n.items()},
                         "skipped_steps": int(trn.skip.item()),
                         "path": "NGPTrainer: kernels of this library only (no library GEMM, no autograd graph); fused query forward, "
                                 "fused recompute backward, fused AdamW + fp16 shadow, Philox draws in the kernels"}
            del o_n, d_n, near_n, far_n, tgt_n, trn
            torch.cuda.empty_cache()
        except Exception as e:  # an auxiliary measurement must not take the headline line down with it
            ngp_train = {"error": f"{type(e).__name__}: {e}"[:300]}
        table16, dw16, cw16 = net._half_params()
        gather = 16 * 8 * 4                            # 512 B of table gathers per point (upper bound)
        ngp = {"metric": "ngp_query_points_per_s", "value": Pn / (ms_q * 1e-3), "unit": "points/s", "nerf_type": "large",
               "points": Pn, "ms": ms_q, "table_mb_fp16": 4 * net.grid.n_entries / 1e6,
               "algorithmic_gather_gbs": Pn * (gather + 24 + 38) / (ms_q * 1e-3) / 1e9,
               "mlp_tflops": Pn * 2 * 9408 / (ms_q * 1e-3) / 1e12,
               "hashgrid_forward": {"points": Pe, "ms": ms_e, "points_per_s": Pe / (ms_e * 1e-3),
                                    "algorithmic_gbs": Pe * (gather + 12 + 64) / (ms_e * 1e-3) / 1e9},
               "hashgrid_backward": {"points": Pe, "ms": ms_b, "points_per_s": Pe / (ms_b * 1e-3),
                                     "algorithmic_gbs": Pe * (16 * 8 * 8 + 12 + 64) / (ms_b * 1e-3) / 1e9},
               "train_step": ngp_train,
               "note": "table gathers are served by L2 (26 MB table), so the GB/s are gather throughput, not HBM traffic; "
                       "uniform random points = worst-case locality; parity unpinned (tiny-cuda-nn absent)",
               "includes": "hash-grid encode + density MLP + SH4 + colour MLP in one launch, fp16 rgb + features written"}
        line["ngp"] = ngp
        del net, pos_n, dir_n, d_enc, d_tab, enc_holder

    # ---- Instant-NGP data-parallel (scripts/train.py:219-222 under DDP): every rank trains its own 2^16 rays,
    # ONE all-reduce of the flat fp32 gradient [table | dmlp | cmlp] per step.  Weak scaling; max over ranks on the device.
    guard.enter("ngp data-parallel")
    if world > 1 and not args.no_extract:
        ngp_dp = None
        try:
            from common import make_rays
            torch.manual_seed(7)                          # identical initial parameters on every rank
            net_dp = D.InstantNGP("large").to(dev)
            with torch.no_grad():
                net_dp.pos_encoding.params.uniform_(-1.0, 1.0)
            Bd = 1 << 16
            o_n, d_n, near_n, far_n, tgt_n = (t.to(dev) for t in make_rays(90 + rank, Bd))
            near_n, far_n = near_n.reshape(-1), far_n.reshape(-1)
            trd = D.NGPTrainer(net_dp, S_COARSE, N_IMPORTANCE, lr=5e-4, total_steps=64)      # default process group
            for _ in range(2):
                trd.step(o_n, d_n, near_n, far_n, tgt_n)
            barrier()
            trd.profile = True
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record()
            for _ in range(5):
                trd.step(o_n, d_n, near_n, far_n, tgt_n)
            b_.record()
            barrier()
            ms_d = reduce_max(a_.elapsed_time(b_)) / 5
            st_d = trd.stage_times_ms()
            trd.profile = False
            fmax, fmin = trd.flat.clone(), trd.flat.clone()
            dist.all_reduce(fmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(fmin, op=dist.ReduceOp.MIN)
            ngp_dp = {"rays_per_gpu": Bd, "ms": ms_d, "rays_per_s": world * Bd / (ms_d * 1e-3), "scaling": "weak",
                      "allreduce_mb": trd.grads.numel() * 4 / 1e6,
                      "stages_ms_rank0": {k: round(sum(v) / len(v), 3) for k, v in st_d.items()},
                      "note": "stage 'adamw' holds the blocking all-reduce of the whole flat gradient + non-finite guard + AdamW",
                      "param_max_minus_min_over_ranks": float((fmax - fmin).abs().max()),
                      "skipped_steps": int(trd.skip.item())}
            del o_n, d_n, near_n, far_n, tgt_n, trd, net_dp, fmax, fmin
            torch.cuda.empty_cache()
        except Exception as e:  # an auxiliary measurement must not take the headline line down with it
            ngp_dp = {"error": f"{type(e).__name__}: {e}"[:300]}
        if line.get("ngp") is None:
            line["ngp"] = {}
        line["ngp"]["data_parallel"] = ngp_dp

    # ---- same-box comparator: the port of the reference step in stock eager PyTorch ON THIS GPU ------------
    # (what the hand-written kernels buy over library kernels on the same device); 4096 rays, coarse + fine, backward,
    # AdamW; bf16 autocast as scripts/train.py:262 (fp16 there) and fp32 with TF32 matmuls as scripts/train.py:22-24.
    eager = None
    guard.enter("gpu_eager_baseline")
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            from oracle import nerf_oracle as O
            torch.backends.cuda.matmul.allow_tf32 = True
            torch.backends.cudnn.allow_tf32 = True
            eager = {"rays": B_RAYS, "path": "oracle/nerf_oracle.py render_losses + backward + torch AdamW, eager PyTorch CUDA (library kernels)"}
            o_e, d_e, near_e, far_e, tgt_e = (t.to(dev) for t in make_rays(77, B_RAYS))
            for tag, ac in (("bf16_autocast", True), ("fp32_tf32", False)):
                sds = [{k: (v.to(dev).requires_grad_("freq_bands" not in k)) for k, v in make_state_dict(sd_seed, 256).items()}
                       for sd_seed in (31, 32)]
                leaves = [v for sd in sds for k, v in sd.items() if "freq_bands" not in k]
                opt_e = torch.optim.AdamW(leaves, lr=5e-4, weight_decay=1e-6)

                def eager_step():
                    t_rand, u = torch.rand(B_RAYS, S_COARSE, device=dev), torch.rand(B_RAYS, N_IMPORTANCE, device=dev)
                    nc, nf = torch.randn(B_RAYS * S_COARSE, 1, device=dev), torch.randn(B_RAYS * T_FINE, 1, device=dev)
                    opt_e.zero_grad(set_to_none=True)
                    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=ac):
                        out = O.render_losses(sds[0], sds[1], o_e, d_e, near_e, far_e, tgt_e, t_rand, u, nc, nf)
                    out["loss"].backward()
                    opt_e.step()

                eager_step(); eager_step()
                torch.cuda.synchronize()
                g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                g0.record()
                for _ in range(5):
                    eager_step()
                g1.record()
                torch.cuda.synchronize()
                ms_g = g0.elapsed_time(g1) / 5
                eager[tag] = {"ms_per_step": ms_g, "rays_per_s": B_RAYS / (ms_g * 1e-3)}
                del sds, leaves, opt_e
            eager["speedup_of_value_over_bf16_autocast"] = value / eager["bf16_autocast"]["rays_per_s"]
            torch.cuda.empty_cache()
        except Exception as e:  # an auxiliary measurement must not take the headline line down with it
            eager = {"error": f"{type(e).__name__}: {e}"[:300]}
        line["gpu_eager_baseline"] = eager

    cpu = None
    guard.enter("cpu_baseline")
    if rank == 0 and world == 1 and not args.no_cpu:
        v, sample, cores, variants = cpu_rate(args.cpu_budget)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "variants": variants}
        line["cpu_baseline"] = cpu

    guard.enter("final barrier")
    barrier()                      # rank 0 alone ran the last sections: leave together, so that the teardown is not skewed
    if guard.finish() and rank == 0:
        print(json.dumps(with_summary(line)), flush=True)
    if world > 1:
        bye = threading.Timer(30.0, lambda: os._exit(0))      # the line is out: a communicator teardown may not hold the job
        bye.daemon = True
        bye.start()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extract", action="store_true", help="skip the extract_vol measurement")
    ap.add_argument("--extract-res", type=int, default=512)
    ap.add_argument("--cpu-budget", type=float, default=12.0)
    ap.add_argument("--aux-timeout", type=float, default=0.0,
                    help="seconds allowed for everything after the headline measurement (0: 420 at N=1, 150 at N>1)")
    ap.add_argument("--min-seconds", type=float, default=2.5,
                    help="the K timed steps are repeated back to back until the timed region lasts this long")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
