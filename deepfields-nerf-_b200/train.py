"""Drop-in for the reference's scripts/train.py hot path: sample_pdf, raw2outputs,
process_rays_in_chunks, forward_pass (same signatures) plus the fused trainer that the `train(cfg)`
entry point and bench.py use (one launch per stage, flat gradient buffer, fused AdamW)."""
from __future__ import annotations

import math
import datetime
import os
import time
from collections import OrderedDict
from types import SimpleNamespace
from typing import Optional

import torch

from . import ops
from .model import OptimizedVanillaNeRF, get_nerf_models
from .sampling import StratifiedSampler

# a collective that cannot complete (a rank that died, a mismatched call) fails after this long instead of holding the GPUs
# for NCCL's default ten minutes
_PG_TIMEOUT = datetime.timedelta(seconds=180)

# forward_pass reads a module-global cfg exactly like the reference (scripts/train.py:148-177)
cfg = SimpleNamespace(N_samples=64, N_importance=128, chunk_size=32768)


def sample_pdf(bins, weights, N_samples, device=None, u=None):
    """scripts/train.py:31-75.  `u` optionally injects the torch.rand draw of :53."""
    if u is None:
        u = torch.rand(list(bins.shape[:-1]) + [N_samples], device=weights.device)
    return ops.sample_pdf(bins, weights, u)


class _Composite(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rgb, sigma, z_vals, white_bg, dist_last):
        rgb_map, weights = ops.composite_forward(rgb, sigma, z_vals, white_bg, dist_last)
        ctx.save_for_backward(rgb, sigma, z_vals)
        ctx.white_bg, ctx.dist_last = white_bg, dist_last
        return rgb_map, weights

    @staticmethod
    def backward(ctx, d_rgb_map, d_weights):
        rgb, sigma, z_vals = ctx.saved_tensors
        d_rgb, d_sigma = ops.composite_backward(rgb, sigma, z_vals, d_rgb_map.contiguous(),
                                                None if d_weights is None else d_weights.contiguous(),
                                                ctx.white_bg, ctx.dist_last)
        return d_rgb.reshape(rgb.shape).to(rgb.dtype), d_sigma.reshape(sigma.shape).to(sigma.dtype), None, None, None


def raw2outputs(rgb, sigma, z_vals, white_bg=False, dist_last=1e10):
    """scripts/train.py:78-106 -> (rgb_map [B,3], weights [B,S]); differentiable w.r.t. rgb and sigma."""
    return _Composite.apply(rgb, sigma, z_vals, bool(white_bg), float(dist_last))


def process_rays_in_chunks(model, pts, view_pts, device=None, chunk_size=8192):
    """scripts/train.py:110-126.  The fused MLP kernel is persistent over any number of points, so the
    chunk loop of the reference collapses to a single launch (chunk_size is accepted and ignored)."""
    return model(pts, view_pts)


def forward_pass(model_coarse, model_fine, rays_o, rays_d, t_near, t_far, image_rgb, sampler, criterion, device,
                 white_bg, draws: Optional[dict] = None):
    """scripts/train.py:129-186, same argument order and return value (total, coarse, fine losses).
    `draws` optionally injects {"t_rand","u","noise_c","noise_f"} for parity tests."""
    d = draws or {}
    S, I = cfg.N_samples, cfg.N_importance
    pts, view_pts, z_vals = sampler(rays_o, rays_d, t_near, t_far, d.get("t_rand"))
    rgb_c, sig_c = model_coarse(pts, view_pts, d.get("noise_c"))
    rgb_map_c, weights = raw2outputs(rgb_c.reshape(-1, S, 3), sig_c.reshape(-1, S, 1), z_vals, white_bg)
    z_mid = 0.5 * (z_vals[..., :-1] + z_vals[..., 1:])
    z_samples = sample_pdf(z_mid, weights[..., 1:-1].detach(), I, device, d.get("u")).detach()
    pts_f, view_f, z_all = sampler.fine_points(rays_o, rays_d, z_samples, z_vals, t_near, t_far, S, I)
    rgb_f, sig_f = model_fine(pts_f, view_f, d.get("noise_f"))
    rgb_map_f, _ = raw2outputs(rgb_f.reshape(-1, S + I, 3), sig_f.reshape(-1, S + I, 1), z_all, white_bg)
    loss_c = criterion(rgb_map_c, image_rgb)
    loss_f = criterion(rgb_map_f, image_rgb)
    return loss_c + loss_f, loss_c, loss_f


def onecycle_tables(total_steps: int, base_lr: float):
    """OneCycleLR(max_lr=10*lr, pct_start=.1, cos, div_factor=10, final_div_factor=100) with the default
    cycle_momentum (beta1 0.95 -> 0.85 -> 0.95), as scripts/train.py:250-259 builds it."""
    max_lr, init_lr = base_lr * 10.0, base_lr
    min_lr = init_lr / 100.0
    up_end, down_end = float(0.1 * total_steps) - 1.0, float(total_steps) - 1.0
    cos = lambda a, b, pct: b + (a - b) / 2.0 * (math.cos(math.pi * pct) + 1.0)  # noqa: E731
    lrs, b1s = [], []
    for s in range(max(total_steps, 1)):
        if s <= up_end:
            pct = s / up_end if up_end > 0 else 1.0
            lrs.append(cos(init_lr, max_lr, pct)); b1s.append(cos(0.95, 0.85, pct))
        else:
            pct = (s - up_end) / (down_end - up_end) if down_end > up_end else 1.0
            lrs.append(cos(max_lr, min_lr, pct)); b1s.append(cos(0.85, 0.95, pct))
    return lrs, b1s


class FusedTrainer:
    """One training step of scripts/train.py:282-323 as ~12 kernel launches and no autograd graph:
    stratified z -> coarse MLP (rays read in-kernel) -> composite+MSE+backward -> pdf/searchsorted ->
    merge -> fine MLP -> composite+MSE+backward -> MLP dgrad/wgrad (both nets) -> [NCCL all-reduce]
    -> fused AdamW(+OneCycleLR) on the flat buffer -> repack bf16 weights.

    Both networks' parameters live in ONE flat fp32 buffer (coarse then fine: the optimiser's parameter
    order at train.py:227-228), so the data-parallel gradient exchange is a single all-reduce."""

    def __init__(self, model_coarse: OptimizedVanillaNeRF, model_fine: OptimizedVanillaNeRF, N_samples=64,
                 N_importance=128, lr=5e-4, total_steps=1000, white_bg=False, dist_last=1e10, weight_decay=1e-6,
                 train_noise=True, process_group=None, rng="kernel", seed=None, use_graph=None):
        """rng: "kernel" = jitter / inverse-CDF draws / sigma noise are drawn inside the kernels (Philox keyed by `seed`
        and the step counter: no draw reaches HBM, graph-replayable), "torch" = torch.rand / torch.randn tensors as in
        round 1; injected draws (parity tests) always win.  seed: defaults to 0x5EED + 7919 * rank (per-rank streams,
        SURVEY 8e).  use_graph: replay the step as CUDA graphs (default: off; DFN_CUDA_GRAPH=1 turns it on)."""
        dev = next(model_coarse.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("FusedTrainer needs CUDA models (no CPU fallback)")
        self.dev, self.S, self.I = dev, N_samples, N_importance
        self.white_bg, self.dist_last, self.wd = white_bg, dist_last, weight_decay
        self.train_noise = train_noise
        # process_group=False: this trainer is local to the process even though a process group exists (no gradient exchange)
        self.local = process_group is False
        self.pg = process_group = None if self.local else process_group
        self.world = torch.distributed.get_world_size(process_group) if self._dist_on() else 1
        rank = torch.distributed.get_rank(process_group) if self._dist_on() else 0
        self.pg = self._grad_group(process_group)
        self.rng = rng
        self.seed = int(seed) if seed is not None else 0x5EED + 7919 * rank
        self.use_graph = (os.environ.get("DFN_CUDA_GRAPH", "0") == "1") if use_graph is None else bool(use_graph)
        self._graphs = {}
        # overlap = (SMs of the fine forward, SMs of the coarse backward): the coarse network's backward only depends on the
        # coarse loss, so it runs on a side stream BESIDE the fine network's forward, each on its share of the SMs
        # (DFN_OVERLAP="100,48"; unset / "0" = everything in sequence on the whole device)
        ov = os.environ.get("DFN_OVERLAP", "0")
        self.overlap = tuple(int(x) for x in ov.split(",")) if "," in ov else None
        self._side = torch.cuda.Stream(device=dev) if self.overlap else None
        # world > 1: SMs the coarse network's data-gradient launch may use while the fine half of the gradient is all-reduced
        # beside it (0 = all of them; then NCCL's CTAs only become resident when that launch drains)
        self._dp_dgrad_sms = int(os.environ.get("DFN_DP_DGRAD_SMS", "0"))
        if self.overlap and ops.bwd_fused_default():
            # the one-launch backward sizes itself to the whole device and both networks would share one ring workspace on
            # two streams at once: corrupted gradients or the ring's time-out trap
            raise RuntimeError("DFN_BWD_FUSED=1 and DFN_OVERLAP cannot be combined: the fused backward needs the whole device "
                               "and a workspace of its own per concurrent launch")
        self._keep = None
        self.models = (model_coarse, model_fine)
        n = ops.MLP_NUM_PARAMS
        # re-home both models into one flat buffer
        self.flat = torch.empty(2 * n, dtype=torch.float32, device=dev)
        for k, m in enumerate(self.models):
            off = k * n
            for p in m._ordered_params():
                c = p.numel()
                self.flat[off:off + c].copy_(p.detach().reshape(-1))
                p.data = self.flat[off:off + c].view_as(p)
                off += c
            m._flat = self.flat[k * n:(k + 1) * n]
            m.mark_dirty()
            m._sync_packed()
        self.grads = torch.zeros_like(self.flat)
        self.m = torch.zeros_like(self.flat)
        self.v = torch.zeros_like(self.flat)
        lrs, b1s = onecycle_tables(total_steps, lr)
        self.lr_table = torch.tensor(lrs, dtype=torch.float32, device=dev)
        self.b1_table = torch.tensor(b1s, dtype=torch.float32, device=dev)
        self.total_steps = total_steps
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.sampler = StratifiedSampler(N_samples, dev)
        self.loss_sum = torch.zeros(2, dtype=torch.float32, device=dev)
        self._B, self._scratch = None, {}
        # optional per-stage CUDA-event timing (bench.py): list of (stage name, event) per recorded step
        self.profile = False
        self._marks = []
        self.launches_per_step = 0
        self._pending = []
        self._static, self.graph_captures = {}, 0   # static input buffers per batch size; graphs captured so far

    # ---- stage timing -----------------------------------------------------------------------------------
    def _mark(self, name, launches=1):
        """Called after every stage of the step (`launches` = kernels of this library the stage launched).  With
        self.profile an event is recorded on the launching (current) stream, so consecutive marks bracket one stage."""
        self._n_launch += launches
        if self.profile:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._marks.append((name, ev))

    def stage_times_ms(self):
        """{stage: [ms per recorded launch]} from the recorded events (call after a synchronize); clears them."""
        out = {}
        prev = None
        for name, ev in self._marks:
            if name != "begin" and prev is not None:
                out.setdefault(name, []).append(prev.elapsed_time(ev))
            prev = ev
        self._marks = []
        return out

    def _dist_on(self):
        return not self.local and torch.distributed.is_available() and torch.distributed.is_initialized()

    def _grad_group(self, parent):
        """The communicator of the gradient exchange.  The two all-reduces of a step are 2.4 MB each - latency class - and run
        BESIDE the coarse network's backward kernels, so NCCL should take few SMs for them: DFN_GRAD_NCCL_CTAS=n (> 0) gives the
        exchange a communicator of its own limited to n CTAs (ncclConfig max_ctas); 0 / unset = the parent group as it is.
        Collective over the parent group: every rank constructs its trainer at the same point."""
        n = int(os.environ.get("DFN_GRAD_NCCL_CTAS", "0"))
        if n <= 0 or self.world == 1 or torch.distributed.get_backend(parent) != "nccl":
            return parent
        opts = torch.distributed.ProcessGroupNCCL.Options()
        opts.config.max_ctas = n
        opts.config.min_ctas = 1
        ranks = torch.distributed.get_process_group_ranks(parent if parent is not None else torch.distributed.group.WORLD)
        return torch.distributed.new_group(ranks=ranks, backend="nccl", pg_options=opts)

    def _alloc(self, B):
        if self._B == B:
            return
        names = ("stash_c", "stash_f", "dz", "dz_c", "out_c", "out_f")
        if self.use_graph:
            # a captured graph holds the addresses of its batch size's scratch: keep one set per batch size alive instead of
            # freeing the full batch's set when the short last batch of an epoch comes by
            if self._B is not None:
                self._scratch[self._B] = tuple(getattr(self, k) for k in names)
            if B in self._scratch:
                for k, v in zip(names, self._scratch[B]):
                    setattr(self, k, v)
                self._B = B
                return
        S, T, dev = self.S, self.S + self.I, self.dev
        Pc, Pf = B * S, B * T
        u8 = lambda n: torch.empty(n, dtype=torch.uint8, device=dev)  # noqa: E731
        self.stash_c, self.stash_f = u8(ops.mlp_stash_bytes(Pc)), u8(ops.mlp_stash_bytes(Pf))
        self.dz = u8(max(ops.mlp_dz_bytes(Pc), ops.mlp_dz_bytes(Pf)))
        self.dz_c = u8(ops.mlp_dz_bytes(Pc)) if self.overlap else self.dz   # its own scratch when the two backward passes overlap
        self.out_c = (torch.empty(Pc, 3, device=dev), torch.empty(Pc, device=dev))
        self.out_f = (torch.empty(Pf, 3, device=dev), torch.empty(Pf, device=dev))
        self._B = B

    def forward_backward(self, rays_o, rays_d, near, far, target, t_rand=None, u=None, noise_c=None, noise_f=None):
        """Losses + parameter gradients (self.grads) for one batch; no optimiser update."""
        B = rays_o.shape[0]
        S, I, T = self.S, self.I, self.S + self.I
        self._alloc(B)
        mc, mf = self.models
        kern = self.rng == "kernel"
        if t_rand is None and not kern:
            t_rand = torch.rand(B, S, device=self.dev)
        if u is None and not kern:
            u = torch.rand(B, I, device=self.dev)
        rng_c = rng_f = None
        if self.train_noise:
            if noise_c is None:
                if kern:
                    rng_c = (self.seed, self.step_dev, ops.RNG_NOISE_COARSE)
                else:
                    noise_c = torch.randn(B * S, device=self.dev)
            if noise_f is None:
                if kern:
                    rng_f = (self.seed, self.step_dev, ops.RNG_NOISE_FINE)
                else:
                    noise_f = torch.randn(B * T, device=self.dev)
        self._n_launch = 0
        self._mark("begin", 0)
        self.loss_sum.zero_()
        self.grads.zero_()
        self._mark("zero", 0)  # torch memsets, not kernels of this library
        n = ops.MLP_NUM_PARAMS
        if t_rand is None:
            z_c = ops.stratified_sample_rng(near, far, self.sampler.t_vals, self.seed, self.step_dev)
        else:
            z_c = ops.stratified_sample(near, far, self.sampler.t_vals, t_rand)
        self._mark("stratified")
        rays_c = (rays_o, rays_d, z_c)
        rgb_c, sig_c = ops.mlp_forward(mc._packed, rays=rays_c, noise=noise_c, stash=self.stash_c, out=self.out_c, noise_rng=rng_c)
        self._mark("mlp_fwd_coarse")
        _, w_c, d_rgb_c, d_sig_c = ops.render_loss(rgb_c, sig_c, z_c, target, self.loss_sum[0:1], self.white_bg,
                                                   self.dist_last)
        self._mark("render_loss_coarse")
        mk_c = lambda nm: self._mark(nm + "_coarse")  # noqa: E731
        mk_f = lambda nm: self._mark(nm + "_fine")  # noqa: E731
        done_c = None
        if self.overlap:
            # the coarse network's backward starts NOW on the side stream, on its share of the SMs
            main = torch.cuda.current_stream()
            self._side.wait_stream(main)
            with torch.cuda.stream(self._side), ops.sm_limit(self.overlap[1]):
                self._bwd(mc, rgb_c, sig_c, d_rgb_c, d_sig_c, self.stash_c, self.grads[:n], rays_c, lambda nm: None, self.dz_c)
                done_c = torch.cuda.Event()
                done_c.record(self._side)
            self._keep = (rgb_c, sig_c, d_rgb_c, d_sig_c, z_c)    # read by the side stream: alive until the optimiser has waited
        # inverse-CDF resampling + sort in one launch; z_new never reaches HBM
        z_all = ops.sample_merge(z_c, w_c, u, n_new=I, rng=(self.seed, self.step_dev))
        self._mark("sample_merge")
        rays_f = (rays_o, rays_d, z_all)
        with ops.sm_limit(self.overlap[0] if self.overlap else 0):
            rgb_f, sig_f = ops.mlp_forward(mf._packed, rays=rays_f, noise=noise_f, stash=self.stash_f, out=self.out_f, noise_rng=rng_f)
        self._mark("mlp_fwd_fine")
        _, _, d_rgb_f, d_sig_f = ops.render_loss(rgb_f, sig_f, z_all, target, self.loss_sum[1:2], self.white_bg,
                                                 self.dist_last, want_weights=False)
        self._mark("render_loss_fine")
        # fine network first: its half of the flat gradient is all-reduced while the coarse network's backward runs
        self._bwd(mf, rgb_f, sig_f, d_rgb_f, d_sig_f, self.stash_f, self.grads[n:], rays_f, mk_f, self.dz)
        self._reduce_async(self.grads[n:])
        if done_c is not None:
            torch.cuda.current_stream().wait_event(done_c)
        else:
            self._bwd(mc, rgb_c, sig_c, d_rgb_c, d_sig_c, self.stash_c, self.grads[:n], rays_c, mk_c, self.dz_c,
                      dgrad_sms=self._dp_dgrad_sms if self.world > 1 else 0)
        self._reduce_async(self.grads[:n])
        self.aux = {"z": z_c, "z_all": z_all, "weights_coarse": w_c}
        return self.loss_sum / (3.0 * B)  # [loss_coarse, loss_fine] (MSELoss mean)

    def _bwd(self, model, rgb, sig, d_rgb, d_sig, stash, grads, rays, mark, dz, dgrad_sms=0):
        if ops.bwd_fused_default():
            P = rgb.shape[0]
            if getattr(self, "_ws", None) is None or self._ws.numel() < ops.mlp_bwd_fused_ws_bytes(P):
                self._ws = torch.empty(ops.mlp_bwd_fused_ws_bytes(P), dtype=torch.uint8, device=self.dev)
            ops.mlp_backward_fused(model._packed, rgb, sig, d_rgb, d_sig, stash, self._ws, grads, rays=rays)
            mark("bwd_fused")
        else:
            ops.mlp_backward(model._packed, rgb, sig, d_rgb, d_sig, stash, dz, grads, rays=rays, mark=mark, dgrad_sms=dgrad_sms)

    def _reduce_async(self, g):
        """Data-parallel exchange of one network's gradient, started as soon as it is complete (NCCL runs on the process
        group's stream; the optimiser waits for it)."""
        if self.world > 1:
            self._pending.append(torch.distributed.all_reduce(g, group=self.pg, async_op=True))

    def optimizer_step(self):
        for w in self._pending:
            w.wait()
        self._pending = []
        ops.adamw_step(self.flat, self.grads, self.m, self.v, self.lr_table, self.b1_table, self.step_dev,
                       weight_decay=self.wd, grad_scale=1.0 / self.world)
        self._mark("adamw")
        for m in self.models:
            m._packed.pack(m._flat)
            m._packed_version = (m._flat._version, sum(p._version for p in m._ordered_params()))
        self._mark("pack_weights", 2)
        self.launches_per_step = self._n_launch

    # ---- CUDA-graph replay of the step ------------------------------------------------------------------------------------
    def _graph_for(self, B):
        """The whole step (world == 1) or its forward/backward and optimiser halves around the eager NCCL all-reduce
        (world > 1), captured ONCE PER BATCH SIZE on static input buffers and replayed afterwards (`step` copies every
        batch into them: 44 B per ray).  Everything that changes from step to step lives in device memory (step counter ->
        OneCycleLR tables and the Philox counter), so a replay IS the next step.  The warm-up run in front of the capture
        (allocations, lazy initialisation) leaves no trace: parameters, Adam moments and the step counter are put back and
        the bf16 weight images repacked, so the number of optimiser steps equals the number of `step()` calls, as in the
        reference loop (scripts/train.py:285-323).  At world > 1 the replayed path exchanges the whole flat gradient with one
        blocking all-reduce between the two graphs: it does NOT keep the eager path's fine/coarse overlap."""
        g = self._graphs.get(B)
        if g is None:
            tensors = self._static[B]
            saved = [t.clone() for t in (self.flat, self.m, self.v, self.step_dev)]
            stream = torch.cuda.Stream(device=self.dev)
            stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(stream):           # warm-up outside the capture, undone below
                self.forward_backward(*tensors)
                self.optimizer_step()
                for dst, src in zip((self.flat, self.m, self.v, self.step_dev), saved):
                    dst.copy_(src)
                for m in self.models:
                    m._packed.pack(m._flat)
                    m._packed_version = (m._flat._version, sum(p._version for p in m._ordered_params()))
            torch.cuda.current_stream().wait_stream(stream)
            torch.cuda.synchronize()
            fb, opt = torch.cuda.CUDAGraph(), None
            if self.world == 1:
                with torch.cuda.graph(fb):
                    losses = self.forward_backward(*tensors)
                    self.optimizer_step()
            else:
                with torch.cuda.graph(fb):
                    losses = self._forward_backward_no_reduce(*tensors)
                opt = torch.cuda.CUDAGraph()
                with torch.cuda.graph(opt):
                    self.optimizer_step()
            g = (fb, opt, losses)
            self._graphs[B] = g
            self.graph_captures += 1
        return g

    def _forward_backward_no_reduce(self, *tensors):
        world, self.world = self.world, 1
        try:
            return self.forward_backward(*tensors)
        finally:
            self.world = world

    def step(self, rays_o, rays_d, near, far, target, **draws):
        if self.use_graph and not draws and not self.profile and self.rng == "kernel":
            tensors = (rays_o, rays_d, near, far, target)
            B = tensors[0].shape[0]
            static = self._static.get(B)
            if static is None:                        # one set of input buffers per batch size, kept for the trainer's life
                static = self._static[B] = tuple(torch.empty_like(t, memory_format=torch.contiguous_format) for t in tensors)
            for dst, src in zip(static, tensors):
                if dst.data_ptr() != src.data_ptr():
                    dst.copy_(src)
            fb, opt, losses = self._graph_for(B)
            fb.replay()
            if opt is not None:
                torch.distributed.all_reduce(self.grads, group=self.pg)
                opt.replay()
            return losses
        losses = self.forward_backward(rays_o, rays_d, near, far, target, **draws)
        self.optimizer_step()
        return losses

    # ---- host-driven step (what scripts/train.py:285-289 does per iteration: H2D of the batch, then the step) ----
    @staticmethod
    def pack_host_batch(rays_o, rays_d, near, far, rgb):
        """One pinned flat fp32 buffer [o(3B) | d(3B) | near(B) | far(B) | rgb(3B)] -> a single H2D copy per step."""
        flat = torch.cat([rays_o.reshape(-1), rays_d.reshape(-1), near.reshape(-1), far.reshape(-1), rgb.reshape(-1)]).float()
        return flat.contiguous().pin_memory()

    def step_from_host(self, host_batch, **draws):
        """Copies a pack_host_batch() buffer to the device, runs one training step and returns the two losses as a
        host tensor (a blocking 8-byte D2H read, the `.item()` of train.py:325-331)."""
        n = host_batch.numel()
        B = n // 11
        if getattr(self, "_hb", None) is None or self._hb.numel() != n:
            self._hb = torch.empty(n, dtype=torch.float32, device=self.dev)
            self._loss_host = torch.empty(2, dtype=torch.float32).pin_memory()
        self._hb.copy_(host_batch, non_blocking=True)
        b = self._hb
        o, d = b[0:3 * B].view(B, 3), b[3 * B:6 * B].view(B, 3)
        near, far, rgb = b[6 * B:7 * B], b[7 * B:8 * B], b[8 * B:11 * B].view(B, 3)
        losses = self.step(o, d, near, far, rgb, **draws)
        self._loss_host.copy_(losses, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self._loss_host

    # ---- checkpoint in the reference layout (scripts/train.py:343-364) ---------------------------------
    def optimizer_state_dict(self, lr_now: float, beta1_now: float):
        state, idx, off = {}, 0, 0
        step = float(self.step_dev.item())
        for m in self.models:
            for p in m._ordered_params():
                c = p.numel()
                state[idx] = {"step": torch.tensor(step), "exp_avg": self.m[off:off + c].view_as(p).clone(),
                              "exp_avg_sq": self.v[off:off + c].view_as(p).clone()}
                idx += 1
                off += c
        base_lr = float(self.lr_table[0])          # OneCycleLR: initial_lr = max_lr / div_factor = cfg.lr (train.py:250-259)
        group = {"lr": lr_now, "betas": (beta1_now, 0.999), "eps": 1e-8, "weight_decay": self.wd, "amsgrad": False,
                 "maximize": False, "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "decoupled_weight_decay": True,
                 # the keys OneCycleLR adds to the group, so that the checkpoint resumes with OneCycleLR(last_epoch=k)
                 "initial_lr": base_lr, "max_lr": base_lr * 10.0, "min_lr": base_lr / 100.0,
                 "base_momentum": 0.85, "max_momentum": 0.95,
                 "params": list(range(idx))}
        return {"state": state, "param_groups": [group]}

    def save_checkpoint(self, path, iteration):
        i = min(int(self.step_dev.item()), self.total_steps - 1)
        torch.save({
            "iteration": iteration,
            "nerf_coarse_state_dict": OrderedDict((k, v.detach().clone()) for k, v in self.models[0].state_dict().items()),
            "nerf_fine_state_dict": OrderedDict((k, v.detach().clone()) for k, v in self.models[1].state_dict().items()),
            "optimizer_state_dict": self.optimizer_state_dict(float(self.lr_table[i]), float(self.b1_table[i])),
        }, path)


def shard_slice(batch_size: int, rank: int, world: int) -> slice:
    """Rays of one global batch owned by `rank` (rank r takes rays [r*B/G, (r+1)*B/G))."""
    per = batch_size // world
    return slice(rank * per, (rank + 1) * per)


class RayStore:
    """GPU-resident replacement for scripts/dataset.py + DataLoader: the whole .npz (keys rgbs, rays_o,
    rays_d, t_near, t_far; dataset.py:32-60) lives in HBM, batches are gathered with an on-device permutation."""

    def __init__(self, rays_file, device, rank=0, world=1):
        import numpy as np
        d = np.load(rays_file)
        f = lambda k: torch.from_numpy(d[k].astype("float32")).to(device)  # noqa: E731
        self.rgb, self.o, self.d = f("rgbs"), f("rays_o"), f("rays_d")
        self.near, self.far = f("t_near").reshape(-1), f("t_far").reshape(-1)
        self.n, self.rank, self.world = self.rgb.shape[0], rank, world

    @classmethod
    def from_tensors(cls, rays, rank=0, world=1):
        """Build the store from device tensors (precompute_rays.precompute_rays_on_device) instead of an .npz file."""
        self = cls.__new__(cls)
        f = lambda k: rays[k].float().contiguous()  # noqa: E731
        self.rgb, self.o, self.d = f("rgbs"), f("rays_o"), f("rays_d")
        self.near, self.far = f("t_near").reshape(-1), f("t_far").reshape(-1)
        self.n, self.rank, self.world = self.rgb.shape[0], rank, world
        return self

    def batches(self, batch_size, generator=None):
        """One shuffled pass (DataLoader(shuffle=True), train.py:209-217).  Like the reference's loader (drop_last=False)
        the final short batch is yielded too.  Every rank must draw the SAME permutation (pass generators seeded alike);
        rank r then takes its shard_slice of each global batch."""
        perm = torch.randperm(self.n, device=self.rgb.device, generator=generator)
        for i in range(0, self.n, batch_size):
            chunk = perm[i:i + batch_size]
            idx = chunk[shard_slice(chunk.numel(), self.rank, self.world)]
            if idx.numel() == 0:
                continue
            yield self.o[idx], self.d[idx], self.near[idx], self.far[idx], self.rgb[idx]

    def steps_per_epoch(self, batch_size):
        """len(dataloader) of the reference: ceil(n / batch_size)."""
        return (self.n + batch_size - 1) // batch_size

    def __len__(self):
        return self.n


def _open_writer(cfg, rank):
    """SummaryWriter(output_dir / "logs") of scripts/train.py:198-203 (rank 0 only); None when tensorboard is absent."""
    if rank != 0:
        return None
    out_dir = getattr(cfg, "output_dir", None) or os.path.dirname(str(cfg.checkpoint_dir).rstrip("/")) or "."
    try:
        from torch.utils.tensorboard import SummaryWriter
    except ImportError:
        return None
    os.makedirs(str(out_dir), exist_ok=True)
    return SummaryWriter(os.path.join(str(out_dir), "logs"))


def _log_scalars(writer, step, loss_c, loss_f, lr_next):
    """The five scalars of scripts/train.py:333-338, same tags."""
    if writer is None:
        return
    writer.add_scalar("Loss/Coarse", loss_c, step)
    writer.add_scalar("Loss/Fine", loss_f, step)
    writer.add_scalar("Loss/Total", loss_c + loss_f, step)
    writer.add_scalar("PSNR", -10.0 * math.log10(max(loss_f, 1e-12)), step)
    writer.add_scalar("Scheduler Step", lr_next, step)


def train(cfg_in):
    """scripts/train.py:189-370 for the vanilla model (cfg.ngp false): same cfg keys, same checkpoint files, same
    TensorBoard tags and step numbering (step = epoch * steps_per_epoch + idx with epoch starting at 1, :333)."""
    global cfg
    cfg = cfg_in
    if getattr(cfg, "ngp", False):
        return train_ngp(cfg)
    dist_on = "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1
    rank, world = 0, 1
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if dist_on:
        if not torch.distributed.is_initialized():
            torch.distributed.init_process_group("nccl", device_id=dev, timeout=_PG_TIMEOUT)
        rank, world = torch.distributed.get_rank(), torch.distributed.get_world_size()
    ckpt_dir = str(cfg.checkpoint_dir)
    os.makedirs(ckpt_dir, exist_ok=True)
    writer = _open_writer(cfg, rank)
    store = RayStore(str(cfg.rays_file), dev, rank, world)
    steps_per_epoch = store.steps_per_epoch(cfg.batch_size)          # len(dataloader), drop_last=False
    torch.manual_seed(0)                                             # identical initial weights on every rank
    mc, mf = get_nerf_models(cfg.pos_L, cfg.dir_L, cfg.hidden_dim, device=dev)
    tr = FusedTrainer(mc, mf, cfg.N_samples, cfg.N_importance, cfg.lr, max(cfg.num_epochs * steps_per_epoch, 1),
                      bool(cfg.white_bg))
    perm_gen = torch.Generator(device=dev)
    perm_gen.manual_seed(0)                                          # the SAME shuffle on every rank ...
    # ... but per-rank jitter / inverse-CDF / noise streams: the kernels draw them from Philox keyed by the trainer's
    # seed 0x5EED + 7919 * rank; torch's generator only matters for FusedTrainer(rng="torch")
    torch.manual_seed(1 + rank)
    t0 = time.time()
    for epoch in range(1, cfg.num_epochs + 1):
        for it, (o, d, near, far, rgb) in enumerate(store.batches(cfg.batch_size, perm_gen)):
            losses = tr.step(o, d, near, far, rgb)
            if it % 100 == 0:
                if world > 1:                                         # the global-batch losses (3 floats every 100 steps)
                    losses = losses.clone()
                    torch.distributed.all_reduce(losses)
                    losses /= world
                if rank == 0:
                    lc, lf = losses.tolist()
                    psnr = -10.0 * math.log10(max(lf, 1e-12))
                    print(f"| Epoch: {epoch} | Iteration: {it} | Loss: {lc + lf:.4f} (Coarse: {lc:.4f}, Fine: {lf:.4f}) | "
                          f"PSNR: {psnr:.2f} | Time: {time.time() - t0:.2f}s |")
                    done = min(int(tr.step_dev.item()), tr.total_steps - 1)
                    _log_scalars(writer, epoch * steps_per_epoch + it, lc, lf, float(tr.lr_table[done]))
        if epoch % 2 == 0 and rank == 0:
            tr.save_checkpoint(os.path.join(ckpt_dir, f"nerf_epoch_{epoch:06d}.pth"), epoch)
    if rank == 0:
        tr.save_checkpoint(os.path.join(ckpt_dir, "nerf_final.pth"), cfg.num_epochs)
        print(f"Training completed in {time.time() - t0:.2f}s")
    if writer is not None:
        writer.close()
    return tr


class NGPTrainer:
    """One training step of scripts/train.py:282-323 with cfg.ngp true (:219-220: ONE InstantNGP is both the coarse and the
    fine model) as launches of this library only - no autograd graph, no library GEMM:
        stratified z (in-kernel jitter) -> points -> fused query (encode + both tiny MLPs) -> compositing + MSE + backward
        -> inverse-CDF resampling + sort -> points -> fused query -> compositing + MSE + backward
        -> fused backward of the query for the fine and the coarse points (recompute + mma.sync dgrad / wgrad + table
        scatter-add, gradients accumulate) -> [ONE NCCL all-reduce of the flat gradient] -> non-finite guard ->
        fused AdamW(+OneCycleLR) on the flat fp32 parameters, which also refreshes their fp16 copy.
    All parameters (hash table, density MLP, colour MLP) live in ONE flat fp32 buffer in the module's parameter order; the
    nn.Parameters are views of it, so state_dict() keeps the reference keys (pos_encoding.params, dmlp.params, cmlp.params).
    Like the reference, world coordinates and view directions go to the encodings unnormalised (instantngp.py:77,98).
    tiny-cuda-nn parity is unpinned; loss_scale = tcnn's default 128 (FIXED: the reference additionally runs under GradScaler's dynamic scale, initial 65536, so at
    small batches its smallest fp16 gradient fragments survive where these flush to zero - pass a larger loss_scale for batches
    of a few thousand rays; d_rgb is bounded by 2 / (3 B) per sample, the guard below catches an overflow), steps with a non-finite gradient are
    skipped as torch.amp.GradScaler.step does (train.py:262-323)."""

    def __init__(self, model, N_samples=64, N_importance=128, lr=5e-4, total_steps=1000, white_bg=False, dist_last=1e10,
                 weight_decay=1e-6, process_group=None, seed=None, loss_scale=128.0):
        dev = model.pos_encoding.params.device
        if dev.type != "cuda":
            raise RuntimeError("NGPTrainer needs a CUDA model (no CPU fallback)")
        self.model, self.dev, self.S, self.I = model, dev, N_samples, N_importance
        self.white_bg, self.dist_last, self.wd, self.loss_scale = white_bg, dist_last, weight_decay, float(loss_scale)
        # process_group=False: this trainer is local to the process even though a process group exists (no gradient exchange)
        local = process_group is False
        self.pg = process_group = None if local else process_group
        dist_on = not local and torch.distributed.is_available() and torch.distributed.is_initialized()
        self.world = torch.distributed.get_world_size(process_group) if dist_on else 1
        rank = torch.distributed.get_rank(process_group) if dist_on else 0
        self.seed = int(seed) if seed is not None else 0x5EED + 7919 * rank
        ps = [model.pos_encoding.params, model.dmlp.params, model.cmlp.params]
        self.sizes = [p.numel() for p in ps]
        n = sum(self.sizes)
        self.flat = torch.empty(n, dtype=torch.float32, device=dev)
        off = 0
        for p in ps:
            c = p.numel()
            self.flat[off:off + c].copy_(p.detach().reshape(-1))
            p.data = self.flat[off:off + c]
            off += c
        self.half = self.flat.half()                      # the operands of the kernels, refreshed by the optimiser
        self.grads = torch.zeros_like(self.flat)
        self.m, self.v = torch.zeros_like(self.flat), torch.zeros_like(self.flat)
        lrs, b1s = onecycle_tables(total_steps, lr)
        self.lr_table = torch.tensor(lrs, dtype=torch.float32, device=dev)
        self.b1_table = torch.tensor(b1s, dtype=torch.float32, device=dev)
        self.total_steps = total_steps
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.skip = torch.zeros(1, dtype=torch.int32, device=dev)
        # counter of the in-kernel draws (jitter, inverse-CDF): advances on EVERY step() call, also on one the non-finite guard
        # skipped, so that the retry sees fresh samples (step_dev, which indexes the schedule and Adam's bias correction, stays)
        self.draw_step = torch.zeros(1, dtype=torch.int32, device=dev)
        self.sampler = StratifiedSampler(N_samples, dev)
        self.loss_sum = torch.zeros(2, dtype=torch.float32, device=dev)
        self.profile, self._marks = False, []

    def _views(self, buf):
        t, d = self.sizes[0], self.sizes[1]
        return buf[:t], buf[t:t + d], buf[t + d:]

    def _mark(self, name):
        if self.profile:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._marks.append((name, ev))

    stage_times_ms = FusedTrainer.stage_times_ms

    def forward_backward(self, rays_o, rays_d, near, far, target, t_rand=None, u=None):
        B, S, I = rays_o.shape[0], self.S, self.I
        T = S + I
        grid = self.model.grid
        tab, dw, cw = self._views(self.half)
        g_t, g_d, g_c = self._views(self.grads)
        self._mark("begin")
        self.loss_sum.zero_()
        self.grads.zero_()
        self._mark("zero")
        if t_rand is None:
            z_c = ops.stratified_sample_rng(near, far, self.sampler.t_vals, self.seed, self.draw_step)
        else:
            z_c = ops.stratified_sample(near, far, self.sampler.t_vals, t_rand)
        pts_c = ops.points_on_rays(rays_o, rays_d, z_c, want_dirs=False)
        self._mark("sample_coarse")
        rgb_c, sig_c = ops.ngp_forward_f32(pts_c, rays_d, tab, grid, dw, cw, dirs_group=S)
        self._mark("query_coarse")
        _, w_c, d_rgb_c, d_sig_c = ops.render_loss(rgb_c, sig_c, z_c, target, self.loss_sum[0:1], self.white_bg, self.dist_last)
        self._mark("render_loss_coarse")
        z_all = ops.sample_merge(z_c, w_c, u, n_new=I, rng=(self.seed, self.draw_step))
        pts_f = ops.points_on_rays(rays_o, rays_d, z_all, want_dirs=False)
        self._mark("sample_fine")
        rgb_f, sig_f = ops.ngp_forward_f32(pts_f, rays_d, tab, grid, dw, cw, dirs_group=T)
        self._mark("query_fine")
        _, _, d_rgb_f, d_sig_f = ops.render_loss(rgb_f, sig_f, z_all, target, self.loss_sum[1:2], self.white_bg, self.dist_last,
                                                 want_weights=False)
        self._mark("render_loss_fine")
        ops.ngp_backward(pts_f, rays_d, tab, grid, dw, cw, d_rgb_f, d_sig_f, g_t, g_d, g_c, dirs_group=T, loss_scale=self.loss_scale)
        self._mark("backward_fine")
        ops.ngp_backward(pts_c, rays_d, tab, grid, dw, cw, d_rgb_c, d_sig_c, g_t, g_d, g_c, dirs_group=S, loss_scale=self.loss_scale)
        self._mark("backward_coarse")
        self.aux = {"z": z_c, "z_all": z_all, "weights_coarse": w_c}
        return self.loss_sum / (3.0 * B)

    def optimizer_step(self):
        if self.world > 1:
            torch.distributed.all_reduce(self.grads, group=self.pg)       # ONE collective: table + both MLPs
        ops.nonfinite_flag(self.grads, self.skip)
        ops.adamw_step_shadow(self.flat, self.grads, self.m, self.v, self.lr_table, self.b1_table, self.step_dev, self.half,
                              skip_flag=self.skip, weight_decay=self.wd, grad_scale=1.0 / self.world)
        self.model._half = None          # the module's cached fp16 copies are stale: the next module call re-derives them
        self.draw_step.add_(1)
        self._mark("adamw")

    def step(self, rays_o, rays_d, near, far, target, **draws):
        losses = self.forward_backward(rays_o, rays_d, near, far, target, **draws)
        self.optimizer_step()
        return losses

    def optimizer_state_dict(self):
        """torch.optim.AdamW layout over model.parameters() = [pos_encoding.params, dir_encoding.params (empty: no state, as
        torch keeps none for a parameter that never had a gradient), dmlp.params, cmlp.params] (instantngp.py:23-64)."""
        i = min(int(self.step_dev.item()), self.total_steps - 1)
        step = float(self.step_dev.item())
        state, off = {}, 0
        for idx, c in zip((0, 2, 3), self.sizes):
            state[idx] = {"step": torch.tensor(step), "exp_avg": self.m[off:off + c].clone(), "exp_avg_sq": self.v[off:off + c].clone()}
            off += c
        base_lr = float(self.lr_table[0])
        group = {"lr": float(self.lr_table[i]), "betas": (float(self.b1_table[i]), 0.999), "eps": 1e-8, "weight_decay": self.wd,
                 "amsgrad": False, "maximize": False, "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "decoupled_weight_decay": True, "initial_lr": base_lr, "max_lr": base_lr * 10.0, "min_lr": base_lr / 100.0,
                 "base_momentum": 0.85, "max_momentum": 0.95, "params": [0, 1, 2, 3]}
        return {"state": state, "param_groups": [group]}

    def save_checkpoint(self, path, iteration):
        sd = OrderedDict((k, v.detach().clone()) for k, v in self.model.state_dict().items())
        torch.save({"iteration": iteration, "nerf_coarse_state_dict": sd, "nerf_fine_state_dict": sd,
                    "optimizer_state_dict": self.optimizer_state_dict()}, path)


def train_ngp(cfg_in):
    """scripts/train.py:189-370 with cfg.ngp true (:219-220): ONE InstantNGP serves as coarse and fine model; AdamW +
    OneCycleLR as :244-259.  The step is NGPTrainer's: only kernels of this library, one flat gradient all-reduce.
    tiny-cuda-nn parity is unpinned."""
    global cfg
    cfg = cfg_in
    from .model import InstantNGP
    dist_on = "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1
    rank, world = 0, 1
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if dist_on:
        if not torch.distributed.is_initialized():
            torch.distributed.init_process_group("nccl", device_id=dev, timeout=_PG_TIMEOUT)
        rank, world = torch.distributed.get_rank(), torch.distributed.get_world_size()
    ckpt_dir = str(cfg.checkpoint_dir)
    os.makedirs(ckpt_dir, exist_ok=True)
    writer = _open_writer(cfg, rank)
    store = RayStore(str(cfg.rays_file), dev, rank, world)
    steps_per_epoch = store.steps_per_epoch(cfg.batch_size)
    torch.manual_seed(0)                                             # identical initial weights on every rank
    model = InstantNGP(getattr(cfg, "nerf_type", "large")).to(dev)
    tr = NGPTrainer(model, cfg.N_samples, cfg.N_importance, cfg.lr, max(cfg.num_epochs * steps_per_epoch, 1), bool(cfg.white_bg))
    perm_gen = torch.Generator(device=dev)
    perm_gen.manual_seed(0)
    t0 = time.time()
    for epoch in range(1, cfg.num_epochs + 1):
        for it, (o, d, near, far, rgb) in enumerate(store.batches(cfg.batch_size, perm_gen)):
            losses = tr.step(o, d, near, far, rgb)
            if it % 100 == 0:
                if world > 1:
                    losses = losses.clone()
                    torch.distributed.all_reduce(losses)
                    losses /= world
                if rank == 0:
                    lc, lf = losses.tolist()
                    print(f"| Epoch: {epoch} | Iteration: {it} | Loss: {lc + lf:.4f} (Coarse: {lc:.4f}, Fine: {lf:.4f}) | "
                          f"PSNR: {-10.0 * math.log10(max(lf, 1e-12)):.2f} | Time: {time.time() - t0:.2f}s |")
                    done = min(int(tr.step_dev.item()), tr.total_steps - 1)
                    _log_scalars(writer, epoch * steps_per_epoch + it, lc, lf, float(tr.lr_table[done]))
        if epoch % 2 == 0 and rank == 0:
            tr.save_checkpoint(os.path.join(ckpt_dir, f"nerf_epoch_{epoch:06d}.pth"), epoch)
    if rank == 0:
        tr.save_checkpoint(os.path.join(ckpt_dir, "nerf_final.pth"), cfg.num_epochs)
        print(f"Training completed in {time.time() - t0:.2f}s")
    if writer is not None:
        writer.close()
    return tr


def main(argv=None):
    """CLI of scripts/train.py:373-380: `--cfg_path cfg/<scene>.yml`."""
    import argparse
    from .config import load_config
    ap = argparse.ArgumentParser(description="Train a NeRF (B200-native hot path)")
    ap.add_argument("--cfg_path", type=str, required=True)
    args = ap.parse_args(argv)
    train(load_config(args.cfg_path))


if __name__ == "__main__":
    main()
