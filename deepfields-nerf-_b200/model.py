"""Drop-in for the reference's scripts/model.py: same module / parameter names (so nerf_final.pth state
dicts load both ways), forward and query_density run the fused tcgen05 MLP kernel.

The fused kernel is specialised to the configuration every shipped cfg uses (cfg/base_cfg.yml:59-61:
hidden_dim 256, pos_L 10, dir_L 4, inputs included); other shapes raise -- there is no fallback path.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops


class OptimizedPositionalEncoding(nn.Module):
    """scripts/model.py:6-53: [x | sin(x 2^k pi) | cos(x 2^k pi)], dim-major / frequency-minor."""

    def __init__(self, L=10, include_input=True):
        super().__init__()
        self.L = L
        self.include_input = include_input
        self.output_dim = 3 + (3 * 2 * L) if include_input else 3 * 2 * L
        self.register_buffer("freq_bands", 2.0 ** torch.arange(L) * torch.pi)

    def forward(self, x):
        return ops.posenc(x, self.freq_bands, self.include_input)


_LAYERS = (  # (attribute path, out, in) in named_parameters() order
    ("pos_layers.0", 256, 63), ("pos_layers.2", 256, 256), ("pos_layers.4", 256, 256), ("pos_layers.6", 256, 256),
    ("skip_layers.0", 256, 319), ("skip_layers.2", 256, 256), ("skip_layers.4", 256, 256), ("skip_layers.6", 256, 256),
    ("density_layer", 1, 256), ("feature_layer", 256, 256), ("color_layer.0", 128, 283), ("color_layer.2", 3, 128),
)


class _MLPFunction(torch.autograd.Function):
    """autograd bridge: forward = fused kernel with activation stash, backward = dgrad + wgrad kernels."""

    @staticmethod
    def forward(ctx, module, pos, view, noise, *params):
        module._sync_packed()
        P = pos.shape[0]
        stash = torch.empty(ops.mlp_stash_bytes(P), dtype=torch.uint8, device=pos.device)
        rgb, sigma = ops.mlp_forward(module._packed, pos=pos, view=view, noise=noise, stash=stash)
        ctx.module, ctx.stash, ctx.pos, ctx.view = module, stash, pos, view
        ctx.save_for_backward(rgb, sigma)
        return rgb, sigma.unsqueeze(-1)

    @staticmethod
    def backward(ctx, d_rgb, d_sigma):
        rgb, sigma = ctx.saved_tensors
        m = ctx.module
        P = rgb.shape[0]
        grads = torch.zeros(ops.MLP_NUM_PARAMS, dtype=torch.float32, device=rgb.device)
        dz = torch.empty(ops.mlp_dz_bytes(P), dtype=torch.uint8, device=rgb.device)
        ops.mlp_backward(m._packed, rgb, sigma, d_rgb.float().contiguous(), d_sigma.float().reshape(-1).contiguous(),
                         ctx.stash, dz, grads, pos=ctx.pos, view=ctx.view)
        outs, off = [], 0
        for p in m._ordered_params():
            outs.append(grads[off:off + p.numel()].view_as(p))
            off += p.numel()
        return (None, None, None, None, *outs)


class OptimizedVanillaNeRF(nn.Module):
    """scripts/model.py:56-187."""

    def __init__(self, pos_L=10, dir_L=4, hidden_dim=256, include_input_pos=True, include_input_dir=True):
        super().__init__()
        self.pos_encoding = OptimizedPositionalEncoding(L=pos_L, include_input=include_input_pos)
        self.dir_encoding = OptimizedPositionalEncoding(L=dir_L, include_input=include_input_dir)
        pos_dim, dir_dim = self.pos_encoding.output_dim, self.dir_encoding.output_dim
        H = hidden_dim
        relu = lambda: nn.ReLU(inplace=True)  # noqa: E731  (parameter-free; keeps the Sequential indices 0,2,4,6)
        self.pos_layers = nn.Sequential(nn.Linear(pos_dim, H), relu(), nn.Linear(H, H), relu(), nn.Linear(H, H), relu(),
                                        nn.Linear(H, H), relu())
        self.skip_layers = nn.Sequential(nn.Linear(H + pos_dim, H), relu(), nn.Linear(H, H), relu(), nn.Linear(H, H),
                                         relu(), nn.Linear(H, H), relu())
        self.density_layer = nn.Linear(H, 1)
        nn.init.constant_(self.density_layer.bias, 0.1)             # model.py:126
        nn.init.normal_(self.density_layer.weight, mean=0.0, std=0.01)  # model.py:127
        self.feature_layer = nn.Linear(H, H)
        self.color_layer = nn.Sequential(nn.Linear(H + dir_dim, H // 2), relu(), nn.Linear(H // 2, 3))
        self._fused_ok = (pos_L == 10 and dir_L == 4 and hidden_dim == 256 and include_input_pos and include_input_dir)
        self._flat = None
        self._packed = None
        self._packed_version = -1

    # ---- flat parameter storage (reference named_parameters() order) ---------------------------------
    def _ordered_params(self):
        out = []
        for path, _, _ in _LAYERS:
            mod = self.get_submodule(path)
            out += [mod.weight, mod.bias]
        return out

    def flatten_parameters(self):
        """Re-homes every parameter as a view of one flat fp32 buffer (kept across optimiser steps)."""
        params = self._ordered_params()
        dev = params[0].device
        if self._flat is not None and self._flat.device == dev and params[0].data_ptr() == self._flat.data_ptr():
            return self._flat
        flat = torch.empty(sum(p.numel() for p in params), dtype=torch.float32, device=dev)
        off = 0
        for p in params:
            n = p.numel()
            flat[off:off + n].copy_(p.detach().reshape(-1))
            p.data = flat[off:off + n].view_as(p)
            off += n
        self._flat = flat
        self._packed_version = -1
        return flat

    def mark_dirty(self):
        """Call after writing the flat parameter buffer through a raw pointer (fused optimiser)."""
        self._packed_version = -1

    def _sync_packed(self):
        if not self._fused_ok:
            raise NotImplementedError("the fused sm_100a MLP supports hidden_dim=256, pos_L=10, dir_L=4 with inputs "
                                      "included (cfg/base_cfg.yml:59-61); no fallback path exists")
        flat = self.flatten_parameters()
        if not flat.is_cuda:
            raise RuntimeError("OptimizedVanillaNeRF: parameters must live on a CUDA device (no CPU fallback)")
        if self._packed is None or self._packed.fwd.device != flat.device:
            self._packed = ops.PackedMLP(flat.device)
            self._packed_version = -1
        # in-place updates (optimizer.step, load_state_dict) bump the parameters' version counters
        ver = (flat._version, sum(p._version for p in self._ordered_params()))
        if self._packed_version != ver:
            self._packed.pack(flat)
            self._packed_version = ver

    # ---- reference API ---------------------------------------------------------------------------------
    def _run(self, pos, view, noise=None):
        lead = pos.shape[:-1]
        pos2 = pos.reshape(-1, 3).float().contiguous()
        view2 = view.reshape(-1, 3).float().contiguous()
        if self.training and noise is None:
            noise = torch.randn(pos2.shape[0], device=pos2.device)  # model.py:160-162
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            rgb, sigma = _MLPFunction.apply(self, pos2, view2, noise, *self._ordered_params())
        else:
            self._sync_packed()
            rgb, sigma = ops.mlp_forward(self._packed, pos=pos2, view=view2, noise=noise)
            sigma = sigma.unsqueeze(-1)
        return rgb.reshape(*lead, 3), sigma.reshape(*lead, 1)

    def forward(self, pos, view, noise=None):
        """model.py:173-187 -> (rgb [...,3], sigma [...,1]).  `noise` optionally injects the sigma-noise draw."""
        return self._run(pos, view, noise)

    def query_density(self, pos, noise=None):
        """model.py:139-171 -> (sigma [...,1], features [...,hidden]); inference-only helper."""
        self._sync_packed()
        lead = pos.shape[:-1]
        pos2 = pos.reshape(-1, 3).float().contiguous()
        P = pos2.shape[0]
        if self.training and noise is None:
            noise = torch.randn(P, device=pos2.device)
        stash = torch.empty(ops.mlp_stash_bytes(P), dtype=torch.uint8, device=pos2.device)
        view = torch.zeros_like(pos2)
        _, sigma = ops.mlp_forward(self._packed, pos=pos2, view=view, noise=noise, stash=stash)
        nt = ops.mlp_num_tiles(P)
        f = stash[8 * nt * 65536: 9 * nt * 65536].view(torch.bfloat16).view(nt, 32, ops.TILE, 8)
        feat = f.permute(0, 2, 1, 3).reshape(nt * ops.TILE, 256)[:P]
        return sigma.reshape(*lead, 1), feat.reshape(*lead, 256)


def get_nerf_models(pos_L=10, dir_L=4, hidden_dim=256, include_input_pos=True, include_input_dir=True, device="cuda"):
    """scripts/model.py:190-216: two independent networks (coarse, fine) on `device`."""
    coarse = OptimizedVanillaNeRF(pos_L, dir_L, hidden_dim, include_input_pos, include_input_dir).to(device)
    fine = OptimizedVanillaNeRF(pos_L, dir_L, hidden_dim, include_input_pos, include_input_dir).to(device)
    return coarse, fine


class _TcnnParams(nn.Module):
    """Stands in for a tcnn.Encoding / tcnn.Network: one flat fp32 `params` tensor (tinycudann/modules.py), so the
    state_dict keys are `pos_encoding.params`, `dir_encoding.params`, `dmlp.params`, `cmlp.params` like the reference."""

    def __init__(self, n_params, n_output_dims):
        super().__init__()
        self.n_output_dims = n_output_dims
        self.params = nn.Parameter(torch.zeros(n_params, dtype=torch.float32))


class _NGPFunction(torch.autograd.Function):
    """InstantNGP.forward under autograd on this library's kernels only: forward = the fused query (encode + both tiny MLPs in
    one launch), backward = dfn_ngp_backward (recompute + mma.sync dgrad / wgrad + table scatter-add in one launch)."""

    @staticmethod
    def forward(ctx, module, pos, view, table_p, dmlp_p, cmlp_p):
        table, dw, cw = module._half_params()
        rgb, features = ops.ngp_forward(pos, view, table, module.grid, dw, cw)
        ctx.module, ctx.pos, ctx.view = module, pos, view
        return rgb, features

    @staticmethod
    def backward(ctx, d_rgb, d_features):
        m = ctx.module
        table, dw, cw = m._half_params()
        if d_features is not None and d_features.shape[1] > 1 and bool(d_features[:, 1:].any()):
            raise NotImplementedError("InstantNGP: only sigma (feature 0) and rgb carry gradients on the fused training path; "
                                      "the other 15 density features are consumed inside the kernel")
        P = ctx.pos.shape[0]
        dev = ctx.pos.device
        d_rgb = torch.zeros(P, 3, device=dev) if d_rgb is None else d_rgb.float().contiguous()
        d_sig = torch.zeros(P, device=dev) if d_features is None else d_features[:, 0].float().contiguous()
        g_t = torch.zeros(2 * m.grid.n_entries, device=dev)
        g_d = torch.zeros(ops.NGP_DMLP_PARAMS, device=dev)
        g_c = torch.zeros(ops.NGP_CMLP_PARAMS, device=dev)
        ops.ngp_backward(ctx.pos, ctx.view, table, m.grid, dw, cw, d_rgb, d_sig, g_t, g_d, g_c)
        return None, None, None, g_t, g_d, g_c


class InstantNGP(nn.Module):
    """scripts/instantngp.py:7-104.  Same constructor contract (ValueError for an unknown nerf_type, :9-21), same
    submodule names, `query_density(pos) -> (sigma, features)` and `forward(pos, viewdirs) -> (rgb, sigma)` returning
    fp16 like tcnn.  The arithmetic of tiny-cuda-nn (un-vendored, commit 74ea173d) is restated: parity unpinned
    (DESIGN.md section 2).  Inference (no grad) is ONE fused launch (`dfn_ngp_forward`); with grad enabled the hash-grid
    backward is ONE more launch of this library (csrc/ngp_train.cu: recompute + mma.sync dgrad / wgrad + table scatter-add):
    no library GEMM anywhere on the path."""

    LOG2_SIZE = {"small": 15, "medium": 17, "large": 19}

    def __init__(self, nerf_type: str = "small"):
        super().__init__()
        if nerf_type not in self.LOG2_SIZE:
            raise ValueError(f"Unkown {nerf_type=}")
        self.nerf_type = nerf_type
        self.grid = ops.hashgrid_levels(self.LOG2_SIZE[nerf_type])
        self.pos_encoding = _TcnnParams(2 * self.grid.n_entries, 32)
        self.dir_encoding = _TcnnParams(0, 16)
        self.dmlp = _TcnnParams(ops.NGP_DMLP_PARAMS, 16)
        self.cmlp = _TcnnParams(ops.NGP_CMLP_PARAMS, 3)
        with torch.no_grad():  # tcnn initial_params: grid U(-1e-4, 1e-4), MLP layers Xavier-uniform
            self.pos_encoding.params.uniform_(-1e-4, 1e-4)
            for mod, shapes in ((self.dmlp, ((64, 32), (16, 64))), (self.cmlp, ((64, 32), (64, 64), (16, 64)))):
                o = 0
                for r, c in shapes:
                    bound = (6.0 / (r + c)) ** 0.5
                    mod.params[o:o + r * c].uniform_(-bound, bound)
                    o += r * c
        self._half = None

    def _half_params(self):
        """fp16 copies of the flat parameters (tcnn casts per call); rebuilt when a parameter changed."""
        ps = (self.pos_encoding.params, self.dmlp.params, self.cmlp.params)
        key = tuple((p.data_ptr(), p._version) for p in ps)
        if self._half is None or self._half[0] != key:
            self._half = (key, tuple(p.detach().half() for p in ps))
        return self._half[1]

    @staticmethod
    def _flat(x, name):
        if not x.is_cuda:
            raise RuntimeError(f"InstantNGP: {name} must be a CUDA tensor (no CPU fallback)")
        return x.reshape(-1, 3).float()

    def _training_path(self):
        return torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())

    def _fused_autograd(self, pos, view):
        return _NGPFunction.apply(self, pos, view, self.pos_encoding.params, self.dmlp.params, self.cmlp.params)

    def query_density(self, pos):
        """instantngp.py:66-80"""
        pos = self._flat(pos, "pos")
        if self._training_path():
            # (gradients flow through sigma = feature 0 only; the colour head is evaluated with a dummy view and discarded)
            _, features = self._fused_autograd(pos, torch.tensor([0.5, 0.5, 0.5], device=pos.device))
        else:
            table, dw, cw = self._half_params()
            _, features = ops.ngp_forward(pos, None, table, self.grid, dw, cw, want_features=True, want_rgb=False)
        return features[:, :1], features

    def forward(self, pos, viewdirs, noise=None):
        """instantngp.py:82-104.  `noise` (the vanilla model's injected sigma noise) is accepted and ignored: this variant
        adds none (:77-80)."""
        pos = self._flat(pos, "pos")
        view = viewdirs.float()
        if view.numel() != 3 and not (view.dim() == 2 and view.stride(0) == 0):
            view = self._flat(view, "viewdirs")
        else:
            view = view.reshape(-1, 3)[0]  # an expanded canonical view (extract_vol.py:100) stays one direction
        if self._training_path():
            rgb, features = self._fused_autograd(pos, view.contiguous())
            return rgb, features[:, :1]
        table, dw, cw = self._half_params()
        rgb, features = ops.ngp_forward(pos, view, table, self.grid, dw, cw)
        return rgb, features[:, :1]
