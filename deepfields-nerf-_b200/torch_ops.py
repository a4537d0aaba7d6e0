"""`torch.library` registration of the hot-path operators (north_star: "a thin C-ABI torch custom-op layer").

The kernels live behind the C ABI of libdeepfields_b200.so (include/deepfields_b200.h: raw pointers, no torch types); this
module makes the differentiable ones visible to PyTorch as operators of the `dfn` namespace - `torch.ops.dfn.raw2outputs`,
`torch.ops.dfn.vanilla_mlp`, `torch.ops.dfn.sample_pdf` - with fake (meta) implementations for tracing and autograd formulas
registered through `register_autograd`, so they compose with `torch.autograd`, `torch.compile` graphs and CUDA-graph capture like
any built-in op.  The Python mirrors of the reference's functions (train.raw2outputs, model.OptimizedVanillaNeRF.forward,
train.sample_pdf) call the same ctypes entry points; the registered ops are the form a maintainer of the reference would bind.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Tuple

import torch
from torch import Tensor

from . import ops

__all__ = ["raw2outputs", "vanilla_mlp", "sample_pdf"]


# ---- raw2outputs (scripts/train.py:78-106) -----------------------------------------------------------------------------------
@torch.library.custom_op("dfn::raw2outputs", mutates_args=(), device_types="cuda")
def raw2outputs(rgb: Tensor, sigma: Tensor, z_vals: Tensor, white_bg: bool, dist_last: float) -> Tuple[Tensor, Tensor]:
    return ops.composite_forward(rgb, sigma, z_vals, white_bg, dist_last)


@raw2outputs.register_fake
def _(rgb, sigma, z_vals, white_bg, dist_last):
    B, S = z_vals.shape
    return z_vals.new_empty(B, 3), z_vals.new_empty(B, S)


@torch.library.custom_op("dfn::raw2outputs_backward", mutates_args=(), device_types="cuda")
def _raw2outputs_backward(rgb: Tensor, sigma: Tensor, z_vals: Tensor, d_rgb_map: Tensor, d_weights: Tensor, white_bg: bool,
                          dist_last: float) -> Tuple[Tensor, Tensor]:
    return ops.composite_backward(rgb, sigma, z_vals, d_rgb_map.contiguous(), d_weights.contiguous(), white_bg, dist_last)


@_raw2outputs_backward.register_fake
def _(rgb, sigma, z_vals, d_rgb_map, d_weights, white_bg, dist_last):
    B, S = z_vals.shape
    return z_vals.new_empty(B, S, 3), z_vals.new_empty(B, S)


def _r2o_setup(ctx, inputs, output):
    rgb, sigma, z_vals, white_bg, dist_last = inputs
    ctx.save_for_backward(rgb, sigma, z_vals)
    ctx.white_bg, ctx.dist_last = white_bg, dist_last


def _r2o_backward(ctx, d_rgb_map, d_weights):
    rgb, sigma, z_vals = ctx.saved_tensors
    d_rgb, d_sigma = torch.ops.dfn.raw2outputs_backward(rgb, sigma, z_vals, d_rgb_map, d_weights, ctx.white_bg, ctx.dist_last)
    return d_rgb.reshape(rgb.shape).to(rgb.dtype), d_sigma.reshape(sigma.shape).to(sigma.dtype), None, None, None


raw2outputs.register_autograd(_r2o_backward, setup_context=_r2o_setup)


# ---- sample_pdf (scripts/train.py:31-75; the draw of :53 is an argument) ---------------------------------------------------------
@torch.library.custom_op("dfn::sample_pdf", mutates_args=(), device_types="cuda")
def sample_pdf(bins: Tensor, weights: Tensor, u: Tensor) -> Tensor:
    return ops.sample_pdf(bins, weights, u)


@sample_pdf.register_fake
def _(bins, weights, u):
    return u.new_empty(u.shape)


# ---- OptimizedVanillaNeRF.forward (scripts/model.py:173-187) on the flat parameter vector ----------------------------------------
# bf16 operand images of a parameter vector, keyed by (storage address, offset, in-place version counter): the forward and the
# backward of one autograd step - and every further call until the parameters are written again (optimiser step, load) - share
# one repack instead of two launches + 2.4 MB each.  A few entries suffice (coarse + fine network, a spare).
_PACKED = OrderedDict()
_PACKED_MAX = 4


def _packed_for(flat_params: Tensor) -> "ops.PackedMLP":
    key = (flat_params.device, flat_params.untyped_storage().data_ptr(), flat_params.storage_offset(), flat_params._version)
    pk = _PACKED.get(key)
    if pk is None:
        # a new version of the same buffer replaces its stale image (and reuses its memory)
        stale = [k for k in _PACKED if k[:3] == key[:3]]
        pk = _PACKED.pop(stale[0]) if stale else ops.PackedMLP(flat_params.device)
        for k in stale[1:]:
            _PACKED.pop(k)
        pk.pack(flat_params)
        _PACKED[key] = pk
        while len(_PACKED) > _PACKED_MAX:
            _PACKED.popitem(last=False)
    else:
        _PACKED.move_to_end(key)
    return pk


@torch.library.custom_op("dfn::vanilla_mlp", mutates_args=(), device_types="cuda")
def vanilla_mlp(flat_params: Tensor, pos: Tensor, view: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
    """(rgb [P,3], sigma [P], activation stash) of the fused tcgen05 forward; flat_params in the reference's named_parameters()
    order (595 844 floats)."""
    pk = _packed_for(flat_params)
    P = pos.shape[0]
    stash = torch.empty(ops.mlp_stash_bytes(P), dtype=torch.uint8, device=pos.device)
    rgb, sigma = ops.mlp_forward(pk, pos=pos.contiguous(), view=view.contiguous(), stash=stash)
    return rgb, sigma, stash


@vanilla_mlp.register_fake
def _(flat_params, pos, view):
    P = pos.shape[0]
    return pos.new_empty(P, 3), pos.new_empty(P), pos.new_empty(ops.mlp_stash_bytes(P), dtype=torch.uint8)


@torch.library.custom_op("dfn::vanilla_mlp_backward", mutates_args=(), device_types="cuda")
def _vanilla_mlp_backward(flat_params: Tensor, pos: Tensor, view: Tensor, rgb: Tensor, sigma: Tensor, stash: Tensor, d_rgb: Tensor,
                          d_sigma: Tensor) -> Tensor:
    pk = _packed_for(flat_params)
    P = pos.shape[0]
    grads = torch.zeros(ops.MLP_NUM_PARAMS, dtype=torch.float32, device=pos.device)
    dz = torch.empty(ops.mlp_dz_bytes(P), dtype=torch.uint8, device=pos.device)
    ops.mlp_backward(pk, rgb, sigma, d_rgb.float().contiguous(), d_sigma.float().contiguous(), stash, dz, grads, pos=pos.contiguous(),
                     view=view.contiguous())
    return grads


@_vanilla_mlp_backward.register_fake
def _(flat_params, pos, view, rgb, sigma, stash, d_rgb, d_sigma):
    return flat_params.new_empty(flat_params.shape)


def _mlp_setup(ctx, inputs, output):
    flat_params, pos, view = inputs
    rgb, sigma, stash = output
    ctx.save_for_backward(flat_params, pos, view, rgb, sigma, stash)


def _mlp_backward(ctx, d_rgb, d_sigma, d_stash):
    flat_params, pos, view, rgb, sigma, stash = ctx.saved_tensors
    d_rgb = torch.zeros_like(rgb) if d_rgb is None else d_rgb
    d_sigma = torch.zeros_like(sigma) if d_sigma is None else d_sigma
    return torch.ops.dfn.vanilla_mlp_backward(flat_params, pos, view, rgb, sigma, stash, d_rgb, d_sigma), None, None


vanilla_mlp.register_autograd(_mlp_backward, setup_context=_mlp_setup)
