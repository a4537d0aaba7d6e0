"""CPU restatement (numpy) of the Instant-NGP variant, scripts/instantngp.py:7-104.  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg, never by the product path.

PARITY UNPINNED.  The arithmetic of this variant lives in tiny-cuda-nn (un-vendored submodule pinned at commit
74ea173d612fe576965f36c2b971a2bbb80af091, CUDA-only, absent here and on the GPU box) and the reference holds no test or
golden vector for it.  What follows restates tcnn's published algorithm:
  * grid_scale / grid_resolution / level offsets / pos_fract / grid_index / coherent_prime_hash
    (include/tiny-cuda-nn/common_device.h, encodings/grid.h),
  * SphericalHarmonics degree 4 (encodings/spherical_harmonics.h; inputs in [0,1] mapped to [-1,1]),
  * FullyFusedMLP (networks/fully_fused_mlp.cu): no biases, fp16 weights row-major [out,in] per layer, output padded to
    16 rows, fp16 activations between layers,
anchored on the reference's call sites instantngp.py:23-64 (configs) and :77-79, :98-102 (use).  Stated deviation: the
interpolation and the matrix products accumulate in fp32 and round to fp16 once (tcnn accumulates in fp16).
"""
from __future__ import annotations

import numpy as np

N_LEVELS = 16
PRIMES = (np.uint32(1), np.uint32(2654435761), np.uint32(805459861))
LOG2_SIZE = {"small": 15, "medium": 17, "large": 19}  # instantngp.py:13-19


def level_table(log2_hashmap_size: int):
    """instantngp.py:23-33 config through tcnn grid_scale / grid_resolution / the offset table of GridEncoding's ctor."""
    f32 = np.float32
    scale = np.zeros(N_LEVELS, f32)
    res = np.zeros(N_LEVELS, np.uint32)
    size = np.zeros(N_LEVELS, np.uint32)
    offset = np.zeros(N_LEVELS, np.uint32)
    off = 0
    for l in range(N_LEVELS):
        # tcnn: exp2(l * log2(1.5)) * 16 - 1 with the device's exp2f; 16 * 1.5**l - 1 is exact in fp32 for l < 16
        s = f32(16.0 * 1.5 ** l - 1.0)
        r = int(np.ceil(s)) + 1
        n = (r ** 3 + 7) // 8 * 8
        n = min(n, 1 << log2_hashmap_size)
        scale[l], res[l], size[l], offset[l] = s, r, n, off
        off += n
    return scale, res, size, offset, off


def _fma32(a, b, c):
    """fp32 fma: exact in float64 for fp32 a, b (48-bit product) then one rounding."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def _locate(pos, scale):
    x = _fma32(np.full_like(pos, scale), pos, np.full_like(pos, 0.5))
    fl = np.floor(x)
    return fl.astype(np.int32).astype(np.uint32), (x - fl).astype(np.float32)


def _grid_index(c, res, size):
    """tcnn grid_index<3>; c uint32 [P,3]."""
    res, size = int(res), int(size)
    with np.errstate(over="ignore"):
        if res ** 3 > size:
            idx = (c[:, 0] * PRIMES[0]) ^ (c[:, 1] * PRIMES[1]) ^ (c[:, 2] * PRIMES[2])
        else:
            idx = c[:, 0] + c[:, 1] * np.uint32(res) + c[:, 2] * np.uint32(res * res)
    return (idx % np.uint32(size)).astype(np.int64)


def _corners(pos, l, tables):
    scale, res, size, offset, _ = tables
    cell, frac = _locate(pos, scale[l])
    for idx in range(8):
        bits = np.array([idx & 1, (idx >> 1) & 1, (idx >> 2) & 1], np.uint32)
        with np.errstate(over="ignore"):
            c = cell + bits[None, :]
        w = np.where(bits[None, :] == 1, frac, np.float32(1.0) - frac).astype(np.float32)
        weight = ((w[:, 0] * w[:, 1]).astype(np.float32) * w[:, 2]).astype(np.float32)
        yield int(offset[l]) + _grid_index(c, res[l], size[l]), weight


def hashgrid_encode(pos, table_f16, log2_hashmap_size):
    """pos [P,3] fp32, table fp16 [n_entries,2] -> fp16 [P,32], feature = 2*level + f."""
    pos = np.asarray(pos, np.float32)
    tables = level_table(log2_hashmap_size)
    out = np.zeros((pos.shape[0], 2 * N_LEVELS), np.float16)
    for l in range(N_LEVELS):
        acc = np.zeros((pos.shape[0], 2), np.float32)
        for j, w in _corners(pos, l, tables):
            v = table_f16[j].astype(np.float32)
            acc = _fma32(np.repeat(w[:, None], 2, 1), v, acc)
        out[:, 2 * l:2 * l + 2] = acc.astype(np.float16)
    return out


def hashgrid_backward(pos, d_out_f16, log2_hashmap_size, grad_scale=1.0):
    """scatter-add of d_out [P,32] into a float64 [n_entries,2] table gradient (order-free reference sum)."""
    pos = np.asarray(pos, np.float32)
    tables = level_table(log2_hashmap_size)
    g = np.zeros((tables[4], 2), np.float64)
    for l in range(N_LEVELS):
        d = d_out_f16[:, 2 * l:2 * l + 2].astype(np.float64) * grad_scale
        for j, w in _corners(pos, l, tables):
            np.add.at(g, j, w[:, None].astype(np.float64) * d)
    return g


def sh4(dirs):
    """tcnn SphericalHarmonics degree 4: dirs [P,3] fp32 in [0,1] -> fp16 [P,16]."""
    d = np.asarray(dirs, np.float32) * np.float32(2) - np.float32(1)
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
    f = np.float32
    o = np.stack([
        np.full_like(x, 0.28209479177387814), f(-0.48860251190291987) * y, f(0.48860251190291987) * z,
        f(-0.48860251190291987) * x, f(1.0925484305920792) * xy, f(-1.0925484305920792) * yz,
        f(0.94617469575755997) * z2 - f(0.31539156525251999), f(-1.0925484305920792) * xz,
        f(0.54627421529603959) * x2 - f(0.54627421529603959) * y2, f(0.59004358992664352) * y * (f(-3) * x2 + y2),
        f(2.8906114426405538) * xy * z, f(0.45704579946446572) * y * (f(1) - f(5) * z2),
        f(0.3731763325901154) * z * (f(5) * z2 - f(3)), f(0.45704579946446572) * x * (f(1) - f(5) * z2),
        f(1.4453057213202769) * z * (x2 - y2), f(0.59004358992664352) * x * (-x2 + f(3) * y2)], 1)
    return o.astype(np.float16)


def _dense(x_f16, w_f16):
    return x_f16.astype(np.float32) @ w_f16.astype(np.float32).T


def split_dmlp(w):
    return w[:2048].reshape(64, 32), w[2048:3072].reshape(16, 64)


def split_cmlp(w):
    return w[:2048].reshape(64, 32), w[2048:6144].reshape(64, 64), w[6144:7168].reshape(16, 64)


def query_density(pos, table_f16, dmlp_w_f16, log2_hashmap_size):
    """instantngp.py:66-80: features fp16 [P,16] (ReLU output activation), sigma = features[:, :1]."""
    w0, w1 = split_dmlp(dmlp_w_f16)
    enc = hashgrid_encode(pos, table_f16, log2_hashmap_size)
    h = np.maximum(_dense(enc, w0), 0).astype(np.float16)
    feat = np.maximum(_dense(h, w1), 0).astype(np.float16)
    return feat[:, :1], feat


def forward(pos, dirs, table_f16, dmlp_w_f16, cmlp_w_f16, log2_hashmap_size):
    """instantngp.py:82-104 -> (rgb fp16 [P,3], sigma fp16 [P,1])."""
    sigma, feat = query_density(pos, table_f16, dmlp_w_f16, log2_hashmap_size)
    w0, w1, w2 = split_cmlp(cmlp_w_f16)
    x = np.concatenate([feat, sh4(dirs)], 1)
    h = np.maximum(_dense(x, w0), 0).astype(np.float16)
    h = np.maximum(_dense(h, w1), 0).astype(np.float16)
    o = _dense(h, w2)[:, :3]
    rgb = (np.float32(1) / (np.float32(1) + np.exp(-o))).astype(np.float16)
    return rgb, sigma
