"""tcgen05 issue-rate / TMEM-drain microbenchmarks.  Usage: python tools/umma_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepfields_nerf_b200 as d
import ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "micro"))
import build as _mb
_mlib = C.CDLL(_mb.LIB if os.path.exists(_mb.LIB) else _mb.build())
_mlib.dfn_umma_bench.restype = C.c_int
_mlib.dfn_umma_bench.argtypes = [C.c_int] * 6 + [C.c_void_p, C.c_void_p]


class _lib:  # the probes live in tools/micro/libdfn_microbench.so, not in the product library
    @staticmethod
    def call(name, *args):
        rc = getattr(_mlib, name)(*args)
        if rc != 0:
            raise RuntimeError(f"{name} failed ({rc})")

def run(mode, iters, group, depth, N, grid):
    out = torch.zeros(grid, dtype=torch.int64, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _lib.call("dfn_umma_bench", mode, 8, group, depth, N, grid, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    e0.record()
    _lib.call("dfn_umma_bench", mode, iters, group, depth, N, grid, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    e1.record()
    torch.cuda.synchronize()
    cyc = out.float().mean().item()
    return cyc, e0.elapsed_time(e1)

iters = 2000
for grid in (1, 148):
    for mode, N in ((8, 256), (8, 64), (9, 256), (9, 64)):
        for group, depth in ((2, 2),):
            cyc, ms = run(mode, iters, group, depth, N, grid)
            per = cyc / (iters * 16)
            tf = grid * iters * 16 * 128 * N * 16 * 2 / (ms * 1e-3) / 1e12
            print(f"grid {grid:3d} mode {mode} N {N:3d} group {group:2d} depth {depth}: {per:7.1f} cyc/MMA  {ms:8.3f} ms  {tf:7.1f} TF/s", flush=True)
    cyc, ms = run(1, iters, 1, 1, 256, grid)
    print(f"grid {grid:3d} drain: {cyc / iters:7.1f} cyc per 128x256 fp32 tile (8 warps)  {ms:.3f} ms", flush=True)
    cyc, ms = run(3, iters, 8, 2, 256, grid)
    print(f"grid {grid:3d} mma+drain: {cyc / (iters * 16):7.1f} cyc/MMA  {ms:.3f} ms", flush=True)
