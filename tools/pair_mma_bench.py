"""cta_group::2 MMA issue-rate probe (csrc/store_bench.cu). Usage: python tools/pair_mma_bench.py"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepfields_nerf_b200 as d
lib = d._lib.load()
fn = lib.dfn_pair_mma_bench; fn.restype = C.c_int
fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
for grid in (2, 148):
    for N in (256, 128):
        cyc = torch.zeros(grid, dtype=torch.int64, device="cuda")
        for rep in range(2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(2000, N, grid, cyc.data_ptr(), torch.cuda.current_stream().cuda_stream)
            e1.record(); torch.cuda.synchronize()
            assert rc == 0, lib.dfn_last_error()
        c = cyc[:grid // 2].float().mean().item()
        ms = e0.elapsed_time(e1)
        tf = (grid // 2) * 2000 * 16 * 256 * N * 16 * 2 / (ms * 1e-3) / 1e12
        print(f"grid {grid} N {N}: {c:.1f} cycles per M256 MMA, {ms:.3f} ms, {tf:.0f} TF/s", flush=True)
