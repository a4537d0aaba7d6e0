"""cta_group::2 MMA issue-rate probe (csrc/store_bench.cu). Usage: python tools/pair_mma_bench.py"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepfields_nerf_b200 as d
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "micro"))
import build as _mb
lib = C.CDLL(_mb.LIB if os.path.exists(_mb.LIB) else _mb.build())   # tools/micro/libdfn_microbench.so, not the product library
fn = lib.dfn_pair_mma_bench; fn.restype = C.c_int
fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p]
buf = torch.empty(2 << 30, dtype=torch.uint8, device='cuda')
for grid, N, mode in ((148, 256, 0), (148, 256, 1), (148, 256, 2), (148, 256, 3), (2, 256, 1)):
    if True:
        cyc = torch.zeros(grid, dtype=torch.int64, device="cuda")
        for rep in range(2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(2000, N, grid, cyc.data_ptr(), buf.data_ptr(), buf.numel(), mode, torch.cuda.current_stream().cuda_stream)
            e1.record(); torch.cuda.synchronize()
            assert rc == 0, lib.dfn_last_error()
        c = cyc[:grid // 2].float().mean().item()
        ms = e0.elapsed_time(e1)
        tf = (grid // 2) * 2000 * 16 * 256 * N * 16 * 2 / (ms * 1e-3) / 1e12
        print(f"grid {grid} N {N} mode {mode} (1 = +STG stream, 2 = +LDG stream): {c:.1f} cycles per M256 MMA, {ms:.3f} ms, {tf:.0f} TF/s", flush=True)
