"""Store-path microbenchmark (see csrc/store_bench.cu). Usage: python tools/store_bench.py"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import deepfields_nerf_b200 as d
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "micro"))
import build as _mb
lib = C.CDLL(_mb.LIB if os.path.exists(_mb.LIB) else _mb.build())   # tools/micro/libdfn_microbench.so, not the product library
fn = lib.dfn_store_bench; fn.restype = C.c_int
fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]
buf = torch.empty(4 << 30, dtype=torch.uint8, device="cuda")
names = {1: "STG + MMA", 129: "STG + MMA + TMA loads", 179: "STG+STS+LDTM+fence+MMA+TMA loads", 243: "STS+LDTM+fence+MMA+TMA loads (no STG)", 17: "STG + STS + MMA", 33: "STG + LDTM + MMA", 49: "STG + STS + LDTM + MMA", 51: "STG+STS+LDTM+fence+MMA",
         113: "STS + LDTM + MMA (no STG)", 115: "STS+LDTM+fence+MMA (no STG)"}
names_old = {0: "STG.128", 1: "STG.128 + MMA stream", 2: "STG.128 + fence/arrive", 3: "STG.128 + fence/arrive + MMA", 4: "smem + bulk s2g",
         5: "bulk s2g + MMA", 8: "STG.cs", 9: "STG.cs + MMA", 11: "STG.cs + fence + MMA"}
for grid in (148,):
    for mode in names:
        iters = 400
        cyc = torch.zeros(3 * grid, dtype=torch.int64, device="cuda")
        for rep in range(2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(mode, iters, grid, buf.data_ptr(), buf.numel(), cyc.data_ptr(), torch.cuda.current_stream().cuda_stream)
            e1.record(); torch.cuda.synchronize()
            assert rc == 0, lib.dfn_last_error()
        c = cyc[:grid].float().mean().item()
        mm = cyc[grid:2*grid].float().mean().item(); tl = cyc[2*grid:].float().mean().item()
        print(f"[{mm:6.1f} cyc/MMA, {tl:6.0f} cyc/16KB load] grid {grid} mode {mode:2d} {names[mode]:32s}: {c/iters:8.0f} cyc per 64 KB tile = {65536/(c/iters):5.1f} B/clk/SM; "
              f"kernel {e0.elapsed_time(e1):.3f} ms -> {grid*iters*65536/e0.elapsed_time(e1)/1e6:.0f} GB/s", flush=True)
