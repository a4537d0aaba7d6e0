"""HBM write / read / copy bandwidth probes with stock torch kernels (context for the stash traffic). Usage: python tools/bw_probe.py"""
import torch
n = 1 << 30  # 4 GiB of fp32
x = torch.empty(n, dtype=torch.float32, device="cuda"); y = torch.empty_like(x)
def t(f, nbytes, name):
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"{name}: {ms:.3f} ms  {nbytes/ms/1e6:.0f} GB/s", flush=True)
t(lambda: x.fill_(1.0), 4 * n, "fill (write only)")
t(lambda: x.zero_(), 4 * n, "zero_ (memset)")
t(lambda: y.copy_(x), 8 * n, "copy (read+write)")
t(lambda: x.sum(), 4 * n, "sum (read only)")
xb = x.view(torch.bfloat16)
t(lambda: torch.relu_(xb), 8 * n, "relu_ in place bf16 (read+write)")
