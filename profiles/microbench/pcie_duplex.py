"""PCIe microbenchmark for the e2e arm: H2D alone, D2H alone, both at once on two streams (pinned host memory).
Run on the GPU box: python profiles/microbench/pcie_duplex.py"""
import torch

gb = 8
n = gb * (1 << 30)
hin = torch.empty(n, dtype=torch.uint8).pin_memory()
hout = torch.empty(n, dtype=torch.uint8).pin_memory()
din = torch.empty(n, dtype=torch.uint8, device="cuda")
dout = torch.zeros(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    fn()
    s1.synchronize(); s2.synchronize()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b)


def h2d():
    with torch.cuda.stream(s1):
        din.copy_(hin, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        hout.copy_(dout, non_blocking=True)


for name, fn, nbytes in (("h2d", h2d, n), ("d2h", d2h, n), ("h2d", h2d, n), ("d2h", d2h, n), ("both", lambda: (h2d(), d2h()), 2 * n), ("both", lambda: (h2d(), d2h()), 2 * n)):
    ms = timed(fn)
    print(f"{name}: {gb} GiB per direction in {ms:.1f} ms = {nbytes / ms / 1e6:.1f} GB/s aggregate")
