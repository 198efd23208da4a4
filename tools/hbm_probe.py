"""Developer probe: HBM read-only, write-only and copy bandwidth (torch kernels, CUDA events, 1 GiB buffers)."""
import torch
n = 1 << 27
a = torch.empty(n, dtype=torch.int64, device="cuda")
b = torch.empty(n, dtype=torch.int64, device="cuda")
big = torch.empty(1 << 28, dtype=torch.int64, device="cuda")  # 2 GiB: flushes the 126 MB L2 between reps
def t(fn, reps=10):
    best = 1e9
    for _ in range(reps):
        big.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
w = t(lambda: a.fill_(7))
r = t(lambda: a.sum())
c = t(lambda: b.copy_(a))
print(f"write-only {n*8/w/1e6:.0f} GB/s  read-only {n*8/r/1e6:.0f} GB/s  copy {2*n*8/c/1e6:.0f} GB/s (read+write)")
