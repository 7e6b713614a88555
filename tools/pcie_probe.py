"""Raw pinned-memory PCIe throughput of the box (context for the bench's e2e leg): H2D, D2H, and both at once."""
import torch

n = 64 << 20
h_a = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_b = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    s1.synchronize(); s2.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def both():
    with torch.cuda.stream(s1):
        d_a.copy_(h_a, non_blocking=True)
    with torch.cuda.stream(s2):
        h_b.copy_(d_b, non_blocking=True)


t = timed(lambda: d_a.copy_(h_a, non_blocking=True))
print(f"H2D {n / t / 1e6:.1f} GB/s")
t = timed(lambda: h_b.copy_(d_b, non_blocking=True))
print(f"D2H {n / t / 1e6:.1f} GB/s")
import time
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10):
    both()
torch.cuda.synchronize(); t = (time.perf_counter() - t0) / 10 * 1e3
print(f"H2D+D2H concurrently: {n / t / 1e6:.1f} GB/s each direction")
