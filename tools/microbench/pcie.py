import torch, time
n = 300_000_000
h_in = torch.empty(n, dtype=torch.int64).pin_memory()
h_out = torch.empty(n, dtype=torch.int64).pin_memory()
d1 = torch.empty(n, dtype=torch.int64, device="cuda")
d2 = torch.empty(n, dtype=torch.int64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=3):
    torch.cuda.synchronize(); best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    return best
gb = n * 8 / 1e9
a = t(lambda: d1.copy_(h_in, non_blocking=True))
b = t(lambda: h_out.copy_(d2, non_blocking=True))
def both():
    with torch.cuda.stream(s1): d1.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2): h_out.copy_(d2, non_blocking=True)
c = t(both)
print(f"H2D {gb/a:.1f} GB/s  D2H {gb/b:.1f} GB/s  both at once: {gb/c:.1f} GB/s each way ({c*1e3:.1f} ms for {gb:.1f} GB each)")
