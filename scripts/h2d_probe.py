import torch, time
n = 3_200_000_000
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(1_500_000_000, dtype=torch.uint8, pin_memory=True); d2 = torch.empty(1_500_000_000, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for name, both in (("H2D alone", False), ("H2D with D2H", True)):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(4):
        with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
        if both:
            with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
    s1.synchronize(); t1 = time.perf_counter() - t
    torch.cuda.synchronize()
    print(name, "%.1f GB/s H2D (%.1f ms per 3.2 GB)" % (4 * n / t1 / 1e9, t1 / 4 * 1e3))
