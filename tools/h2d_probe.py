import torch, time
n = 1 << 30
a = torch.empty(n, dtype=torch.float64).pin_memory(); b = torch.empty(n, dtype=torch.float64).pin_memory()
da = torch.empty(n, dtype=torch.float64, device="cuda"); db = torch.empty(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(mode):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    if mode == "one":
        with torch.cuda.stream(s1):
            da.copy_(a, non_blocking=True); db.copy_(b, non_blocking=True)
    elif mode == "two":
        with torch.cuda.stream(s1): da.copy_(a, non_blocking=True)
        with torch.cuda.stream(s2): db.copy_(b, non_blocking=True)
    elif mode == "slabs16":
        m = n // 16
        with torch.cuda.stream(s1):
            for i in range(16):
                da[i*m:(i+1)*m].copy_(a[i*m:(i+1)*m], non_blocking=True); db[i*m:(i+1)*m].copy_(b[i*m:(i+1)*m], non_blocking=True)
    elif mode == "slabs16two":
        m = n // 16
        for i in range(16):
            with torch.cuda.stream(s1): da[i*m:(i+1)*m].copy_(a[i*m:(i+1)*m], non_blocking=True)
            with torch.cuda.stream(s2): db[i*m:(i+1)*m].copy_(b[i*m:(i+1)*m], non_blocking=True)
    torch.cuda.synchronize(); return time.perf_counter() - t0
for mode in ("one", "two", "slabs16", "slabs16two", "one", "two"):
    run(mode); dt = min(run(mode) for _ in range(3))
    print(mode, round(dt*1e3, 1), "ms", round(16 * n / dt / 1e9, 1), "GB/s")
