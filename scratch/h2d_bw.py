"""Raw pinned host->device bandwidth of this box (the ceiling of bench.py's e2e leg)."""
import torch, time
for gb in (1, 4):
    n = gb * (1 << 30) // 8
    h = torch.empty(n, dtype=torch.float64).pin_memory()
    d = torch.empty(n, dtype=torch.float64, device="cuda")
    for nstream in (1, 2, 4):
        streams = [torch.cuda.Stream() for _ in range(nstream)]
        chunk = n // nstream
        best = 1e9
        for _ in range(3):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for i, s in enumerate(streams):
                with torch.cuda.stream(s):
                    d[i * chunk:(i + 1) * chunk].copy_(h[i * chunk:(i + 1) * chunk], non_blocking=True)
            torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
        print("H2D %d GB over %d stream(s): %.1f GB/s" % (gb, nstream, gb * 1.073741824 / best))
    best = 1e9
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        h.copy_(d, non_blocking=True)
        torch.cuda.synchronize(); best = min(best, time.perf_counter() - t0)
    print("D2H %d GB: %.1f GB/s" % (gb, gb * 1.073741824 / best))
