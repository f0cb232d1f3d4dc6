"""D2H / H2D copy time by size on this box (pinned memory, CUDA events): the ceiling of the e2e path."""
import torch, time
dev = torch.device("cuda", 0)
for size in [4096, 65536, 262144, 1 << 20, 6 << 20, 7 << 20, 64 << 20]:
    d = torch.empty(size, dtype=torch.uint8, device=dev)
    h = torch.empty(size, dtype=torch.uint8).pin_memory()
    for direction in ("d2h", "h2d"):
        for _ in range(3):
            (h.copy_(d, non_blocking=True) if direction == "d2h" else d.copy_(h, non_blocking=True))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            (h.copy_(d, non_blocking=True) if direction == "d2h" else d.copy_(h, non_blocking=True))
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        t0 = time.perf_counter()
        for _ in range(reps):
            (h.copy_(d, non_blocking=True) if direction == "d2h" else d.copy_(h, non_blocking=True))
            torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e6 / reps
        print("%s %9d B  %8.1f us back-to-back (%5.1f GB/s)   %8.1f us issue+sync" % (direction, size, us, size / us / 1e3, wall))
