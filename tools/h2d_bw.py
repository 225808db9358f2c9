"""Pinned host -> device copy bandwidth on this box (the ceiling of bench.py's e2e leg)."""
import torch
for mb in (2, 8, 32, 256):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device="cuda")
    for ns in (1, 4):
        ss = [torch.cuda.Stream() for _ in range(ns)]
        hs = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(ns)]; ds = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(ns)]
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = max(4, 2048 // mb // ns)
        e0.record()
        for r in range(reps):
            for s, a, b in zip(ss, hs, ds):
                s.wait_event(e0) if r == 0 else None
                with torch.cuda.stream(s):
                    b.copy_(a, non_blocking=True)
        for s in ss:
            torch.cuda.current_stream().wait_stream(s)
        e1.record(); torch.cuda.synchronize()
        print(f"{mb:4d} MB x {ns} streams: {reps * ns * n / e0.elapsed_time(e1) / 1e6:.1f} GB/s")
