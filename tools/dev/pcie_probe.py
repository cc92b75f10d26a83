"""Developer probe: pinned H2D / D2H copy time for the e2e step's transfer sizes."""
import time, torch
for nbytes, kind in ((393216, "h2d"), (1228800, "d2h"), (1064960, "d2h"), (32768, "d2h")):
    h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep in range(3):
        torch.cuda.synchronize()
        e0.record()
        for _ in range(50):
            if kind == "h2d": d.copy_(h, non_blocking=True)
            else: h.copy_(d, non_blocking=True)
        e1.record(); torch.cuda.synchronize()
    dev_us = e0.elapsed_time(e1) * 1e3 / 50
    t0 = time.perf_counter()
    for _ in range(50):
        if kind == "h2d": d.copy_(h, non_blocking=True)
        else: h.copy_(d, non_blocking=True)
        torch.cuda.synchronize()
    wall_us = (time.perf_counter() - t0) * 1e6 / 50
    print(f"{kind} {nbytes} B: {dev_us:.1f} us back-to-back on the device ({nbytes / dev_us / 1e3:.1f} GB/s), {wall_us:.1f} us copy+sync wall")
