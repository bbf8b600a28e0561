"""PCIe probe for the e2e numbers in bench.py: pinned H2D alone, D2H alone, both at once (two
streams).  Prints GB/s; the bidirectional figure bounds the pipelined e2e throughput."""
import json
import time

import torch

n_in, n_out = 86_844_958, 135_125_269
hi = torch.empty(n_in, dtype=torch.uint8).pin_memory()
ho = torch.empty(n_out, dtype=torch.uint8).pin_memory()
di = torch.empty(n_in, dtype=torch.uint8, device="cuda")
do = torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, reps=20):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                di.copy_(hi, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                ho.copy_(do, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


run(True, True, 3)
a, b, c = run(True, False), run(False, True), run(True, True)
print(json.dumps({"h2d_GBps": n_in / a / 1e9, "d2h_GBps": n_out / b / 1e9,
                  "both_ms": c * 1e3, "h2d_ms": a * 1e3, "d2h_ms": b * 1e3,
                  "both_total_GBps": (n_in + n_out) / c / 1e9}))
