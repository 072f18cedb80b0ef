#!/usr/bin/env python
"""Context for the roofline: what plain library kernels reach on this GPU for (a) a streaming read, (b) a copy,
(c) a random gather of 4 KB rows.  Prints one JSON line.  Not part of the product or of bench.py."""
import json
import torch

dev = torch.device("cuda", 0)
n = 4 << 30  # 4 Gi floats = 16 GB
x = torch.empty(n, dtype=torch.float32, device=dev).normal_()


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best / 1e3


out = {}
out["read_sum_gbs"] = n * 4 / timed(lambda: x.sum()) / 1e9
y = torch.empty(n // 2, dtype=torch.float32, device=dev)
out["copy_gbs"] = 2 * (n // 2) * 4 / timed(lambda: y.copy_(x[: n // 2])) / 1e9
rows = x.view(-1, 1024)  # 4 KB rows
idx = torch.randint(0, rows.shape[0], (1 << 20,), device=dev)
dst = torch.empty((1 << 20, 1024), dtype=torch.float32, device=dev)
t = timed(lambda: torch.index_select(rows, 0, idx, out=dst))
out["gather_4k_rows_read_gbs"] = (1 << 20) * 4096 / t / 1e9
out["gather_4k_rows_read_plus_write_gbs"] = 2 * (1 << 20) * 4096 / t / 1e9
print(json.dumps(out))
