#!/usr/bin/env python3
"""HBM bandwidth probes that put the measured copy peak (MEASURED_PEAKS.json, 50/50 read/write) next to the traffic mix
of the unprojection kernel (24 % read / 76 % write): pure write (fill_), copy, and a 1:3 read:write expand."""
import json, torch
def t(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(it):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e-3
n = 1 << 29
a = torch.empty(n, dtype=torch.float32, device="cuda"); b = torch.empty(n, dtype=torch.float32, device="cuda")
out = {}
out["fill_gbs"] = 4 * n / t(lambda: a.fill_(1.0)) / 1e9
out["copy_gbs"] = 8 * n / t(lambda: b.copy_(a)) / 1e9
s = a[: n // 3]
big = b[: 3 * (n // 3)].view(n // 3, 3)
out["expand_1r3w_gbs"] = (4 * (n // 3) * 4) / t(lambda: big.copy_(s[:, None].expand(-1, 3))) / 1e9
out["read_sum_gbs"] = 4 * n / t(lambda: a.sum()) / 1e9
print(json.dumps(out))
