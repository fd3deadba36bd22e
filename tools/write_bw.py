#!/usr/bin/env python
"""Pure-write vs copy HBM bandwidth on this GPU (context for conv1, whose only significant traffic is its output)."""
import json, torch
def timeit(fn, n=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
res = {}
for gb in (1, 5):
    x = torch.empty(gb * (1 << 30) // 2, dtype=torch.float16, device="cuda")
    y = torch.empty_like(x)
    ms = timeit(lambda: x.zero_())
    res[f"memset_{gb}GB_TBps"] = round(x.numel() * 2 / ms / 1e9, 3)
    ms = timeit(lambda: x.fill_(1.5))
    res[f"fill_{gb}GB_TBps"] = round(x.numel() * 2 / ms / 1e9, 3)
    ms = timeit(lambda: y.copy_(x))
    res[f"copy_{gb}GB_TBps_rw"] = round(2 * x.numel() * 2 / ms / 1e9, 3)
    ms = timeit(lambda: torch.relu_(x))
    res[f"inplace_relu_{gb}GB_TBps_rw"] = round(2 * x.numel() * 2 / ms / 1e9, 3)
    del x, y
print(json.dumps(res))
