"""PCIe copy rates on the GPU box: what an arp_contacts call can hope for (scratch measurement, torch for plumbing only)."""
import json
import time

import torch

dev = torch.device("cuda:0")
res = {}


def timed(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return best


for mb in (1, 4, 8, 20, 64):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    t = timed(lambda: d.copy_(h, non_blocking=True))
    res[f"h2d_{mb}MB_GBps"] = n / t / 1e9
    t = timed(lambda: h.copy_(d, non_blocking=True))
    res[f"d2h_{mb}MB_GBps"] = n / t / 1e9

# 20 MB as 5 arrays of 4 MB: one stream, two streams, four streams
parts_h = [torch.empty(4 << 20, dtype=torch.uint8).pin_memory() for _ in range(5)]
parts_d = [torch.empty(4 << 20, dtype=torch.uint8, device=dev) for _ in range(5)]
streams = [torch.cuda.Stream() for _ in range(4)]


def multi(ns):
    def f():
        for k, (hh, dd) in enumerate(zip(parts_h, parts_d)):
            with torch.cuda.stream(streams[k % ns]):
                dd.copy_(hh, non_blocking=True)
    return f


for ns in (1, 2, 4):
    res[f"h2d_5x4MB_{ns}streams_GBps"] = 5 * (4 << 20) / timed(multi(ns)) / 1e9

# both directions at once: 20 MB up on one stream, 14 MB down on another
h_up = torch.empty(20 << 20, dtype=torch.uint8).pin_memory()
d_up = torch.empty(20 << 20, dtype=torch.uint8, device=dev)
h_dn = torch.empty(14 << 20, dtype=torch.uint8).pin_memory()
d_dn = torch.empty(14 << 20, dtype=torch.uint8, device=dev)


def both():
    with torch.cuda.stream(streams[0]):
        d_up.copy_(h_up, non_blocking=True)
    with torch.cuda.stream(streams[1]):
        h_dn.copy_(d_dn, non_blocking=True)


res["up20_down14_together_ms"] = 1e3 * timed(both)
res["up20_alone_ms"] = 1e3 * timed(lambda: d_up.copy_(h_up, non_blocking=True))
res["down14_alone_ms"] = 1e3 * timed(lambda: h_dn.copy_(d_dn, non_blocking=True))
# write-combined / numa: which node is the pinned memory on is not controllable from here; report the cpu affinity
import os
res["affinity"] = sorted(os.sched_getaffinity(0))[:64]
res["props"] = str(torch.cuda.get_device_properties(0))
print(json.dumps(res, indent=1))
