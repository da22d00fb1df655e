"""Latency of one structure through the C ABI (configs 1 and 2): staged run (device-resident), one-shot arp_contacts, and
the full table call."""
import time, json, sys
import numpy as np
import arpeggia_b200 as A

ctx = A.Context(0)
out = {}
for name in ("1ubq", "6bft"):
    atoms = A.parse_pdb(f"tests/data/{name}.pdb")
    p = A.Packed(atoms)
    ctx.upload(p)
    for _ in range(20):
        ctx.run(0.1, 6.5); ctx.stats()
    ms = []
    for _ in range(200):
        ctx.run(0.1, 6.5)
        ms.append(ctx.stats()["ms_run"])
    t0 = time.perf_counter()
    for _ in range(200):
        ctx.run(0.1, 6.5)
        ctx.download()
    wall_staged = (time.perf_counter() - t0) / 200 * 1e3
    t0 = time.perf_counter()
    for _ in range(200):
        ctx.contacts_rows(p, 0.1, 6.5)
    wall_oneshot = (time.perf_counter() - t0) / 200 * 1e3
    t0 = time.perf_counter()
    for _ in range(50):
        ctx.contacts_table(atoms, "/", 0.1, 6.5)
    wall_table = (time.perf_counter() - t0) / 50 * 1e3
    out[name] = {"atoms": len(atoms), "ms_run_device_median": float(np.median(ms)), "ms_run_device_min": float(np.min(ms)),
                 "ms_wall_run_plus_download": wall_staged, "ms_wall_arp_contacts": wall_oneshot, "ms_wall_arph_contacts_table": wall_table}
print(json.dumps(out))
