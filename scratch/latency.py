#!/usr/bin/env python
"""Configs 1 and 2 (too small for throughput -- latency): one structure through arp_contacts, wall clock of the
call from packed host arrays to host rows, device time of the run, and the CPU restatement on the same records."""
import ctypes as C, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import arpeggia_b200 as A
from arpeggia_b200 import _lib as AL
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L = AL.lib()
ctx = A.Context(0)
out = {}
for name in ("1ubq", "6bft"):
    atoms = A.parse_pdb(os.path.join(ROOT, "tests", "data", name + ".pdb"))
    t0 = time.perf_counter(); packed = A.Packed(atoms, "/"); t_pack = time.perf_counter() - t0
    ptrs = (C.POINTER(AL.Structure) * 1)(packed.structure_ptr)
    row = {"atoms": int(packed.structure.n_atoms), "pack_ms": 1e3 * t_pack}
    for flags, label in ((0, "counting"), (1, "emit_radius")):
        prm = AL.Params(0.1, 6.5, flags, 0)
        walls = []
        for it in range(60):
            rp = C.POINTER(AL.Result)()
            t0 = time.perf_counter()
            AL.check(L.arp_contacts_batch(ctx._h, ptrs, 1, C.byref(prm), C.byref(rp)), ctx._h)
            walls.append(time.perf_counter() - t0)
            n = rp.contents.n_rows
            L.arp_result_free(ctx._h, rp)
        st = ctx.stats()
        walls = np.array(walls[10:]) * 1e3
        row[label] = {"rows": int(n), "call_ms_median": float(np.median(walls)), "call_ms_best": float(walls.min()),
                      "device_run_ms": st["ms_run"], "h2d_ms": st["ms_h2d"], "d2h_ms": st["ms_d2h"], "kernel_launches": st["kernel_launches"]}
    ws = []
    for _ in range(10):  # the first call sizes the buffers
        t0 = time.perf_counter(); tab = ctx.contacts_table(atoms, "/", 0.1, 6.5, flags=4); ws.append(1e3 * (time.perf_counter() - t0))
    row["records_to_table_ms_first_call"] = ws[0]
    row["records_to_table_ms"] = float(np.median(ws[2:]))
    row["records_to_table_split_ms"] = {k: 1e3 * tab[k] for k in ("t_setup_s", "t_classify_s", "t_table_s")}  # pack + side tables | device incl. frame | wrap
    ts = []
    for _ in range(5):
        t0 = time.perf_counter(); want = O.contacts(atoms.view(O.ATOM_DTYPE), "/", 0.1, 6.5); ts.append(time.perf_counter() - t0)
    row["cpu_port_records_to_table_ms"] = 1e3 * min(ts)
    row["table_rows"] = int(tab["n_rows"]); row["cpu_rows"] = int(want["n_rows"])
    out[name] = row
print(json.dumps(out))
