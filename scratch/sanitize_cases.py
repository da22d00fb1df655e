"""The cases compute-sanitizer runs over (SURVEY 5 "Race detection"): configs 1 and 2 in both screen
modes with the frame, a dense-overlap structure (crowded cells, row-buffer growth), a protonated structure (hydrogen
lists), a batch job, the neighbour sum (staged and static), and a two-slab exchange on one GPU (peer-copy transport).
Every result is checked against the oracle, so a sanitizer run is also a parity run.
usage: compute-sanitizer --tool <memcheck|racecheck|initcheck|synccheck> python scratch/sanitize_cases.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import arpeggia_b200 as A  # noqa: E402
from arpeggia_b200 import core, distributed as D, sap, synth  # noqa: E402
from oracle import oracle as O  # noqa: E402
from tests import fixtures as F  # noqa: E402
from tests.util import assert_tables_equal  # noqa: E402

os.environ.setdefault("ARP_NO_GRAPH", "1")  # every kernel as its own launch
ubq = A.parse_pdb(os.path.join(ROOT, "tests", "data", "1ubq.pdb"))
bft = A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb"))
ctx = A.Context(0)
KEY = ["from", "to", "label", "distance"]
done = []

for name, atoms in (("1ubq", ubq), ("6bft", bft)):
    want = O.contacts(atoms.view(O.ATOM_DTYPE), "/", 0.1, 6.5, num_threads=0)
    for flags in (0, 2, 4):  # counting mode with both counters / counting mode / emit-radius screen; frame on the device
        assert_tables_equal(ctx.contacts_table(atoms, "/", 0.1, 6.5, flags=flags), want, f"{name} flags {flags}")
    done.append(f"{name}: 3 modes + frame")

dense = synth.tile(bft, 2, 1, 1, pitch=(20.0, 15.0, 10.0))
want = O.contacts(dense.view(O.ATOM_DTYPE), "/", 0.1, 6.5, flags=O.FLAG_SKIP_LOAD_FILTER, num_threads=0)
for flags in (2, 2 | 4):
    assert_tables_equal(ctx.contacts_table(dense, "/", 0.1, 6.5, flags=flags), want, f"dense flags {flags}")
done.append("dense overlap: 2 modes")

prot = F.all_cases(ubq.view(O.ATOM_DTYPE))["protonated"].view(A.ATOM_DTYPE)
want = O.contacts(prot.view(O.ATOM_DTYPE), "/", 0.1, 6.5, num_threads=0)
for flags in (0, 4):
    assert_tables_equal(ctx.contacts_table(prot, "/", 0.1, 6.5, flags=flags), want, f"protonated flags {flags}")
done.append("protonated: 2 modes")

packs = [A.Packed(synth.jittered_complex(bft, k), "/", flags=2) for k in range(3)] + [A.Packed(ubq)]
for flags in (0, 1):
    parts = A.split_rows(ctx.contacts_rows(packs, 0.1, 6.5, flags))
    for p, part in zip(packs, parts):
        single = ctx.contacts_rows(p, 0.1, 6.5, flags)["rows"]
        assert np.array_equal(np.sort(part, order=KEY), np.sort(single, order=KEY))
done.append("batch of 4: 2 modes")

atoms = sap.prepare_atoms(bft)
pk = A.Packed(atoms, "/", flags=2)
a = pk.arrays()
sel = sap.sidechain_mask(atoms)
w = np.random.default_rng(1).integers(-4, 5, size=len(atoms)).astype(np.float32)
ctx.upload(pk)
got = ctx.neighbor_sum(w, sel, 5.0)
want = O.neighbor_sum(a["x"], a["y"], a["z"], w, sel, 5.0)
assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
import torch  # noqa: E402  (device arrays for the entry point that allocates nothing)
import ctypes as C  # noqa: E402
t = lambda v, dt: torch.from_numpy(np.ascontiguousarray(v, dtype=dt)).cuda()
dx, dy, dz, dw, ds = t(a["x"], np.float64), t(a["y"], np.float64), t(a["z"], np.float64), t(w, np.float32), t(sel.astype(np.uint8), np.uint8)
out = torch.zeros(len(w), dtype=torch.float32, device="cuda")
torch.cuda.synchronize()
for _ in range(2):
    core.check(core.lib().arp_neighbor_sum_static(0, None, dx.data_ptr(), dy.data_ptr(), dz.data_ptr(), dw.data_ptr(), ds.data_ptr(), len(w),
                                                  C.c_float(5.0), out.data_ptr()))
assert np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32))
done.append("neighbour sum: staged + static")

whole = A.Packed(bft).arrays()
single = ctx.contacts_rows(A.Packed(bft), 0.1, 6.5)["rows"]
ctxs = core.create_multi([0, 0])
for r, cx in enumerate(ctxs):
    cx.upload([D.ArrayStructure(D.cut_slab(whole, r, 2))])
for flags in (0, 1):
    core.slab_exchange_multi(ctxs, 0.1, 6.5)
    rows = []
    for cx in ctxs:
        cx.run(0.1, 6.5, flags)
        rows.append(cx.download()["rows"])
    assert np.array_equal(np.sort(np.concatenate(rows), order=KEY), np.sort(single, order=KEY))
for cx in ctxs:
    cx.close()
done.append("two slabs, halo exchange: 2 modes")
ctx.close()
print("sanitize_cases ok:", "; ".join(done))
