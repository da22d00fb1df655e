"""Config 4: a batch of jittered ~5 k-atom complexes (chains A, B, C, G of 6bft) through arp_contacts_batch on one GPU."""
import sys, time, json
import numpy as np
import arpeggia_b200 as A
from arpeggia_b200 import synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
bft = A.parse_pdb("tests/data/6bft.pdb")
t0 = time.perf_counter()
packs = [A.Packed(synth.jittered_complex(bft, k), "/", flags=2) for k in range(n)]
t_pack = time.perf_counter() - t0
ctx = A.Context(0)
t0 = time.perf_counter(); ctx.upload(packs); t_up = time.perf_counter() - t0
for _ in range(3):
    ctx.run(0.1, 6.5, 0)
    st = ctx.stats()
res = ctx.download()
st = ctx.stats()
parts = A.split_rows(res)
out = {"structures": n, "atoms": int(st["n_atoms"]), "rows": int(len(res["rows"])), "pairs_in_cutoff": int(res["pairs_in_cutoff"]),
       "ms_run": st["ms_run"], "ms_grid": st["ms_grid"], "ms_pairs": st["ms_pairs"], "cells": int(st["n_cells"]),
       "pairs_per_s": res["pairs_in_cutoff"] / (st["ms_run"] * 1e-3), "structures_per_s": n / (st["ms_run"] * 1e-3),
       "s_pack": t_pack, "s_upload": t_up}
# property: a structure's rows in the batch are the rows of the structure alone
key = ["from", "to", "label"]
rng = np.random.default_rng(4)
for k in rng.choice(n, size=min(n, 8), replace=False):
    single = ctx.contacts_rows(packs[int(k)], 0.1, 6.5)["rows"]
    assert np.array_equal(np.sort(parts[int(k)], order=key), np.sort(single, order=key)), k
out["spot_checked"] = min(n, 8)
ctx.upload(packs)
for _ in range(3):
    ctx.run(0.1, 6.5, 1)
    st1 = ctx.stats()
res1 = ctx.download()
k3 = ["from", "to", "label", "distance"]
out["emit_radius"] = {"ms_run": st1["ms_run"], "ms_grid": st1["ms_grid"], "ms_pairs": st1["ms_pairs"], "cells": int(st1["n_cells"]),
                      "rows": int(len(res1["rows"])), "pairs_per_s": res["pairs_in_cutoff"] / (st1["ms_run"] * 1e-3),
                      "structures_per_s": n / (st1["ms_run"] * 1e-3),
                      "rows_identical": bool(np.array_equal(np.sort(res1["rows"], order=k3), np.sort(res["rows"], order=k3)))}
print(json.dumps(out))
