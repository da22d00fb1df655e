import numpy as np, arpeggia_b200 as A
from arpeggia_b200 import synth
bft = A.parse_pdb("tests/data/6bft.pdb")
ctx = A.Context(0)
p = A.Packed(synth.tile(bft, 3, 2, 2), "/", flags=2)
a = p.arrays()
print(a["n_atoms"], a["n_rings"], a["n_residues"])
res = ctx.contacts_rows(p, 0.1, 6.5)
rows = res["rows"]
ring = (rows["from"] >> 31) == 1
print("ring rows", ring.sum(), "rows", len(rows))
copy_of = np.where(ring, (rows["from"] & 0x7FFFFFFF) // 139, rows["from"] // 8180)
print(np.bincount(copy_of, minlength=12))
print(np.bincount(copy_of[ring], minlength=12), np.bincount(copy_of[~ring], minlength=12))
ck = a["chain_key"]; print(ck[:5], ck[8179], ck[8180], np.unique(ck[:8180]).size, np.unique(ck).size)
