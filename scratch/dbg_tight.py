import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import arpeggia_b200 as A
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
bft = A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb"))
ctx = A.Context(0)
p = A.Packed(bft)
key = ["from", "to", "label", "distance"]
a = np.sort(ctx.contacts_rows(p, 0.1, 6.5, flags=0)["rows"], order=key)
b = np.sort(ctx.contacts_rows(p, 0.1, 6.5, flags=1)["rows"], order=key)
print(len(a), len(b))
sa = set(map(tuple, a.tolist())); sb = set(map(tuple, b.tolist()))
print("only in tight:", sorted(sb - sa)[:10]); print("only in default:", sorted(sa - sb)[:10])
from collections import Counter
cb = Counter(map(tuple, b.tolist())); print("dups in tight:", [(k, v) for k, v in cb.items() if v > 1][:10])
arr = p.arrays()
for r in list(sb - sa)[:3] + [k for k, v in cb.items() if v > 1][:3]:
    f, t = r[0], r[1]
    if f < 2**31 and t < 2**31:
        print(r, "from xyz", arr["x"][f], arr["y"][f], arr["z"][f], "to xyz", arr["x"][t], arr["y"][t], arr["z"][t], "res", arr["res_id"][f], arr["res_id"][t])
