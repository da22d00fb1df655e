"""How long does the frame's column download take, call after call? (scratch)"""
import sys
sys.path.insert(0, ".")
import arpeggia_b200 as A
from arpeggia_b200 import synth

bft = A.parse_pdb("tests/data/6bft.pdb")
atoms = synth.assembly_1m(bft)
packed = A.Packed(atoms, "/", flags=2)
ctx = A.Context(0)
ctx.upload(packed)
ctx.run(0.1, 6.5, 1)
for k in range(8):
    fr = ctx.frame(packed, device_columns=(k % 2 == 1))
    fs = fr["frame_stats"]
    print(k, "device columns" if k % 2 else "compact", {q: round(v, 3) for q, v in fs.items() if q.startswith("ms_")}, fr["n_rows"])
