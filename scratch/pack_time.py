import ctypes as C
import sys
import time

sys.path.insert(0, ".")
import arpeggia_b200 as A
from arpeggia_b200 import _lib, synth

bft = A.parse_pdb("tests/data/6bft.pdb")
atoms = synth.assembly_1m(bft)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    t0 = time.perf_counter()
    p = A.Packed(atoms, "/", flags=2)
    t1 = time.perf_counter()
    tp = C.POINTER(_lib.FrameTables)()
    _lib.check(_lib.lib().arph_packed_frame_tables(p._h, C.byref(tp)))
    t2 = time.perf_counter()
    print(f"pack {t1 - t0:.3f}s tables {t2 - t1:.3f}s", file=sys.stderr)
