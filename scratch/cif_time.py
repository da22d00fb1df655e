import os
import sys
import tempfile
import time

sys.path.insert(0, ".")
import arpeggia_b200 as A
from arpeggia_b200 import synth

bft = A.parse_pdb("tests/data/6bft.pdb")
atoms = synth.assembly_1m(bft)
with tempfile.TemporaryDirectory() as td:
    path = os.path.join(td, "c3.cif")
    synth.write_cif(atoms, path)
    print("bytes", os.path.getsize(path), file=sys.stderr)
    for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
        t0 = time.perf_counter()
        a = A.parse_cif(path)
        t1 = time.perf_counter()
        print(f"parse_cif {t1 - t0:.3f}s {len(a)} records", file=sys.stderr)
