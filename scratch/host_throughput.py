"""Host ingest throughput: PDB reader and packer on a 1 M-atom structure; OMP_NUM_THREADS selects the threads."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import arpeggia_b200 as A
from arpeggia_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = open(os.path.join(ROOT, "tests", "data", "6bft.pdb")).read().splitlines()
lines = [l for l in src if l.startswith(("ATOM", "HETATM"))]
path = "/tmp/big_models.pdb"
if not os.path.exists(path):
    with open(path, "w") as f:
        for m in range(1, 126):
            f.write("MODEL     %4d\n" % m + "\n".join(lines) + "\nENDMDL\n")
out = {"threads": os.environ.get("OMP_NUM_THREADS", "all"), "cores": os.cpu_count()}
t = time.perf_counter(); a = A.parse_pdb(path); out["parse_s"] = time.perf_counter() - t; out["records"] = int(len(a))
cif = "/tmp/big_models.cif"
if not os.path.exists(cif):
    from pathlib import Path
    from tests.test_host_cpu import write_cif
    small = Path("/tmp/one.cif"); write_cif(A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb")), small)
    L = small.read_text().split("\n")
    rows = [l for l in L if l.startswith(("ATOM", "HETATM"))]
    first = L.index(rows[0])
    with open(cif, "w") as f:
        f.write("\n".join(L[:first]) + "\n")
        for m in range(125):
            f.write("\n".join(rows) + "\n")
        f.write("#\n")
t = time.perf_counter(); c = A.parse_cif(cif); out["parse_cif_s"] = time.perf_counter() - t; out["cif_records"] = int(len(c))
atoms = synth.assembly_1m(A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb")))
t = time.perf_counter(); p = A.Packed(atoms, "/", flags=2); out["pack_s"] = time.perf_counter() - t
out["atoms"] = int(p.structure.n_atoms)
import ctypes as C
from arpeggia_b200 import _lib
tp = C.POINTER(_lib.FrameTables)()
t = time.perf_counter(); _lib.check(_lib.lib().arph_packed_frame_tables(p._h, C.byref(tp))); out["frame_tables_s"] = time.perf_counter() - t
print(json.dumps(out))
