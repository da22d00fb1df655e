"""Regenerates tests/golden/1ubq_contacts.csv: the default output of `arpeggia contacts -i 1ubq.pdb` as this repository
writes it -- the ORACLE's table (oracle/, pinned by the reference's own golden values: 532 rows) through the writer of the
CLI (arpeggia_b200.frame.write_table, csv).  The reference itself cannot run here (Rust), so the file pins the table's
content and this repository's CSV formatting (header, column order, quoting only where needed, nulls as empty fields);
polars' float formatting of the reference's own CsvWriter is not pinned by it.      python tests/golden/make_golden.py"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

import arpeggia_b200 as A  # noqa: E402
from arpeggia_b200.frame import to_arrow, write_table  # noqa: E402
from oracle import oracle as O  # noqa: E402


def golden_table():
    atoms = A.parse_pdb(ROOT / "tests" / "data" / "1ubq.pdb")
    return to_arrow(O.contacts(atoms.view(O.ATOM_DTYPE)))


if __name__ == "__main__":
    out = ROOT / "tests" / "golden" / "1ubq_contacts.csv"
    write_table(golden_table(), out, "csv")
    print(out, out.stat().st_size, "bytes")
