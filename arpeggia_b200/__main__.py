"""`python -m arpeggia_b200 contacts ...`: the flag surface of `arpeggia contacts` (src/cli/contacts.rs:9-52)."""
from __future__ import annotations

import argparse
import logging
import os
import sys


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="arpeggia")
    sub = ap.add_subparsers(dest="cmd", required=True)
    c = sub.add_parser("contacts", help="Analyze atomic and ring contacts in a PDB or mmCIF file")
    c.add_argument("-i", "--input", required=True, help="Path to the PDB or mmCIF file to be analyzed")
    c.add_argument("-o", "--output", required=True, help="Output directory")
    c.add_argument("-g", "--groups", default="/", help="Group chains for interactions, e.g. A,B/C,D")
    c.add_argument("-f", "--filename", default="contacts", help="Name of the output file")
    c.add_argument("-t", "--output-format", default="csv", choices=["csv", "parquet", "json", "ndjson"])
    c.add_argument("-c", "--vdw-comp", type=float, default=0.1)
    c.add_argument("-d", "--dist-cutoff", type=float, default=6.5)
    c.add_argument("-j", "--num-threads", type=int, default=1)
    c.add_argument("--ignore-zero-occupancy", action="store_true")
    c.add_argument("--device", type=int, default=0, help="CUDA device ordinal")
    a = ap.parse_args(argv)

    logging.basicConfig(level=os.environ.get("RUST_LOG", "info").upper(), format="%(levelname)s %(message)s")
    log = logging.getLogger("arpeggia")
    from . import contacts, load_atoms
    from .frame import write_table

    inp = os.path.realpath(a.input)
    if not os.path.exists(inp):  # cli/contacts.rs:58-64: log and return
        log.error("Failed to retrieve input file: %s", a.input)
        return 0
    out_dir = os.path.abspath(a.output)
    atoms = load_atoms(inp)
    if not (atoms["element"] == b"H").any():  # cli/contacts.rs:92-101
        log.warning("No hydrogen atoms found in the structure. This may affect the accuracy of the results.")
    df = contacts(inp, a.groups, a.vdw_comp, a.dist_cutoff, a.ignore_zero_occupancy, a.num_threads, device=a.device)
    os.makedirs(out_dir, exist_ok=True)
    col = df["interaction"]
    if hasattr(col, "to_pylist"):  # pyarrow: counted in Arrow's C++, no Python list of a million strings
        import pyarrow.compute as pc

        n_clash = int(pc.sum(pc.equal(col, "StericClash")).as_py() or 0)
    else:  # polars
        n_clash = int((col == "StericClash").sum())
    if n_clash:  # cli/contacts.rs:116-132
        log.warning("Found %d steric %s", n_clash, "clash" if n_clash == 1 else "clashes")
    # Path::with_extension (cli/contacts.rs:111-113): an extension already in the name is REPLACED, not appended to
    path = os.path.join(out_dir, os.path.splitext(a.filename)[0] + "." + a.output_format)
    write_table(df, path, a.output_format)
    log.info("Results saved to %s", path)
    return 0


if __name__ == "__main__":
    sys.exit(main())
