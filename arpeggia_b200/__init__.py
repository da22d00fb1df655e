"""arpeggia_b200: the `contacts` hot path of y1zhou/arpeggia, rebuilt for B200 (sm_100a).

Drop-in for the reference's Python surface of that path (python/arpeggia/__init__.py:58-69,
src/python.rs:31-56): `contacts(input_file, groups, vdw_comp, dist_cutoff, ignore_zero_occupancy, num_threads)`
returns the same 20-column, identically sorted frame -- a polars.DataFrame when polars is importable,
otherwise a pyarrow.Table with the same schema.  The search and classification run on the GPU through the
C ABI in include/arpeggia_cuda.h; there is no CPU fallback.
"""
from __future__ import annotations

from ._lib import ArpeggiaError, ATOM_DTYPE, HOST_FLAG_NO_COUNTERS as FLAG_NO_COUNTERS, LABELS, ROW_DTYPE
from .core import COLUMNS, Context, Packed, default_context, load_atoms, pack_many, parse_cif, parse_pdb, read_table, split_rows
from .frame import to_arrow, to_frame
from . import sap  # noqa: F401  (the neighbour sum of the SAP score)

__version__ = "0.1.0"
__all__ = ["contacts", "Context", "Packed", "ArpeggiaError", "parse_pdb", "parse_cif", "load_atoms", "to_arrow", "to_frame", "COLUMNS"]


def contacts(input_file: str, groups: str = "/", vdw_comp: float = 0.1, dist_cutoff: float = 6.5,
             ignore_zero_occupancy: bool = False, num_threads: int = 1, *, device: int = 0):
    """Load a PDB or mmCIF file and calculate atomic and ring contacts (src/python.rs:31-56).

    Args:
        input_file: Path to the PDB or mmCIF file
        groups: Chain groups specification. Defaults to "/" (all-to-all).
            Examples: "A,B/C,D" for chains A,B vs C,D; "A/" for chain A vs all others.
        vdw_comp: VdW radii compensation factor. Defaults to 0.1.
        dist_cutoff: Distance cutoff for neighbor searches in Angstroms. Defaults to 6.5.
        ignore_zero_occupancy: If True, ignore atoms with zero occupancy. Defaults to False.
        num_threads: accepted for signature compatibility; results never depended on it (utils.rs:8-30)
            and the device path does not use host threads.
        device: CUDA device ordinal (extension; keyword only).

    Returns:
        polars.DataFrame (pyarrow.Table when polars is not installed) with columns model, interaction, distance,
        from_{chain,resn,resi,insertion,altloc,atomn,atomi}, to_{...}, sc_centroid_dist, sc_dihedral, sc_centroid_angle.

    Raises:
        ArpeggiaError: with the reference's panic messages for the same conditions (bad / empty chain groups,
        no ring residue at all, element without a van der Waals radius, ...), or when no CUDA device exists.
    """
    del num_threads
    # one C call, file -> sorted frame (arph_contacts_file: reader, packer, device search + classification, frame built on
    # the device); the reference's DataFrame carries no workload counters, so the device screens at the emit radius (same
    # rows).  The columns reach Arrow without a Python list (core.table_to_arrow).
    at, _ = default_context(device).contacts_file_arrow(input_file, groups, float(vdw_comp), float(dist_cutoff), bool(ignore_zero_occupancy),
                                                        flags=FLAG_NO_COUNTERS)
    try:
        import polars as pl
    except ImportError:
        return at
    return pl.from_arrow(at)
