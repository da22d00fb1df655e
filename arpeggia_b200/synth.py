"""Synthetic protein-density workloads (BASELINE.json configs 3-5).  No RNG except the seeded jitter.

All generators work on atom records (ATOM_DTYPE), i.e. on what a parser yields, so the same arrays feed the
product and the oracle.  Coordinates stay on the PDB 0.001 A lattice: shifts are applied in integer
milli-angstrom and converted back with one correctly rounded division, exactly like parsing the decimal text.
"""
from __future__ import annotations

import numpy as np

from ._lib import ATOM_DTYPE

# 6bft bounding box is 180.254 x 86.988 x 111.932 A; the pitch leaves a ~3 A gap so neighbouring copies touch
PITCH_6BFT = (183.300, 90.000, 114.900)
# shrunk pitch: copies interpenetrate until the mean occupancy of occupied 6.5 A cells is >= 14 atoms
PITCH_6BFT_DENSE = (91.650, 60.000, 57.450)


def _milli(v: np.ndarray) -> np.ndarray:
    return np.rint(v * 1000.0).astype(np.int64)


def copy_numbers(nx: int, ny: int, nz: int, x_range=None) -> list[int]:
    """Copy numbers c = (k*ny + j)*nx + i of a tiling, in generation order, restricted to x-indices i in x_range."""
    i0, i1 = (0, nx) if x_range is None else x_range
    return [(k * ny + j) * nx + i for k in range(nz) for j in range(ny) for i in range(i0, i1)]


def tile(atoms: np.ndarray, nx: int, ny: int, nz: int, pitch=PITCH_6BFT, serial_stride: int | None = None, x_range=None) -> np.ndarray:
    """nx*ny*nz translated copies of `atoms`; copy (i,j,k) is shifted by (i,j,k)*pitch and its chains are
    renamed "<id><5-digit copy number>" so every chain id stays unique (model 0).  x_range = (i0, i1) generates only
    the copies with i0 <= i < i1 -- one x-slab of the assembly -- with the copy numbers, chain ids and serials they
    have in the whole tiling (copy_numbers() lists them)."""
    atoms = np.ascontiguousarray(atoms, dtype=ATOM_DTYPE)
    n = len(atoms)
    copies = nx * ny * nz
    i0, i1 = (0, nx) if x_range is None else x_range
    if serial_stride is None:
        serial_stride = int(atoms["serial"].max()) + 1
    out = np.empty(n * (i1 - i0) * ny * nz, dtype=ATOM_DTYPE)
    mx, my, mz = _milli(atoms["x"]), _milli(atoms["y"]), _milli(atoms["z"])
    px, py, pz = (int(round(p * 1000)) for p in pitch)
    chains = np.unique(atoms["chain"])
    masks = [(ch, atoms["chain"] == ch) for ch in chains]
    q = 0
    for k in range(nz):
        for j in range(ny):
            for i in range(i0, i1):
                c = (k * ny + j) * nx + i
                blk = out[q * n:(q + 1) * n]
                blk[:] = atoms
                blk["x"] = (mx + i * px) / 1000.0
                blk["y"] = (my + j * py) / 1000.0
                blk["z"] = (mz + k * pz) / 1000.0
                blk["serial"] = atoms["serial"] + c * serial_stride
                if copies > 1:
                    for ch, m in masks:
                        blk["chain"][m] = ch + b"%05d" % c
                q += 1
    return out


def assembly_1m(base: np.ndarray, dense: bool = False) -> np.ndarray:
    """Config 3: 6bft tiled 5x5x5 = 1 022 500 atoms (dense=True: overlapping copies, interior density)."""
    return tile(base, 5, 5, 5, PITCH_6BFT_DENSE if dense else PITCH_6BFT)


def assembly_100m(base: np.ndarray) -> np.ndarray:
    """Config 5: 6bft tiled 16x32x24 = 100 515 840 atoms."""
    return tile(base, 16, 32, 24, PITCH_6BFT)


def jittered_complex(base: np.ndarray, k: int, chains=(b"A", b"B", b"C", b"G"), amplitude_milli: int = 50) -> np.ndarray:
    """Config 4, complex k: chains A,B,C,G of 6bft with a per-coordinate jitter uniform in [-0.050, +0.050] A on
    the 0.001 A lattice, PCG64 stream seeded 20260000 + k."""
    sub = np.ascontiguousarray(base[np.isin(base["chain"], list(chains))], dtype=ATOM_DTYPE).copy()
    rng = np.random.Generator(np.random.PCG64(20260000 + k))
    for ax in ("x", "y", "z"):
        sub[ax] = (_milli(sub[ax]) + rng.integers(-amplitude_milli, amplitude_milli + 1, size=len(sub))) / 1000.0
    return sub


def write_cif(atoms: np.ndarray, path) -> None:
    """Atom records -> a minimal PDBx/mmCIF file (one _atom_site row per line, label ids = author ids).  The tiled
    assemblies carry 6-character chain ids, which the one-column chain field of the PDB format cannot hold."""
    cols = ["group_PDB", "id", "type_symbol", "label_atom_id", "label_alt_id", "label_comp_id", "label_asym_id", "label_seq_id",
            "pdbx_PDB_ins_code", "Cartn_x", "Cartn_y", "Cartn_z", "occupancy", "auth_seq_id", "auth_asym_id", "pdbx_PDB_model_num"]

    def txt(field, blank="."):
        out = []
        for b in atoms[field].tolist():
            v = b.decode().strip()
            out.append(blank if not v else (f'"{v}"' if ("'" in v or " " in v) else v))
        return out

    het = ["HETATM" if h else "ATOM" for h in atoms["hetero"].tolist()]
    el, name, alt, resn, chain, ic = txt("element", "?"), txt("name"), txt("altloc"), txt("resn"), txt("chain"), txt("icode", "?")
    ser, resi, mod = atoms["serial"].tolist(), atoms["resi"].tolist(), np.maximum(atoms["model"], 1).tolist()
    x, y, z, occ = atoms["x"].tolist(), atoms["y"].tolist(), atoms["z"].tolist(), atoms["occupancy"].tolist()
    line = [f"{het[i]} {ser[i]} {el[i]} {name[i]} {alt[i]} {resn[i]} {chain[i]} {resi[i]} {ic[i]} {x[i]:.3f} {y[i]:.3f} {z[i]:.3f} "
            f"{occ[i]:.2f} {resi[i]} {chain[i]} {mod[i]}" for i in range(len(atoms))]
    with open(path, "w") as f:
        f.write("data_SYNTH\n#\nloop_\n" + "".join(f"_atom_site.{c}\n" for c in cols))
        f.write("\n".join(line))
        f.write("\n#\n")
