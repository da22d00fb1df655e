"""ctypes binding of the CPU oracle (oracle/arpeggia_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "_build" / "liboracle.so"

# byte-identical to orc_atom_record (oracle.h) and arp_atom_record (include/arpeggia_host.h)
ATOM_DTYPE = np.dtype(
    [
        ("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("occupancy", "<f8"),
        ("serial", "<i8"), ("resi", "<i8"), ("model", "<i8"),
        ("chain", "S8"), ("resn", "S8"), ("name", "S8"), ("element", "S4"),
        ("altloc", "S2"), ("icode", "S2"), ("hetero", "u1"), ("pad_", "V7"),
    ],
    align=False,
)
assert ATOM_DTYPE.itemsize == 96

INTERACTIONS = [
    "StericClash", "CovalentBond", "Disulfide", "VanDerWaalsContact", "IonicBond", "HydrogenBond",
    "WeakHydrogenBond", "PolarContact", "WeakPolarContact", "IonicRepulsion", "SaltBridge",
    "PiDisplacedStacking", "PiTStacking", "PiSandwichStacking", "PiParallelInPlaneStacking",
    "PiTiltedStacking", "PiLStacking", "CationPi", "HydrophobicContact",
]

FLAG_NO_MULTIMODEL_QUIRK = 1
FLAG_SKIP_LOAD_FILTER = 2
FLAG_NO_TABLE = 4

_I32P = C.POINTER(C.c_int32)
_F32P = C.POINTER(C.c_float)


class _Table(C.Structure):
    _fields_ = (
        [("n_rows", C.c_int64), ("model", C.POINTER(C.c_uint32)), ("interaction", C.POINTER(C.c_uint8)),
         ("distance", _F32P)]
        + [(f"{s}_{c}", _I32P) for s in ("from", "to")
           for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi")]
        + [("sc_centroid_dist", _F32P), ("sc_dihedral", _F32P), ("sc_centroid_angle", _F32P),
           ("sc_valid", C.POINTER(C.c_uint8)), ("n_strings", C.c_int32), ("strings", C.POINTER(C.c_char_p)),
           ("n_atoms", C.c_int64), ("pairs_in_cutoff", C.c_int64), ("pairs_classified", C.c_int64),
           ("t_setup_s", C.c_double), ("t_search_s", C.c_double), ("t_classify_s", C.c_double),
           ("t_rings_s", C.c_double), ("t_table_s", C.c_double)]
    )


STRING_COLS = ["chain", "resn", "insertion", "altloc", "atomn"]
COLUMNS = (
    ["model", "interaction", "distance"]
    + [f"from_{c}" for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi")]
    + [f"to_{c}" for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi")]
    + ["sc_centroid_dist", "sc_dihedral", "sc_centroid_angle"]
)


def build(force: bool = False) -> Path:
    """Compile the oracle with oracle/Makefile (g++; no GPU, no reference sources needed)."""
    src = _HERE / "arpeggia_oracle.cpp"
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < max(
        src.stat().st_mtime, (_HERE / "oracle.h").stat().st_mtime
    ):
        subprocess.run(["make", "-C", str(_HERE)], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(_LIB_PATH))
        L.orc_last_error.restype = C.c_char_p
        L.orc_interaction_name.restype = C.c_char_p
        L.orc_parse_pdb.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_contacts.argtypes = [C.c_void_p, C.c_int64, C.c_char_p, C.c_double, C.c_double, C.c_int,
                                   C.c_uint32, C.c_int, C.POINTER(C.POINTER(_Table))]
        L.orc_table_free.argtypes = [C.POINTER(_Table)]
        L.orc_plane_fit.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_plane_point_angle.argtypes = [C.c_void_p] * 3
        L.orc_plane_point_angle.restype = C.c_double
        L.orc_plane_point_dist.argtypes = [C.c_void_p] * 2
        L.orc_plane_point_dist.restype = C.c_double
        L.orc_plane_dihedral.argtypes = [C.c_void_p] * 2
        L.orc_plane_dihedral.restype = C.c_double
        L.orc_neighbor_sum.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        L.orc_parse_groups.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64]
        L.orc_residue_plane.argtypes = [C.c_void_p, C.c_int64, C.c_char_p, C.c_int64, C.c_int,
                                        C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


class OracleError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


def _check(rc: int):
    if rc != 0:
        raise OracleError(rc, lib().orc_last_error().decode())


def parse_pdb(path: str | os.PathLike) -> np.ndarray:
    L = lib()
    p = C.c_void_p()
    n = C.c_int64()
    _check(L.orc_parse_pdb(str(path).encode(), C.byref(p), C.byref(n)))
    try:
        buf = (C.c_char * (n.value * ATOM_DTYPE.itemsize)).from_address(p.value) if n.value else b""
        return np.frombuffer(bytes(buf), dtype=ATOM_DTYPE).copy()
    finally:
        L.orc_free(p)


def read_table(t, strings_free=None) -> dict:
    """Copy an orc_table / arph_table (same layout) into a dict of numpy arrays + python string lists."""
    n = t.n_rows
    out = {"n_rows": n}
    strings = [t.strings[i].decode() for i in range(t.n_strings)]
    sarr = np.array(strings, dtype=object) if strings else np.array([], dtype=object)

    def arr(ptr, dt):
        if n == 0 or not ptr:
            return np.zeros(0, dtype=dt)
        return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True)

    has_cols = bool(t.model) or n == 0
    out["has_columns"] = has_cols
    if has_cols:
        out["model"] = arr(t.model, np.uint32)
        code = arr(t.interaction, np.uint8)
        out["interaction_code"] = code
        out["interaction"] = np.array(INTERACTIONS, dtype=object)[code] if n else np.array([], dtype=object)
        out["distance"] = arr(t.distance, np.float32)
        for side in ("from", "to"):
            for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi"):
                v = arr(getattr(t, f"{side}_{c}"), np.int32)
                out[f"{side}_{c}"] = sarr[v] if c in STRING_COLS and n else (v if c not in STRING_COLS else np.array([], dtype=object))
        valid = arr(t.sc_valid, np.uint8).astype(bool)
        out["sc_valid"] = valid
        for c in ("sc_centroid_dist", "sc_dihedral", "sc_centroid_angle"):
            out[c] = arr(getattr(t, c), np.float32)
    for c in ("n_atoms", "pairs_in_cutoff", "pairs_classified", "t_setup_s", "t_search_s", "t_classify_s",
              "t_rings_s", "t_table_s"):
        out[c] = getattr(t, c)
    return out


def contacts(atoms: np.ndarray, groups: str = "/", vdw_comp: float = 0.1, dist_cutoff: float = 6.5,
             ignore_zero_occupancy: bool = False, flags: int = 0, num_threads: int = 1) -> dict:
    """get_contacts on atom records -> dict of columns (sorted by the reference's 10 keys)."""
    L = lib()
    atoms = np.ascontiguousarray(atoms, dtype=ATOM_DTYPE)
    tp = C.POINTER(_Table)()
    _check(L.orc_contacts(atoms.ctypes.data, len(atoms), groups.encode(), vdw_comp, dist_cutoff,
                          int(ignore_zero_occupancy), flags, num_threads, C.byref(tp)))
    try:
        return read_table(tp.contents)
    finally:
        L.orc_table_free(tp)


def contacts_file(path, **kw) -> dict:
    return contacts(parse_pdb(path), **kw)


def plane_fit(points: np.ndarray):
    pts = np.ascontiguousarray(points, dtype=np.float64)
    c = np.zeros(3)
    nrm = np.zeros(3)
    _check(lib().orc_plane_fit(pts.ctypes.data, len(pts), c.ctypes.data, nrm.ctypes.data))
    return c, nrm


def plane_point_angle(center, normal, point) -> float:
    a, b, c = (np.ascontiguousarray(v, dtype=np.float64) for v in (center, normal, point))
    return lib().orc_plane_point_angle(a.ctypes.data, b.ctypes.data, c.ctypes.data)


def plane_point_dist(center, point) -> float:
    a, c = (np.ascontiguousarray(v, dtype=np.float64) for v in (center, point))
    return lib().orc_plane_point_dist(a.ctypes.data, c.ctypes.data)


def plane_dihedral(n1, n2) -> float:
    a, b = (np.ascontiguousarray(v, dtype=np.float64) for v in (n1, n2))
    return lib().orc_plane_dihedral(a.ctypes.data, b.ctypes.data)


def parse_groups(all_chains, groups: str):
    lig = C.create_string_buffer(1 << 16)
    rec = C.create_string_buffer(1 << 16)
    _check(lib().orc_parse_groups(",".join(sorted(all_chains)).encode(), groups.encode(), lig, rec, 1 << 16))
    f = lambda b: set(x for x in b.value.decode().split(",") if x)
    return f(lig), f(rec)


def residue_plane(atoms: np.ndarray, chain: str, resi: int, kind: str = "ring"):
    atoms = np.ascontiguousarray(atoms, dtype=ATOM_DTYPE)
    n = C.c_int64()
    c = np.zeros(3)
    nrm = np.zeros(3)
    rc = lib().orc_residue_plane(atoms.ctypes.data, len(atoms), chain.encode(), resi, 0 if kind == "ring" else 1,
                                 C.byref(n), c.ctypes.data, nrm.ctypes.data)
    if rc != 0:
        return n.value, None, None
    return n.value, c, nrm


def neighbor_sum(x, y, z, weight, select, radius: float) -> np.ndarray:
    """src/sap.rs:181-215 on caller-supplied weights: per selected atom, the f32 sum of weight over the selected atoms
    within `radius` (itself included, inclusive comparison), in increasing atom order; 0 where not selected."""
    x, y, z = (np.ascontiguousarray(v, dtype=np.float64) for v in (x, y, z))
    w = np.ascontiguousarray(weight, dtype=np.float32)
    sel = np.ascontiguousarray(select, dtype=np.uint8)
    out = np.zeros(len(x), dtype=np.float32)
    _check(lib().orc_neighbor_sum(x.ctypes.data, y.ctypes.data, z.ctypes.data, len(x), w.ctypes.data, sel.ctypes.data,
                                  C.c_float(radius), out.ctypes.data))
    return out
