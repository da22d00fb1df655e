"""Host-side objects over the C ABI: parsed atom records, packed structures, device contexts, result tables.

Mirrors the reference's host flow for the contacts path (src/python.rs:31-56, src/contacts/mod.rs:61-136):
load -> (zero-occupancy filter) -> pack -> device search + classification -> 20-column table.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

from . import _lib
from ._lib import (ATOM_DTYPE, LABELS, ROW_DTYPE, ArpeggiaError, Frame, FrameTables, Params, Result, Stats, Structure, Table,
                   check, lib)

STRING_COLS = ("chain", "resn", "insertion", "altloc", "atomn")
COLUMNS = (
    ["model", "interaction", "distance"]
    + [f"from_{c}" for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi")]
    + [f"to_{c}" for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi")]
    + ["sc_centroid_dist", "sc_dihedral", "sc_centroid_angle"]
)


def _parse_with(fn_name: str, path, *extra) -> np.ndarray:
    L = lib()
    p = C.c_void_p()
    n = C.c_int64()
    check(getattr(L, fn_name)(os.fspath(path).encode(), *extra, C.byref(p), C.byref(n)))
    try:
        if n.value == 0:
            return np.zeros(0, dtype=ATOM_DTYPE)
        buf = (C.c_char * (n.value * ATOM_DTYPE.itemsize)).from_address(p.value)
        return np.frombuffer(bytes(buf), dtype=ATOM_DTYPE).copy()
    finally:
        L.arph_free(p)


def parse_pdb(path) -> np.ndarray:
    """ATOM/HETATM records of a PDB file as a structured array (ATOM_DTYPE)."""
    return _parse_with("arph_parse_pdb", path)


def parse_cif(path, auth_ids: bool = False) -> np.ndarray:
    """The _atom_site loop of an mmCIF file as a structured array (ATOM_DTYPE).  Chain ids / residue numbers are the label
    identifiers (pdbtbx's choice as far as it is known); auth_ids=True takes the author identifiers instead."""
    return _parse_with("arph_parse_cif_ids", path, int(bool(auth_ids)))


def load_atoms(path) -> np.ndarray:
    """pdbtbx-style dispatch on the file extension (src/utils.rs:51-57): .cif / .mmcif -> mmCIF, else PDB."""
    return _parse_with("arph_parse_file", path)


class Packed:
    """arph_packed: device SoA + decode dictionaries for one structure."""

    def __init__(self, atoms: np.ndarray, groups: str = "/", ignore_zero_occupancy: bool = False, flags: int = 0):
        L = lib()
        self._atoms = np.ascontiguousarray(atoms, dtype=ATOM_DTYPE)
        self._h = C.c_void_p()
        check(L.arph_pack(self._atoms.ctypes.data, len(self._atoms), groups.encode(), int(ignore_zero_occupancy), flags,
                          C.byref(self._h)))

    @property
    def structure(self) -> Structure:
        return lib().arph_packed_structure(self._h).contents

    @property
    def structure_ptr(self):
        return lib().arph_packed_structure(self._h)

    @property
    def warnings(self) -> list[str]:
        L = lib()
        return [L.arph_packed_warning(self._h, i).decode() for i in range(L.arph_packed_n_warnings(self._h))]

    def arrays(self, copy: bool = True) -> dict:
        """The SoA as numpy arrays (copy=False: views that are only valid while this object lives)."""
        s = self.structure
        out = {"n_atoms": s.n_atoms, "n_residues": s.n_residues, "n_rings": s.n_rings, "n_h": s.n_h}

        def view(ptr, n, dt):
            v = np.ctypeslib.as_array(ptr, shape=(n,)).view(dt) if n else np.zeros(0, dtype=dt)
            return v.copy() if copy else v

        for k in ("x", "y", "z"):
            out[k] = view(getattr(s, k), s.n_atoms, np.float64)
        for k in ("attr", "chain_key", "res_ordinal", "res_id"):
            out[k] = view(getattr(s, k), s.n_atoms, np.uint32)
        out["res_h_begin"] = view(s.res_h_begin, s.n_residues + 1, np.uint32)
        for k in ("h_x", "h_y", "h_z"):
            out[k] = view(getattr(s, k), s.n_h, np.float64)
        for k in ("res_cb", "res_sg"):
            out[k] = view(getattr(s, k), s.n_residues, np.int32)
        for k in ("res_chain_key", "res_res_ordinal"):
            out[k] = view(getattr(s, k), s.n_residues, np.uint32)
        out["res_atom_begin"] = view(s.res_atom_begin, s.n_residues + 1, np.uint32) if s.res_atom_begin else None
        for k in ("x_milli", "y_milli", "z_milli"):  # on-lattice form of the coordinates, None when a coordinate is off it
            out[k] = view(getattr(s, k), s.n_atoms, np.int32) if getattr(s, k) else None
        # the most compact forms: 21-bit lattice words relative to xyz_milli21_origin, 16-bit attribute words
        out["xyz_milli21"] = view(s.xyz_milli21, s.n_atoms, np.uint64) if s.xyz_milli21 else None
        out["xyz_milli21_origin"] = np.array(list(s.xyz_milli21_origin), dtype=np.int32)
        out["attr16"] = view(s.attr16, s.n_atoms, np.uint16) if s.attr16 else None
        for k in ("ring_center", "ring_normal"):
            out[k] = view(getattr(s, k), 3 * s.n_rings, np.float64).reshape(-1, 3)
        for k in ("ring_attr", "ring_chain_key", "ring_res_ordinal"):
            out[k] = view(getattr(s, k), s.n_rings, np.uint32)
        out["res_tag"] = view(s.res_tag, s.n_residues, np.uint32) if s.res_tag else None
        out["atom_gid"] = view(s.atom_gid, s.n_atoms, np.uint32) if s.atom_gid else None
        out["ring_gid"] = view(s.ring_gid, s.n_rings, np.uint32) if s.ring_gid else None
        return out

    # ---- this structure as one part of a larger one (slab jobs) ------------------------------------------------
    def chain_ids(self) -> list[bytes]:
        """Chain ids in byte order = the packer's local chain ranks (low 24 bits of chain_key)."""
        L = lib()
        return [L.arph_packed_chain_id(self._h, k) for k in range(L.arph_packed_n_chains(self._h))]

    def remap_chain_ranks(self, new_rank) -> None:
        """Replace the local chain ranks by the ranks of the same chains in the undivided structure (complex.rs:129)."""
        m = np.ascontiguousarray(new_rank, dtype=np.uint32)
        check(lib().arph_packed_remap_chain_ranks(self._h, m.ctypes.data_as(_lib._U32P), len(m)))

    def set_global_ids(self, atom_gid, ring_gid, res_tag=None) -> None:
        """Ids of the atoms / rings in the undivided structure (rows then carry them) and residue tags valid across parts."""
        a = np.ascontiguousarray(atom_gid, dtype=np.uint32)
        g = np.ascontiguousarray(ring_gid, dtype=np.uint32)
        s = self.structure
        assert len(a) == s.n_atoms and len(g) == s.n_rings
        t = None if res_tag is None else np.ascontiguousarray(res_tag, dtype=np.uint32)
        assert t is None or len(t) == s.n_residues
        check(lib().arph_packed_set_global_ids(self._h, a.ctypes.data_as(_lib._U32P), g.ctypes.data_as(_lib._U32P),
                                               t.ctypes.data_as(_lib._U32P) if t is not None else None))

    def atom_serials(self) -> np.ndarray:
        """Serial number of every packed atom (atoms_with_hierarchy() order can differ from the record order)."""
        import ctypes as _C

        from ._lib import FrameTables
        tp = _C.POINTER(FrameTables)()
        check(lib().arph_packed_frame_tables(self._h, _C.byref(tp)))
        n = self.structure.n_atoms
        return np.ctypeslib.as_array(tp.contents.ent_atomi, shape=(max(n, 1),))[:n].astype(np.int64)

    def build_table(self, rows: np.ndarray) -> dict:
        L = lib()
        rows = np.ascontiguousarray(rows, dtype=ROW_DTYPE)
        tp = C.POINTER(Table)()
        check(L.arph_build_table(self._h, rows.ctypes.data, len(rows), C.byref(tp)))
        try:
            return read_table(tp.contents)
        finally:
            L.arph_table_free(tp)

    def close(self):
        if getattr(self, "_h", None):
            try:
                lib().arph_packed_free(self._h)
            except Exception:  # interpreter teardown
                pass
            self._h = None

    __del__ = close


def pack_many(atom_arrays, groups: str = "/", ignore_zero_occupancy: bool = False, flags: int = 0) -> list:
    """arph_pack_many: a list of atom-record arrays -> a list of Packed, packed on all host threads (one structure each)."""
    L = lib()
    arrs = [np.ascontiguousarray(a, dtype=ATOM_DTYPE) for a in atom_arrays]
    count = len(arrs)
    ptrs = (C.c_void_p * count)(*[a.ctypes.data for a in arrs])
    ns = (C.c_int64 * count)(*[len(a) for a in arrs])
    out = (C.c_void_p * count)()
    check(L.arph_pack_many(ptrs, ns, count, groups.encode(), int(ignore_zero_occupancy), flags, out))
    packs = []
    for k in range(count):
        p = Packed.__new__(Packed)
        p._atoms, p._h = arrs[k], C.c_void_p(out[k])
        packs.append(p)
    return packs


def read_table(t: Table) -> dict:
    """Copy an arph_table into numpy arrays (string columns as object arrays)."""
    n = t.n_rows
    strings = np.array([t.strings[i].decode() for i in range(t.n_strings)] + [""], dtype=object)

    def arr(ptr, dt):
        return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True) if n else np.zeros(0, dtype=dt)

    out = {"n_rows": n, "model": arr(t.model, np.uint32)}
    code = arr(t.interaction, np.uint8)
    out["interaction_code"] = code
    out["interaction"] = np.array(LABELS, dtype=object)[code] if n else np.zeros(0, dtype=object)
    out["distance"] = arr(t.distance, np.float32)
    for side in ("from", "to"):
        for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi"):
            v = arr(getattr(t, f"{side}_{c}"), np.int32)
            out[f"{side}_{c}"] = strings[v] if c in STRING_COLS else v
    out["sc_valid"] = arr(t.sc_valid, np.uint8).astype(bool)
    out["sc_validity_bits"] = np.ctypeslib.as_array(t.sc_validity, shape=((n + 7) // 8,)).copy() if n else np.zeros(0, dtype=np.uint8)
    out["sc_null_count"] = int(t.sc_null_count)
    out["strings"] = [str(v) for v in strings[:-1]]  # the table's string dictionary (ids of the string columns index it)
    for c in ("sc_centroid_dist", "sc_dihedral", "sc_centroid_angle"):
        out[c] = arr(getattr(t, c), np.float32)
    for c in ("n_atoms", "pairs_in_cutoff", "pairs_classified", "t_setup_s", "t_classify_s", "t_table_s"):
        out[c] = getattr(t, c)
    return out


class _TableOwner:
    """Keeps an arph_table alive for the Arrow buffers that point into it; frees it when the last of them is gone."""

    def __init__(self, tp):
        self.tp = tp

    def __del__(self):
        try:
            lib().arph_table_free(self.tp)
        except Exception:  # interpreter teardown
            pass


def table_to_arrow(tp, strings: str = "string"):
    """POINTER(arph_table) -> pyarrow.Table with the reference's schema (src/contacts/mod.rs:140-214).  TAKES OWNERSHIP of
    the table.  No Python list and no copy of a column: every numeric column is an Arrow array over the table's own
    (page-locked) memory, and the int32 id columns are the INDEX buffers of Arrow dictionary arrays over the table's string
    dictionary.  strings="dictionary" leaves those columns dictionary-encoded (polars: Categorical); "string" decodes them
    to large_utf8 in Arrow's C++ (polars: String, the reference's dtype)."""
    import pyarrow as pa

    t = tp.contents
    n = int(t.n_rows)
    owner = _TableOwner(tp)
    dic = pa.array([t.strings[i].decode() for i in range(t.n_strings)], type=pa.large_string())  # a few thousand entries
    lab = pa.array(LABELS, type=pa.large_string())

    def wrap(ptr, typ, width, validity=None, nulls=0):
        if n == 0:
            return pa.array([], type=typ)
        buf = pa.foreign_buffer(C.addressof(ptr.contents), n * width, base=owner)
        return pa.Array.from_buffers(typ, n, [validity, buf], null_count=nulls)

    def text(idx, d):
        if n == 0:
            return pa.array([], type=pa.large_string()) if strings != "dictionary" else pa.DictionaryArray.from_arrays(pa.array([], type=pa.int32()), d)
        return pa.DictionaryArray.from_arrays(idx, d, safe=False)  # the library only ever writes ids of its own dictionary

    bitmap, n_null = None, 0
    if n:  # the library hands the validity of the sc_* columns over in Arrow's own layout
        n_null = int(t.sc_null_count)
        bitmap = pa.foreign_buffer(C.addressof(t.sc_validity.contents), (n + 7) // 8, base=owner)
    cols = []
    for name in COLUMNS:
        if name == "model":
            a = wrap(t.model, pa.uint32(), 4)
        elif name == "interaction":
            a = text(wrap(t.interaction, pa.uint8(), 1), lab)
        elif name == "distance":
            a = wrap(t.distance, pa.float32(), 4)
        elif name.startswith("sc_"):
            a = wrap(getattr(t, name), pa.float32(), 4, bitmap, n_null)
        elif name.split("_", 1)[-1] in STRING_COLS:
            a = text(wrap(getattr(t, name), pa.int32(), 4), dic)
        else:
            a = wrap(getattr(t, name), pa.int32(), 4)
        cols.append(a)
    if strings != "dictionary" and n:
        # String columns (the reference's dtype): the eleven decodes are independent and Arrow's C++ releases the GIL, so
        # on a large frame they run side by side
        todo = [k for k, a in enumerate(cols) if pa.types.is_dictionary(a.type)]
        if n >= 100_000:
            from concurrent.futures import ThreadPoolExecutor

            with ThreadPoolExecutor(max_workers=len(todo)) as ex:
                done = list(ex.map(lambda k: cols[k].dictionary_decode(), todo))
        else:
            done = [cols[k].dictionary_decode() for k in todo]
        for k, a in zip(todo, done):
            cols[k] = a
    return pa.table(cols, names=list(COLUMNS))


class Context:
    """arp_ctx: one per (device, host thread).  Creating it without a CUDA device raises ArpeggiaError."""

    def __init__(self, device: int = 0):
        L = lib()
        self._h = C.c_void_p()
        check(L.arp_ctx_create(C.byref(self._h), device))
        self.device = device

    @property
    def stream(self) -> int:
        """cudaStream_t (as int) every kernel of this context is launched on."""
        return lib().arp_ctx_stream(self._h) or 0

    # ---- one-shot ---------------------------------------------------------------------------------
    def contacts_rows(self, structures, vdw_comp=0.1, dist_cutoff=6.5, flags=0) -> dict:
        """arp_contacts / arp_contacts_batch on Packed structures (or raw Structure pointers) -> raw rows."""
        L = lib()
        if isinstance(structures, Packed):
            structures = [structures]
        ptrs = (C.POINTER(Structure) * len(structures))(*[s.structure_ptr if hasattr(s, "structure_ptr") else s for s in structures])
        prm = Params(vdw_comp, dist_cutoff, flags, 0)
        rp = C.POINTER(Result)()
        check(L.arp_contacts_batch(self._h, ptrs, len(structures), C.byref(prm), C.byref(rp)), self._h)
        try:
            return _read_result(rp.contents)
        finally:
            L.arp_result_free(self._h, rp)

    def contacts_table(self, atoms: np.ndarray, groups="/", vdw_comp=0.1, dist_cutoff=6.5, ignore_zero_occupancy=False,
                       flags=0) -> dict:
        """arph_contacts: get_contacts on atom records -> sorted 20-column table (dict of numpy arrays)."""
        L = lib()
        atoms = np.ascontiguousarray(atoms, dtype=ATOM_DTYPE)
        tp = C.POINTER(Table)()
        check(L.arph_contacts(self._h, atoms.ctypes.data, len(atoms), groups.encode(), vdw_comp, dist_cutoff,
                              int(ignore_zero_occupancy), flags, C.byref(tp)))
        try:
            return read_table(tp.contents)
        finally:
            L.arph_table_free(tp)

    def contacts_file_arrow(self, path, groups="/", vdw_comp=0.1, dist_cutoff=6.5, ignore_zero_occupancy=False, flags=0,
                            strings: str = "string"):
        """arph_contacts_file: file -> (parse, pack, device, frame) -> pyarrow.Table, plus the host-side timings."""
        L = lib()
        tp = C.POINTER(Table)()
        check(L.arph_contacts_file(self._h, os.fspath(path).encode(), groups.encode(), vdw_comp, dist_cutoff, int(ignore_zero_occupancy),
                                   flags, C.byref(tp)))
        t = tp.contents
        info = {k: getattr(t, k) for k in ("n_rows", "n_atoms", "t_setup_s", "t_classify_s", "t_table_s")}
        return table_to_arrow(tp, strings), info  # the Arrow table owns the arph_table from here on

    # ---- staged -------------------------------------------------------------------------------------
    def upload(self, structures):
        if isinstance(structures, Packed):
            structures = [structures]
        self._keep = structures  # host arrays must stay alive during the call only, but keep refs anyway
        ptrs = (C.POINTER(Structure) * len(structures))(*[s.structure_ptr if hasattr(s, "structure_ptr") else s for s in structures])
        check(lib().arp_stage_upload(self._h, ptrs, len(structures)), self._h)

    def run(self, vdw_comp=0.1, dist_cutoff=6.5, flags=0):
        prm = Params(vdw_comp, dist_cutoff, flags, 0)
        check(lib().arp_stage_run(self._h, C.byref(prm)), self._h)

    def neighbor_sum(self, weight: np.ndarray, select: np.ndarray, radius: float = 5.0) -> np.ndarray:
        """arp_stage_neighbor_sum on the uploaded job (a batch: every structure on its own): the radius query of the SAP score
        (src/sap.rs:181-215).  out[x] = sum of weight[y] over the selected atoms y within `radius` of the selected atom x
        (x included), 0 where x is not selected.  Arrays are in the job's atom order (Packed.arrays())."""
        w = np.ascontiguousarray(weight, dtype=np.float32)
        sel = np.ascontiguousarray(select, dtype=np.uint8)
        out = np.zeros(len(w), dtype=np.float32)
        check(lib().arp_stage_neighbor_sum(self._h, w.ctypes.data, sel.ctypes.data, C.c_float(radius), out.ctypes.data), self._h)
        return out

    def neighbor_sum_stats(self) -> dict:
        g, s_, n = C.c_float(), C.c_float(), C.c_int64()
        check(lib().arp_neighbor_sum_stats(self._h, C.byref(g), C.byref(s_), C.byref(n)), self._h)
        return {"ms_grid": g.value, "ms_sum": s_.value, "n_cells": n.value}

    def download(self) -> dict:
        L = lib()
        rp = C.POINTER(Result)()
        check(L.arp_stage_download(self._h, C.byref(rp)), self._h)
        try:
            return _read_result(rp.contents)
        finally:
            L.arp_result_free(self._h, rp)

    def frame(self, packed: "Packed", arrow: str | None = None, device_columns: bool = False):
        """arp_stage_frame_compact after run(): sort + sc join on the device, 26 bytes per row back, the per-entity columns
        gathered by arph_table_from_frame on the host's threads -> the table of get_contacts, plus the device timings of
        the frame stage under "frame_stats".  device_columns=True: arp_stage_frame (every column decoded on the device,
        86 bytes per row back)."""
        L = lib()
        tp = C.POINTER(FrameTables)()
        check(L.arph_packed_frame_tables(packed._h, C.byref(tp)))
        fp = C.POINTER(Frame)()
        check((L.arp_stage_frame if device_columns else L.arp_stage_frame_compact)(self._h, tp, C.byref(fp)), self._h)
        fs = {k: getattr(fp.contents, k) for k in ("ms_sort", "ms_columns", "ms_d2h", "sort_passes", "key_bits")}
        tb = C.POINTER(Table)()
        check(L.arph_table_from_frame(packed._h, fp, C.byref(tb)))  # owns the frame from here on
        if arrow:
            return table_to_arrow(tb, arrow), fs  # owns the table (and the frame's page-locked columns under it)
        try:
            out = read_table(tb.contents)
        finally:
            L.arph_table_free(tb)
        out["frame_stats"] = fs
        return out

    # ---- slab jobs ----------------------------------------------------------------------------------------
    def comm_init(self, comm_id: "_lib.CommId", rank: int, world: int) -> None:
        """arp_ctx_comm_init: join the communicator whose id one rank made with comm_get_id() (one process per GPU)."""
        check(lib().arp_ctx_comm_init(self._h, C.byref(comm_id), rank, world), self._h)

    def slab_exchange(self, vdw_comp=0.1, dist_cutoff=6.5) -> dict:
        """arp_slab_exchange: complete the uploaded (owned) part with halo copies from the other ranks."""
        prm = Params(vdw_comp, dist_cutoff, 0, 0)
        info = _lib.SlabInfo()
        check(lib().arp_slab_exchange(self._h, C.byref(prm), C.byref(info)), self._h)
        return {k: getattr(info, k) for k, _ in _lib.SlabInfo._fields_}

    def stats(self) -> dict:
        st = Stats()
        check(lib().arp_stage_stats(self._h, C.byref(st)), self._h)
        return {k: getattr(st, k) for k, _ in Stats._fields_ if not k.startswith("reserved")}

    def close(self):
        if getattr(self, "_h", None):
            try:
                lib().arp_ctx_destroy(self._h)
            except Exception:  # interpreter teardown: the module globals are already gone
                pass
            self._h = None

    __del__ = close


def _read_result(r: Result) -> dict:
    n = r.n_rows
    if n:
        buf = (C.c_char * (n * ROW_DTYPE.itemsize)).from_address(r.rows)
        rows = np.frombuffer(buf, dtype=ROW_DTYPE).copy()
    else:
        rows = np.zeros(0, dtype=ROW_DTYPE)
    ns = r.n_structures
    return {
        "rows": rows,
        "atom_begin": np.ctypeslib.as_array(r.atom_begin, shape=(ns + 1,)).copy(),
        "ring_begin": np.ctypeslib.as_array(r.ring_begin, shape=(ns + 1,)).copy(),
        "pairs_screened": r.pairs_screened,
        "pairs_in_cutoff": r.pairs_in_cutoff,
        "pairs_classified": r.pairs_classified,
    }


def split_rows(result: dict) -> list[np.ndarray]:
    """Rows of a batch job -> one array per structure with structure-local atom / ring indices."""
    rows, ab, rb = result["rows"], result["atom_begin"], result["ring_begin"]
    is_ring_f = (rows["from"] & _lib.RING_FLAG) != 0
    idx_f = (rows["from"] & ~np.uint32(_lib.RING_FLAG)).astype(np.int64)
    s = np.where(is_ring_f, np.searchsorted(rb, idx_f, side="right") - 1, np.searchsorted(ab, idx_f, side="right") - 1)
    out = []
    order = np.argsort(s, kind="stable")
    bounds = np.searchsorted(s[order], np.arange(len(ab)))
    for k in range(len(ab) - 1):
        part = rows[order[bounds[k]:bounds[k + 1]]].copy()
        for col in ("from", "to"):
            ring = (part[col] & _lib.RING_FLAG) != 0
            v = (part[col] & ~np.uint32(_lib.RING_FLAG)).astype(np.int64)
            v = np.where(ring, v - rb[k], v - ab[k]).astype(np.uint32)
            part[col] = np.where(ring, v | np.uint32(_lib.RING_FLAG), v)
        out.append(part)
    return out


def comm_get_id() -> "_lib.CommId":
    """arp_comm_get_id: a fresh communicator id (ncclUniqueId); distribute its 128 bytes to every rank."""
    cid = _lib.CommId()
    check(lib().arp_comm_get_id(C.byref(cid)))
    return cid


def create_multi(devices) -> list[Context]:
    """arp_ctx_create_multi: one context per entry of `devices`, driven by this process.  Distinct devices talk over NCCL
    (ncclCommInitAll); a repeated device makes the set use peer copies (the single-GPU test configuration)."""
    n = len(devices)
    hs = (C.c_void_p * n)()
    dv = (C.c_int * n)(*devices)
    check(lib().arp_ctx_create_multi(hs, dv, n))
    out = []
    for k in range(n):
        cx = Context.__new__(Context)
        cx._h = C.c_void_p(hs[k])
        cx.device = devices[k]
        out.append(cx)
    return out


def slab_exchange_multi(ctxs, vdw_comp=0.1, dist_cutoff=6.5) -> list[dict]:
    """arp_slab_exchange_multi on every context of a create_multi() set."""
    n = len(ctxs)
    hs = (C.c_void_p * n)(*[cx._h for cx in ctxs])
    prm = Params(vdw_comp, dist_cutoff, 0, 0)
    info = (_lib.SlabInfo * n)()
    check(lib().arp_slab_exchange_multi(hs, n, C.byref(prm), info), ctxs[0]._h)
    return [{k: getattr(info[i], k) for k, _ in _lib.SlabInfo._fields_} for i in range(n)]


_default = threading.local()


def default_context(device: int = 0) -> Context:
    """Per-thread cached context for the convenience API."""
    cache = getattr(_default, "ctx", None)
    if cache is None:
        cache = _default.ctx = {}
    if device not in cache:
        cache[device] = Context(device)
    return cache[device]
