"""ctypes binding of libarpeggia_b200.so (include/arpeggia_cuda.h + include/arpeggia_host.h).

The library is built in-tree by `make -C arpeggia_b200/csrc` (see __graft_entry__.build()).  There is no
fallback of any kind: a missing library raises ImportError, a missing GPU makes every compute call raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_PKG = Path(__file__).resolve().parent
# ARPEGGIA_B200_LIB: kernel-tuning builds of the same library (scratch/variants.sh); never a different implementation
LIB_PATH = Path(os.environ.get("ARPEGGIA_B200_LIB") or (_PKG / "_build" / "libarpeggia_b200.so"))

# byte-identical to arp_atom_record (include/arpeggia_host.h)
ATOM_DTYPE = np.dtype(
    [
        ("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("occupancy", "<f8"),
        ("serial", "<i8"), ("resi", "<i8"), ("model", "<i8"),
        ("chain", "S8"), ("resn", "S8"), ("name", "S8"), ("element", "S4"),
        ("altloc", "S2"), ("icode", "S2"), ("hetero", "u1"), ("pad_", "V7"),
    ]
)
assert ATOM_DTYPE.itemsize == 96

ROW_DTYPE = np.dtype([("from", "<u4"), ("to", "<u4"), ("label", "<u4"), ("distance", "<f4")])
RING_FLAG = 0x80000000
FLAG_EMIT_RADIUS_SCREEN = 1
FLAG_COUNT_CLASSIFIED = 2
STATIC_MAX_ATOMS = 131072  # ARP_STATIC_MAX_ATOMS
HOST_FLAG_NO_MULTIMODEL_QUIRK = 1
HOST_FLAG_SKIP_LOAD_FILTER = 2
HOST_FLAG_NO_COUNTERS = 4
HOST_FLAG_DEVICE_COLUMNS = 8

LABELS = [
    "StericClash", "CovalentBond", "Disulfide", "VanDerWaalsContact", "IonicBond", "HydrogenBond",
    "WeakHydrogenBond", "PolarContact", "WeakPolarContact", "IonicRepulsion", "SaltBridge",
    "PiDisplacedStacking", "PiTStacking", "PiSandwichStacking", "PiParallelInPlaneStacking",
    "PiTiltedStacking", "PiLStacking", "CationPi", "HydrophobicContact",
]

_F64P = C.POINTER(C.c_double)
_U32P = C.POINTER(C.c_uint32)
_I32P = C.POINTER(C.c_int32)
_I64P = C.POINTER(C.c_int64)
_F32P = C.POINTER(C.c_float)


class Structure(C.Structure):
    """arp_structure"""
    _fields_ = [
        ("n_atoms", C.c_int64), ("n_residues", C.c_int64), ("n_rings", C.c_int64), ("n_h", C.c_int64),
        ("x", _F64P), ("y", _F64P), ("z", _F64P),
        ("attr", _U32P), ("chain_key", _U32P), ("res_ordinal", _U32P), ("res_id", _U32P), ("res_h_begin", _U32P),
        ("h_x", _F64P), ("h_y", _F64P), ("h_z", _F64P),
        ("res_cb", _I32P), ("res_sg", _I32P),
        ("ring_center", _F64P), ("ring_normal", _F64P),
        ("ring_attr", _U32P), ("ring_chain_key", _U32P), ("ring_res_ordinal", _U32P),
        ("bbox_min", C.c_double * 3), ("bbox_max", C.c_double * 3), ("has_bbox", C.c_int32), ("reserved_", C.c_int32),
        ("res_chain_key", _U32P), ("res_res_ordinal", _U32P), ("res_atom_begin", _U32P),
        ("x_milli", _I32P), ("y_milli", _I32P), ("z_milli", _I32P),
        ("atom_gid", _U32P), ("ring_gid", _U32P), ("res_tag", _U32P),
        ("xyz_milli21", C.POINTER(C.c_uint64)), ("xyz_milli21_origin", C.c_int32 * 3), ("reserved2_", C.c_int32),
        ("attr16", C.POINTER(C.c_uint16)),
    ]


class Params(C.Structure):
    """arp_params"""
    _fields_ = [("vdw_comp", C.c_double), ("dist_cutoff", C.c_double), ("flags", C.c_uint32), ("reserved_", C.c_uint32)]


class Result(C.Structure):
    """arp_result"""
    _fields_ = [
        ("n_rows", C.c_int64), ("rows", C.c_void_p), ("n_structures", C.c_int64),
        ("atom_begin", _I64P), ("ring_begin", _I64P),
        ("pairs_screened", C.c_int64), ("pairs_in_cutoff", C.c_int64), ("pairs_classified", C.c_int64),
    ]


class Stats(C.Structure):
    """arp_stats"""
    _fields_ = [
        ("n_atoms", C.c_int64), ("n_rings", C.c_int64), ("n_cells", C.c_int64), ("n_nonempty_cells", C.c_int64),
        ("n_rows", C.c_int64), ("kernel_launches", C.c_int64), ("pair_kernel_retries", C.c_int32), ("upload_copies", C.c_int32),
        ("n_screened_pairs", C.c_int64), ("n_filtered_pairs", C.c_int64),
        ("ms_h2d", C.c_float), ("ms_grid", C.c_float), ("ms_pairs", C.c_float), ("ms_rings", C.c_float),
        ("ms_d2h", C.c_float), ("ms_run", C.c_float),
        ("ms_screen", C.c_float), ("ms_filter", C.c_float), ("ms_exact", C.c_float), ("reserved2_", C.c_float),
        ("ms_wall_upload", C.c_float), ("ms_wall_run", C.c_float), ("ms_wall_download", C.c_float), ("ms_wall_call", C.c_float),
    ]


class Table(C.Structure):
    """arph_table"""
    _fields_ = (
        [("n_rows", C.c_int64), ("model", _U32P), ("interaction", C.POINTER(C.c_uint8)), ("distance", _F32P)]
        + [(f"{s}_{c}", _I32P) for s in ("from", "to")
           for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi")]
        + [("sc_centroid_dist", _F32P), ("sc_dihedral", _F32P), ("sc_centroid_angle", _F32P),
           ("sc_valid", C.POINTER(C.c_uint8)), ("n_strings", C.c_int32), ("strings", C.POINTER(C.c_char_p)),
           ("n_atoms", C.c_int64), ("pairs_in_cutoff", C.c_int64), ("pairs_classified", C.c_int64),
           ("t_setup_s", C.c_double), ("t_search_s", C.c_double), ("t_classify_s", C.c_double),
           ("t_rings_s", C.c_double), ("t_table_s", C.c_double),
           ("sc_validity", C.POINTER(C.c_uint8)), ("sc_null_count", C.c_int64)]
    )


class FrameSide(C.Structure):
    """arp_frame_side"""
    _fields_ = [("chain", C.c_int32), ("resn", C.c_int32), ("resi", C.c_int32), ("icode", C.c_int32), ("altloc", C.c_int32),
                ("model", C.c_uint32), ("chain_rank", C.c_uint32), ("has_sc", C.c_uint32)]


class FrameTables(C.Structure):
    """arp_frame_tables"""
    _fields_ = [
        ("n_atoms", C.c_int64), ("n_rings", C.c_int64), ("n_sides", C.c_int64),
        ("sides", C.POINTER(FrameSide)), ("side_sc", _F64P), ("ent_side", _U32P), ("ent_atomn", _I32P), ("ent_atomi", _I32P),
        ("ent_rank", _U32P), ("ent_tie", _U32P), ("label_rank", C.c_uint8 * 32),
        ("n_chain_ranks", C.c_uint32), ("n_ent_ranks", C.c_uint32), ("max_tie", C.c_uint32),
    ]


class Frame(C.Structure):
    """arp_frame"""
    _fields_ = (
        [("n_rows", C.c_int64), ("model", _U32P), ("interaction", C.POINTER(C.c_uint8)), ("distance", _F32P)]
        + [(f"{s}_{c}", _I32P) for s in ("from", "to")
           for c in ("chain", "resn", "resi", "insertion", "altloc", "atomn", "atomi")]
        + [("sc_centroid_dist", _F32P), ("sc_dihedral", _F32P), ("sc_centroid_angle", _F32P),
           ("sc_valid", C.POINTER(C.c_uint8)), ("from", _U32P), ("to", _U32P),
           ("pairs_in_cutoff", C.c_int64), ("pairs_classified", C.c_int64),
           ("ms_sort", C.c_float), ("ms_columns", C.c_float), ("ms_d2h", C.c_float),
           ("sort_passes", C.c_int32), ("key_bits", C.c_int32)]
    )


class CommId(C.Structure):
    """arp_comm_id (an ncclUniqueId)"""
    _fields_ = [("bytes", C.c_char * 128)]


class SlabInfo(C.Structure):
    """arp_slab_info"""
    _fields_ = [
        ("own_atoms", C.c_int64), ("own_residues", C.c_int64), ("own_rings", C.c_int64),
        ("ghost_atoms", C.c_int64), ("ghost_residues", C.c_int64), ("ghost_rings", C.c_int64), ("ghost_h", C.c_int64),
        ("sent_atoms", C.c_int64), ("sent_residues", C.c_int64), ("sent_rings", C.c_int64),
        ("bytes_sent", C.c_int64), ("bytes_received", C.c_int64),
        ("n_peers", C.c_int32), ("transport", C.c_int32),
        ("ms_select", C.c_float), ("ms_transfer", C.c_float), ("ms_unpack", C.c_float), ("ms_total", C.c_float),
    ]


class ArpeggiaError(RuntimeError):
    """Raised for every non-zero arp_status; `.code` is the status (include/arpeggia_cuda.h)."""

    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


# every symbol include/*.h declares: (name, restype, argtypes)
_P = C.c_void_p
SIGNATURES = {
    # arpeggia_cuda.h
    "arp_label_name": (C.c_char_p, [C.c_int]),
    "arp_element_class": (C.c_int, [C.c_char_p]),
    "arp_element_radii": (C.c_int, [C.c_int, _F64P, _F64P]),
    "arp_ctx_create": (C.c_int, [C.POINTER(_P), C.c_int]),
    "arp_ctx_destroy": (None, [_P]),
    "arp_last_error": (C.c_char_p, [_P]),
    "arp_ctx_stream": (_P, [_P]),
    "arp_contacts": (C.c_int, [_P, C.POINTER(Structure), C.POINTER(Params), C.POINTER(C.POINTER(Result))]),
    "arp_contacts_batch": (C.c_int, [_P, C.POINTER(C.POINTER(Structure)), C.c_int64, C.POINTER(Params),
                                     C.POINTER(C.POINTER(Result))]),
    "arp_result_free": (None, [_P, C.POINTER(Result)]),
    "arp_stage_upload": (C.c_int, [_P, C.POINTER(C.POINTER(Structure)), C.c_int64]),
    "arp_stage_run": (C.c_int, [_P, C.POINTER(Params)]),
    "arp_stage_download": (C.c_int, [_P, C.POINTER(C.POINTER(Result))]),
    "arp_stage_stats": (C.c_int, [_P, C.POINTER(Stats)]),
    "arp_stage_neighbor_sum": (C.c_int, [_P, _P, _P, C.c_float, _P]),
    "arp_neighbor_sum_stats": (C.c_int, [_P, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_int64)]),
    "arp_stage_frame": (C.c_int, [_P, C.POINTER(FrameTables), C.POINTER(C.POINTER(Frame))]),
    "arp_stage_frame_compact": (C.c_int, [_P, C.POINTER(FrameTables), C.POINTER(C.POINTER(Frame))]),
    "arp_contacts_frame_compact": (C.c_int, [_P, C.POINTER(Structure), C.POINTER(Params), C.POINTER(FrameTables),
                                             C.POINTER(C.POINTER(Frame))]),
    "arp_contacts_frame": (C.c_int, [_P, C.POINTER(Structure), C.POINTER(Params), C.POINTER(FrameTables),
                                     C.POINTER(C.POINTER(Frame))]),
    "arp_frame_free": (None, [_P, C.POINTER(Frame)]),
    "arp_comm_get_id": (C.c_int, [C.POINTER(CommId)]),
    "arp_ctx_comm_init": (C.c_int, [_P, C.POINTER(CommId), C.c_int, C.c_int]),
    "arp_ctx_create_multi": (C.c_int, [C.POINTER(_P), C.POINTER(C.c_int), C.c_int]),
    "arp_slab_exchange": (C.c_int, [_P, C.POINTER(Params), C.POINTER(SlabInfo)]),
    "arp_slab_exchange_multi": (C.c_int, [C.POINTER(_P), C.c_int, C.POINTER(Params), C.POINTER(SlabInfo)]),
    "arp_neighbor_sum_static": (C.c_int, [C.c_int, _P, _P, _P, _P, _P, _P, C.c_int64, C.c_float, _P]),
    "arp_host_alloc": (C.c_int, [C.POINTER(_P), C.c_int64]),
    "arp_host_free": (None, [_P]),
    # arpeggia_host.h
    "arph_last_error": (C.c_char_p, []),
    "arph_parse_pdb": (C.c_int, [C.c_char_p, C.POINTER(_P), C.POINTER(C.c_int64)]),
    "arph_parse_cif": (C.c_int, [C.c_char_p, C.POINTER(_P), C.POINTER(C.c_int64)]),
    "arph_parse_cif_ids": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(_P), C.POINTER(C.c_int64)]),
    "arph_parse_file": (C.c_int, [C.c_char_p, C.POINTER(_P), C.POINTER(C.c_int64)]),
    "arph_free": (None, [_P]),
    "arph_pack": (C.c_int, [_P, C.c_int64, C.c_char_p, C.c_int, C.c_uint32, C.POINTER(_P)]),
    "arph_pack_many": (C.c_int, [C.POINTER(_P), C.POINTER(C.c_int64), C.c_int64, C.c_char_p, C.c_int, C.c_uint32, C.POINTER(_P)]),
    "arph_packed_structure": (C.POINTER(Structure), [_P]),
    "arph_packed_n_warnings": (C.c_int64, [_P]),
    "arph_packed_warning": (C.c_char_p, [_P, C.c_int64]),
    "arph_packed_free": (None, [_P]),
    "arph_packed_frame_tables": (C.c_int, [_P, C.POINTER(C.POINTER(FrameTables))]),
    "arph_packed_n_chains": (C.c_int64, [_P]),
    "arph_packed_chain_id": (C.c_char_p, [_P, C.c_int64]),
    "arph_packed_remap_chain_ranks": (C.c_int, [_P, _U32P, C.c_int64]),
    "arph_packed_set_global_ids": (C.c_int, [_P, _U32P, _U32P, _U32P]),
    "arph_table_from_frame": (C.c_int, [_P, C.POINTER(Frame), C.POINTER(C.POINTER(Table))]),
    "arph_build_table": (C.c_int, [_P, _P, C.c_int64, C.POINTER(C.POINTER(Table))]),
    "arph_table_free": (None, [C.POINTER(Table)]),
    "arph_contacts": (C.c_int, [_P, _P, C.c_int64, C.c_char_p, C.c_double, C.c_double, C.c_int, C.c_uint32,
                                C.POINTER(C.POINTER(Table))]),
    "arph_contacts_file": (C.c_int, [_P, C.c_char_p, C.c_char_p, C.c_double, C.c_double, C.c_int, C.c_uint32,
                                     C.POINTER(C.POINTER(Table))]),
    "arph_plane_fit": (C.c_int, [_P, C.c_int64, _P, _P]),
    "arph_plane_point_angle": (C.c_double, [_P, _P, _P]),
    "arph_plane_dihedral": (C.c_double, [_P, _P]),
    "arph_parse_groups": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64]),
}


def build(force: bool = False) -> Path:
    """Compile the library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", str(_PKG / "csrc"), "-j8"]
    if force:
        subprocess.run(cmd + ["clean"], check=True, capture_output=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libarpeggia_b200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return LIB_PATH


_lib = None


def lib():
    """The loaded C-ABI library.  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C arpeggia_b200/csrc). There is no CPU or pure-Python fallback."
            )
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int, ctx=None):
    if rc != 0:
        L = lib()
        msg = (L.arp_last_error(ctx) if ctx else L.arph_last_error()) or b""
        if not msg:
            msg = L.arph_last_error() or b""
        raise ArpeggiaError(rc, msg.decode(errors="replace"))
