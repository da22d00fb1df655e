"""The C-ABI library loads without a GPU, exports every symbol include/*.h declares, and fails loudly
(no CPU fallback) when a compute entry point is called without a CUDA device."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

import arpeggia_b200 as A
from arpeggia_b200 import _lib

ROOT = Path(__file__).resolve().parent.parent


def declared_functions():
    names = set()
    for h in (ROOT / "include").glob("*.h"):
        text = re.sub(r"/\*.*?\*/", "", h.read_text(), flags=re.S)
        for m in re.finditer(r"\b(arph?_[a-z0-9_]+)\s*\(", text):
            names.add(m.group(1))
    return names


def test_every_declared_symbol_is_exported_and_bound():
    L = _lib.lib()
    decl = declared_functions()
    assert len(decl) >= 30
    for name in decl:
        assert hasattr(L, name), f"{name} is declared in include/ but not exported by the library"
    assert decl == set(_lib.SIGNATURES), decl ^ set(_lib.SIGNATURES)


def test_struct_sizes_match_the_headers():
    assert C.sizeof(_lib.Params) == 24 and _lib.ROW_DTYPE.itemsize == 16 and _lib.ATOM_DTYPE.itemsize == 96
    assert C.sizeof(_lib.Structure) == 4 * 8 + 18 * 8 + 6 * 8 + 2 * 4 + 3 * 8 + 3 * 8 + 3 * 8 + (8 + 4 * 4 + 8)  # counts, pointers, bbox hint, compact forms, lattice form, ids + tags, 21-bit lattice words + 16-bit attr
    assert C.sizeof(_lib.SlabInfo) == 12 * 8 + 2 * 4 + 4 * 4 and C.sizeof(_lib.CommId) == 128
    assert C.sizeof(_lib.Result) == 8 * 8


def test_labels_and_elements():
    L = _lib.lib()
    assert [L.arp_label_name(i).decode() for i in range(19)] == _lib.LABELS
    assert L.arp_element_class(b"H") == 0 and L.arp_element_class(b"XX") == -1
    cov, vdw = C.c_double(), C.c_double()
    for sym, (c, v) in {"C": (0.75, 1.77), "N": (0.71, 1.66), "O": (0.63, 1.50), "S": (1.03, 1.89), "H": (0.32, 1.20)}.items():
        assert L.arp_element_radii(L.arp_element_class(sym.encode()), C.byref(cov), C.byref(vdw)) == 0
        assert (cov.value, vdw.value) == (c, v)


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(A.ArpeggiaError, match="no CUDA device") as e:
        A.Context(0)
    assert e.value.code == 9  # ARP_ERR_CUDA
    with pytest.raises(A.ArpeggiaError):
        A.contacts(str(ROOT / "tests" / "data" / "1ubq.pdb"))
    L = _lib.lib()
    tp = C.POINTER(_lib.Table)()
    atoms = A.parse_pdb(ROOT / "tests" / "data" / "1ubq.pdb")
    rc = L.arph_contacts(None, atoms.ctypes.data, len(atoms), b"/", 0.1, 6.5, 0, 0, C.byref(tp))
    assert rc == 9 and b"no CPU fallback" in L.arph_last_error()


def test_rust_bindings_declare_the_device_abi():
    """bindings/rust/arpeggia-cuda-sys cannot be compiled here (no Rust toolchain): keep it in step with the header by
    text -- every function of arpeggia_cuda.h is declared, every struct has the header's field count."""
    text = (ROOT / "bindings" / "rust" / "arpeggia-cuda-sys" / "src" / "lib.rs").read_text()
    header = (ROOT / "include" / "arpeggia_cuda.h").read_text()
    funcs = set(re.findall(r"\b(arp_[a-z0-9_]+)\s*\(", header))
    assert len(funcs) >= 19
    for name in funcs:
        assert re.search(rf"pub fn {name}\(", text), f"{name} missing from the Rust bindings"
    for struct, ctype in (("arp_structure", _lib.Structure), ("arp_params", _lib.Params), ("arp_result", _lib.Result),
                          ("arp_stats", _lib.Stats), ("arp_frame_side", _lib.FrameSide), ("arp_frame_tables", _lib.FrameTables),
                          ("arp_frame", _lib.Frame), ("arp_slab_info", _lib.SlabInfo)):
        body = re.search(rf"pub struct {struct} \{{(.*?)\n\}}", text, re.S).group(1)
        rust_fields = re.findall(r"pub (\w+):", body)
        assert rust_fields == [f for f, _ in ctype._fields_], struct
