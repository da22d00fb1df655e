"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same inputs.  Needs a B200."""
import ctypes
import os
import numpy as np
import pytest

import arpeggia_b200 as A
from oracle import oracle as O
from tests.util import assert_tables_equal

ROOT = __import__("pathlib").Path(__file__).resolve().parents[1]

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    return A.Context(0)


@pytest.fixture(scope="module")
def ubq(ubq_path):
    return A.parse_pdb(ubq_path)


@pytest.fixture(scope="module")
def bft(bft_path):
    return A.parse_pdb(bft_path)


def test_config1_1ubq_golden_532(ctx, ubq):
    got = ctx.contacts_table(ubq, "/", 0.1, 6.5)
    assert got["n_rows"] == 532  # python/tests/test_arpeggia.py:35
    want = O.contacts(ubq.view(O.ATOM_DTYPE), "/", 0.1, 6.5)
    assert_tables_equal(got, want, "1ubq")
    assert got["pairs_in_cutoff"] == want["pairs_in_cutoff"] == 29028
    assert got["pairs_classified"] == want["pairs_classified"] == 9128


def test_config2_6bft_bit_exact(ctx, bft):
    got = ctx.contacts_table(bft, "/", 0.1, 6.5)
    want = O.contacts(bft.view(O.ATOM_DTYPE), "/", 0.1, 6.5)
    assert_tables_equal(got, want, "6bft")
    assert got["n_rows"] == 7236
    assert got["pairs_in_cutoff"] == want["pairs_in_cutoff"] and got["pairs_classified"] == want["pairs_classified"]


@pytest.mark.parametrize("groups", ["A,B/G", "H,L/", "/C,G", "A,B,H/B,H,L", "A/A"])
def test_group_selection(ctx, bft, groups):
    got = ctx.contacts_table(bft, groups, 0.1, 6.5)
    want = O.contacts(bft.view(O.ATOM_DTYPE), groups, 0.1, 6.5)
    assert_tables_equal(got, want, groups)


@pytest.mark.parametrize("vdw_comp,rc", [(0.1, 4.0), (0.3, 5.0), (0.0, 3.0), (0.1, 9.0), (1.0, 6.5), (0.1, 12.5)])
def test_parameters(ctx, bft, vdw_comp, rc):
    got = ctx.contacts_table(bft, "/", vdw_comp, rc)
    want = O.contacts(bft.view(O.ATOM_DTYPE), "/", vdw_comp, rc)
    assert_tables_equal(got, want, f"comp={vdw_comp} rc={rc}")
    assert got["pairs_in_cutoff"] == want["pairs_in_cutoff"]
    fast = ctx.contacts_table(bft, "/", vdw_comp, rc, flags=4)  # emit-radius screen where r_emit < rc, else the same path
    assert_tables_equal(fast, want, f"comp={vdw_comp} rc={rc} (emit-radius screen)")


def test_screen_modes_give_identical_rows(ctx, bft):
    """flags only change which counters are produced, never the rows."""
    p = A.Packed(bft)
    key = ["from", "to", "label"]
    full = ctx.contacts_rows(p, 0.1, 6.5, flags=2)   # every in-cutoff pair through the exact filter
    dflt = ctx.contacts_rows(p, 0.1, 6.5, flags=0)   # default: exact stage only where a row is possible
    tight = ctx.contacts_rows(p, 0.1, 6.5, flags=1)  # screen at the emit radius
    assert np.array_equal(np.sort(full["rows"], order=key), np.sort(dflt["rows"], order=key))
    assert np.array_equal(np.sort(full["rows"], order=key), np.sort(tight["rows"], order=key))
    assert full["pairs_in_cutoff"] == dflt["pairs_in_cutoff"] == 396474
    assert full["pairs_classified"] == 124047 and dflt["pairs_classified"] == -1
    assert tight["pairs_in_cutoff"] == -1 and tight["pairs_screened"] < dflt["pairs_screened"]


def test_frame_without_counters_equals_frame_with_them(ctx, bft, ubq):
    """ARPH_FLAG_NO_COUNTERS (what arpeggia.contacts() passes): emit-radius screen on the device, the same table."""
    for atoms, tag in ((ubq, "1ubq"), (bft, "6bft")):
        full = ctx.contacts_table(atoms, "/", 0.1, 6.5)
        fast = ctx.contacts_table(atoms, "/", 0.1, 6.5, flags=4)
        assert fast["pairs_in_cutoff"] == -1 and fast["pairs_classified"] == -1 and full["pairs_in_cutoff"] > 0
        assert_tables_equal(fast, full, tag)


def test_errors_match_reference_panics(ctx, bft):
    with pytest.raises(A.ArpeggiaError, match="Invalid chain groups format"):
        ctx.contacts_table(bft, "A,B", 0.1, 6.5)
    with pytest.raises(A.ArpeggiaError, match="Empty chain groups!"):
        ctx.contacts_table(bft, "A,B,C,G,H,L/", 0.1, 6.5)
    gly = bft[np.isin(bft["resn"], [b"GLY", b"ALA"])]
    with pytest.raises(A.ArpeggiaError, match="Error building ring positions"):
        ctx.contacts_table(gly, "/", 0.1, 6.5)


def test_python_dropin_surface(ubq_path):
    df = A.contacts(ubq_path, groups="/", vdw_comp=0.1, dist_cutoff=6.5)
    n = df.height if hasattr(df, "height") else df.num_rows
    assert n == 532
    cols = list(df.columns) if not hasattr(df, "column_names") else df.column_names
    assert cols == A.COLUMNS and len(cols) == 20
    df2 = A.contacts(ubq_path, ignore_zero_occupancy=True)
    assert (df2.height if hasattr(df2, "height") else df2.num_rows) == 532


def test_python_dropin_frame_equals_the_oracle_table(bft_path, tmp_path):
    """arpeggia.contacts(file): the Arrow columns (built from the table's id columns as dictionary indices, no Python
    list) carry exactly the oracle's table -- values, nulls, dtypes of mod.rs:140-214 -- for a PDB and an mmCIF file."""
    import pyarrow as pa

    from arpeggia_b200 import synth

    bft = A.parse_pdb(bft_path)
    want = O.contacts(bft.view(O.ATOM_DTYPE), "/", 0.1, 6.5, num_threads=0)
    cif = tmp_path / "s.cif"
    synth.write_cif(bft, cif)
    for path in (bft_path, str(cif)):
        df = A.contacts(path)
        at = df if isinstance(df, pa.Table) else df.to_arrow()
        assert at.num_rows == want["n_rows"] and at.column_names == A.COLUMNS
        for c in A.COLUMNS:
            col = at[c].combine_chunks()
            if c == "model":
                assert col.type == pa.uint32()
                assert np.array_equal(col.to_numpy(), want[c] if path == bft_path else np.ones_like(want[c]))  # the cif carries model 1
            elif c.startswith("sc_"):
                assert col.type == pa.float32() and col.null_count == int((~want["sc_valid"]).sum())
                v = col.to_numpy(zero_copy_only=False)
                assert np.array_equal(np.isnan(v), ~want["sc_valid"] | np.isnan(want[c]))
                m = want["sc_valid"]
                assert np.array_equal(v[m].view(np.uint32), want[c][m].view(np.uint32))
            elif c == "distance":
                assert col.type == pa.float32() and np.array_equal(col.to_numpy().view(np.uint32), want[c].view(np.uint32))
            elif want[c].dtype == object:
                assert col.type in (pa.large_string(), pa.string()) and col.to_pylist() == want[c].tolist(), c
            else:
                assert col.type == pa.int32() and np.array_equal(col.to_numpy(), want[c]), c
    # dictionary-encoded form of the same frame
    at2, _ = A.default_context(0).contacts_file_arrow(bft_path, flags=A.FLAG_NO_COUNTERS, strings="dictionary")
    assert pa.types.is_dictionary(at2["from_chain"].type) and at2["from_chain"].to_pylist() == want["from_chain"].tolist()


def test_config3_1m_atoms_bit_exact(ctx, bft):
    """BASELINE config 3 at full size: 6bft tiled 5x5x5 = 1 022 500 atoms, table identical to the oracle's."""
    from arpeggia_b200 import synth

    atoms = synth.assembly_1m(bft)
    assert len(atoms) == 1022500
    got = ctx.contacts_table(atoms, "/", 0.1, 6.5, flags=2)
    want = O.contacts(atoms.view(O.ATOM_DTYPE), "/", 0.1, 6.5, flags=O.FLAG_SKIP_LOAD_FILTER, num_threads=0)
    assert_tables_equal(got, want, "1M")
    assert got["pairs_in_cutoff"] == want["pairs_in_cutoff"] and got["pairs_classified"] == want["pairs_classified"]


@pytest.mark.skipif(not os.environ.get("ARP_BIG_TESTS"), reason="8 M atoms against the oracle: a minute of host time; run with ARP_BIG_TESTS=1 "
                    "(last run: profiles/r2c_big_oracle_diff.txt)")
def test_8m_atoms_touching_copies_bit_exact(ctx, bft):
    """The largest inputs diffed against the ORACLE, both screen modes: 6bft tiled 10x10x10 at config
    3's pitch (8.18 M atoms; at that pitch the copies do not reach each other: every total is 1000 x the single structure)
    and 8x8x8 at the dense pitch (4.19 M atoms, neighbouring copies interpenetrate: inter-copy contacts and clashes in
    every direction).  Beyond these sizes the checks are properties (totals of config 5 = multiples of the single
    structure, slab unions = the one-GPU rows)."""
    from arpeggia_b200 import synth

    for name, atoms in (("8.18M, config-3 pitch", synth.tile(bft, 10, 10, 10, synth.PITCH_6BFT)),
                        ("4.19M, dense pitch", synth.tile(bft, 8, 8, 8, synth.PITCH_6BFT_DENSE))):
        want = O.contacts(atoms.view(O.ATOM_DTYPE), "/", 0.1, 6.5, flags=O.FLAG_SKIP_LOAD_FILTER, num_threads=0)
        got = ctx.contacts_table(atoms, "/", 0.1, 6.5, flags=2)
        assert_tables_equal(got, want, name)
        assert got["pairs_in_cutoff"] == want["pairs_in_cutoff"] and got["pairs_classified"] == want["pairs_classified"]
        emit = ctx.contacts_table(atoms, "/", 0.1, 6.5, flags=2 | A.FLAG_NO_COUNTERS)
        assert_tables_equal(emit, want, name + ", emit-radius screen")
        per_copy = got["n_rows"] / (len(atoms) / len(bft))
        print(f"{name}: {len(atoms)} atoms, {got['n_rows']} rows ({per_copy:.1f} per copy; 7236 for a lone copy), "
              f"{got['pairs_in_cutoff']} pairs in cutoff: both screen modes equal to the oracle")
        del want, got, emit


def test_dense_overlapping_copies_clashes(ctx, bft):
    """Interpenetrating copies: steric clashes, crowded cells (> 32 atoms), row-buffer growth."""
    from arpeggia_b200 import synth

    atoms = synth.tile(bft, 2, 2, 2, pitch=(20.0, 15.0, 10.0))
    got = ctx.contacts_table(atoms, "/", 0.1, 6.5, flags=2)
    want = O.contacts(atoms.view(O.ATOM_DTYPE), "/", 0.1, 6.5, flags=O.FLAG_SKIP_LOAD_FILTER, num_threads=0)
    assert_tables_equal(got, want, "dense")
    assert (got["interaction"] == "StericClash").sum() > 1000


def test_batch_equals_singles(ctx, bft):
    """arp_contacts_batch: rows of every structure equal the rows of a single-structure call."""
    from arpeggia_b200 import synth

    packs = [A.Packed(synth.jittered_complex(bft, k), "/", flags=2) for k in range(5)] + [A.Packed(bft)]
    res = ctx.contacts_rows(packs, 0.1, 6.5)
    parts = A.split_rows(res)
    assert len(parts) == len(packs)
    for p, part in zip(packs, parts):
        single = ctx.contacts_rows(p, 0.1, 6.5)["rows"]
        key = ["from", "to", "label"]
        assert np.array_equal(np.sort(part, order=key), np.sort(single, order=key))


EDGE_CASES = ["protonated", "salt_bridges", "altloc", "altloc_protonated", "insertion", "two_models", "zero_occupancy", "ties"]


@pytest.mark.parametrize("case", EDGE_CASES)
def test_edge_cases_equal_oracle(ctx, ubq, case):
    """Behaviour no reference fixture pins (hydrogens, altlocs, icodes, models, occupancy, ties): GPU == oracle."""
    from tests import fixtures as F

    atoms = F.all_cases(ubq.view(O.ATOM_DTYPE))[case]
    zero = case == "zero_occupancy"
    got = ctx.contacts_table(atoms.view(A.ATOM_DTYPE), "/", 0.1, 6.5, ignore_zero_occupancy=zero)
    want = O.contacts(atoms, "/", 0.1, 6.5, ignore_zero_occupancy=zero)
    assert_tables_equal(got, want, case)
    assert got["pairs_in_cutoff"] == want["pairs_in_cutoff"] and got["pairs_classified"] == want["pairs_classified"]
    # the path arpeggia.contacts() takes: emit-radius screen (k_screen_atoms), no counters
    fast = ctx.contacts_table(atoms.view(A.ATOM_DTYPE), "/", 0.1, 6.5, ignore_zero_occupancy=zero, flags=4)
    assert_tables_equal(fast, want, case + " (emit-radius screen)")
    assert fast["pairs_in_cutoff"] == -1
    if case == "two_models":  # and without reproducing complex.rs:447-449
        got = ctx.contacts_table(atoms.view(A.ATOM_DTYPE), "/", 0.1, 6.5, flags=1)
        want = O.contacts(atoms, "/", 0.1, 6.5, flags=O.FLAG_NO_MULTIMODEL_QUIRK)
        assert_tables_equal(got, want, case + " (no quirk)")


def test_bounding_box_hint_is_verified(ctx, bft):
    """arp_structure.bbox_* is a hint: a box that misses atoms, or no box at all, must give the same rows."""
    import ctypes as C

    from arpeggia_b200 import _lib

    packed = A.Packed(bft, "/")
    assert packed.structure.has_bbox == 1
    want = ctx.contacts_rows(packed, 0.1, 6.5)
    for variant in ("none", "too_small", "shifted"):
        st = _lib.Structure.from_buffer_copy(packed.structure)  # same pointers, own scalar fields
        if variant == "none":
            st.has_bbox = 0
        elif variant == "too_small":
            for a in range(3):
                st.bbox_max[a] = st.bbox_min[a] + 0.25 * (st.bbox_max[a] - st.bbox_min[a])
        else:
            for a in range(3):
                st.bbox_min[a] += 1000.0
                st.bbox_max[a] += 1000.0
        got = ctx.contacts_rows([C.pointer(st)], 0.1, 6.5)
        assert got["pairs_in_cutoff"] == want["pairs_in_cutoff"], variant
        a = np.sort(got["rows"], order=["from", "to", "label"])
        b = np.sort(want["rows"], order=["from", "to", "label"])
        assert np.array_equal(a, b), variant


# ---- the frame built on the device (arp_stage_frame) ------------------------------------------------
def _staged_tables(ctx, atoms, flags=0, groups="/", device_columns=False):
    """(frame built on the device, table built on the host from the same run's rows)"""
    p = A.Packed(atoms, groups, flags=flags)
    ctx.upload(p)
    ctx.run(0.1, 6.5, flags=2)
    host = p.build_table(ctx.download()["rows"])
    dev = ctx.frame(p, device_columns=device_columns)
    return dev, host, p


@pytest.mark.parametrize("columns", ["host", "device"])
@pytest.mark.parametrize("case", ["1ubq", "6bft", "altloc", "insertion", "two_models", "protonated"])
def test_device_frame_equals_host_table(ctx, ubq, bft, case, columns):
    """Sort + decode + side-chain join on the device give the very frame results_to_df / collect_sc_stats / sort give on
    the host (mod.rs:61-214): same order, same dictionary ids, same f32 bits -- with the fifteen per-entity columns
    decoded on the device (arp_stage_frame) or gathered by the host from the compact frame (arp_stage_frame_compact)."""
    from tests import fixtures as F

    atoms = {"1ubq": ubq, "6bft": bft}.get(case)
    if atoms is None:
        atoms = F.all_cases(ubq.view(O.ATOM_DTYPE))[case].view(A.ATOM_DTYPE)
    dev, host, _ = _staged_tables(ctx, atoms, device_columns=columns == "device")
    assert_tables_equal(dev, host, case)
    assert np.array_equal(dev["interaction_code"], host["interaction_code"]) and np.array_equal(dev["model"], host["model"])
    fs = dev["frame_stats"]
    assert fs["key_bits"] <= 64 and fs["sort_passes"] == (fs["key_bits"] + 7) // 8


def test_device_frame_wide_keys(ctx, bft):
    """Keys wider than 64 bits take the two-word path of the radix sort: same frame."""
    import ctypes as C
    from arpeggia_b200 import _lib

    p = A.Packed(bft)
    ctx.upload(p)
    ctx.run(0.1, 6.5, flags=2)
    host = p.build_table(ctx.download()["rows"])
    L = _lib.lib()
    tp = C.POINTER(_lib.FrameTables)()
    _lib.check(L.arph_packed_frame_tables(p._h, C.byref(tp)))
    wide = _lib.FrameTables.from_buffer_copy(tp.contents)
    wide.n_chain_ranks, wide.n_ent_ranks, wide.max_tie = 1 << 24, 1 << 28, 200  # legal upper bounds, 117-bit keys
    fp = C.POINTER(_lib.Frame)()
    _lib.check(L.arp_stage_frame(ctx._h, C.byref(wide), C.byref(fp)), ctx._h)
    assert fp.contents.key_bits == 2 * 24 + 2 * 28 + 5 + 2 * 8 and fp.contents.sort_passes == 16
    tb = C.POINTER(_lib.Table)()
    _lib.check(L.arph_table_from_frame(p._h, fp, C.byref(tb)))
    try:
        dev = A.read_table(tb.contents)
    finally:
        L.arph_table_free(tb)
    assert_tables_equal(dev, host, "wide keys")


def test_device_frame_1m_rows_sorted(ctx, bft):
    """Config 3 at full size: 0.9 M rows sorted on the device equal the host's comparator sort of the same rows."""
    from arpeggia_b200 import synth

    dev, host, p = _staged_tables(ctx, synth.assembly_1m(bft), flags=2)
    assert dev["n_rows"] == host["n_rows"] > 900000
    assert_tables_equal(dev, host, "1M frame")
    # the same rows with every column decoded on the device, and straight into Arrow (validity bitmap from the library)
    full = ctx.frame(p, device_columns=True)
    assert_tables_equal(full, host, "1M frame, device columns")
    assert full["frame_stats"]["ms_d2h"] > dev["frame_stats"]["ms_d2h"]  # 86 vs 26 bytes per row
    at, _ = ctx.frame(p, arrow="dictionary")
    assert at.num_rows == host["n_rows"] and at["sc_dihedral"].null_count == int((~host["sc_valid"]).sum())
    assert np.array_equal(at["from_atomi"].to_numpy(), host["from_atomi"]) and np.array_equal(at["to_resi"].to_numpy(), host["to_resi"])
    valid = np.asarray(at["sc_centroid_dist"].is_valid())
    assert np.array_equal(valid, host["sc_valid"])


def test_frame_narrowing_error(ctx, ubq):
    """mod.rs:150-178: an atom serial beyond i32 panics in the reference when a row carries it."""
    got = ctx.contacts_table(ubq, "/", 0.1, 6.5)
    atoms = ubq.copy()
    k = int(np.nonzero(atoms["serial"] == got["from_atomi"][0])[0][0])
    atoms["serial"][k] = (1 << 31) + 5
    with pytest.raises(A.ArpeggiaError) as e:
        ctx.contacts_table(atoms, "/", 0.1, 6.5)
    assert e.value.code == 6


def test_per_atom_and_per_residue_keys_same_rows(ctx, bft):
    """chain_key / res_ordinal per atom (slab mode, hand-packed callers) or per residue (the packer's compact form)."""
    from arpeggia_b200 import _lib

    p = A.Packed(bft)
    want = ctx.contacts_rows(p, 0.1, 6.5)  # per-residue form
    s = _lib.Structure.from_buffer_copy(p.structure)
    s.res_chain_key = s.res_res_ordinal = s.res_atom_begin = None
    got = ctx.contacts_rows([ctypes.pointer(s)], 0.1, 6.5)
    key = ["from", "to", "label"]
    assert np.array_equal(np.sort(got["rows"], order=key), np.sort(want["rows"], order=key))
    s.chain_key = None
    with pytest.raises(A.ArpeggiaError) as e:
        ctx.contacts_rows([ctypes.pointer(s)], 0.1, 6.5)
    assert e.value.code == 8


def test_lattice_coordinates_same_rows(ctx, bft):
    """x / y / z as f64, or on the 0.001 A lattice as int32 (arp_structure.x_milli, what the packer offers for anything
    read from PDB columns): the device rebuilds the same doubles, rows and counters are identical."""
    from arpeggia_b200 import _lib

    p = A.Packed(bft)
    assert p.structure.x_milli and p.structure.y_milli and p.structure.z_milli
    a = p.arrays()
    for k in "xyz":
        assert np.array_equal(a[k + "_milli"] / 1000.0, a[k])  # one IEEE division gives back the parsed double
    assert p.structure.xyz_milli21 and p.structure.attr16
    s = _lib.Structure.from_buffer_copy(p.structure)
    s.x = s.y = s.z = None  # only the lattice forms: the 21-bit words win
    s.x_milli = s.y_milli = s.z_milli = None
    s.attr = None  # ... and only the 16-bit attribute words
    lat21 = ctx.contacts_rows([ctypes.pointer(s)], 0.1, 6.5)
    s = _lib.Structure.from_buffer_copy(p.structure)
    s.x = s.y = s.z = None  # only the 32-bit lattice form
    s.xyz_milli21 = None
    s.attr16 = None
    lat = ctx.contacts_rows([ctypes.pointer(s)], 0.1, 6.5)
    s = _lib.Structure.from_buffer_copy(p.structure)
    s.x_milli = s.y_milli = s.z_milli = None  # only f64
    s.xyz_milli21 = None
    s.attr16 = None
    f64 = ctx.contacts_rows([ctypes.pointer(s)], 0.1, 6.5)
    key = ["from", "to", "label", "distance"]
    assert np.array_equal(np.sort(lat["rows"], order=key), np.sort(f64["rows"], order=key))
    assert np.array_equal(np.sort(lat21["rows"], order=key), np.sort(f64["rows"], order=key))
    assert lat["pairs_in_cutoff"] == f64["pairs_in_cutoff"] == lat21["pairs_in_cutoff"] == 396474
    s.x = None
    with pytest.raises(A.ArpeggiaError) as e:
        ctx.contacts_rows([ctypes.pointer(s)], 0.1, 6.5)
    assert e.value.code == 8


def test_emit_radius_screen_rows_at_scale(ctx, bft):
    """ARP_FLAG_EMIT_RADIUS_SCREEN (k_screen_atoms, cells at r_emit): the same rows as the default screen on config 3,
    on interpenetrating copies (crowded cells), on a batch job and for parameters that move r_emit."""
    from arpeggia_b200 import synth

    key = ["from", "to", "label", "distance"]

    def same(packed, comp=0.1, rc=6.5):
        a = ctx.contacts_rows(packed, comp, rc, flags=0)
        b = ctx.contacts_rows(packed, comp, rc, flags=1)
        assert len(a["rows"]) == len(b["rows"]) > 0
        assert np.array_equal(np.sort(a["rows"], order=key), np.sort(b["rows"], order=key))
        return a, b

    a, b = same(A.Packed(synth.assembly_1m(bft), "/", flags=2))
    assert len(a["rows"]) == 904500 and b["pairs_in_cutoff"] == -1 and b["pairs_screened"] < a["pairs_screened"] / 2
    same(A.Packed(synth.tile(bft, 2, 2, 2, pitch=(20.0, 15.0, 10.0)), "/", flags=2))
    same([A.Packed(synth.jittered_complex(bft, k), "/", flags=2) for k in range(7)] + [A.Packed(bft)])
    p = A.Packed(bft)
    for comp, rc in [(0.3, 5.0), (1.0, 6.5), (0.1, 12.5), (0.1, 4.6), (0.0, 3.0)]:
        same(p, comp, rc)
    for groups in ["A,B/G", "H,L/", "A/A"]:
        same(A.Packed(bft, groups))


def test_config4_batch_of_1000_complexes(ctx, bft):
    """BASELINE config 4 at a tenth of its size (scratch/batch.py runs the 10 000): one arp_contacts_batch job; a
    structure's rows in the batch are exactly the rows of that structure alone, the counters add up."""
    from arpeggia_b200 import synth

    packs = [A.Packed(synth.jittered_complex(bft, k), "/", flags=2) for k in range(1000)]
    res = ctx.contacts_rows(packs, 0.1, 6.5)
    parts = A.split_rows(res)
    assert len(parts) == 1000 and sum(len(p) for p in parts) == len(res["rows"])
    key = ["from", "to", "label"]
    total_in = 0
    for k in (0, 1, 499, 998, 999):
        single = ctx.contacts_rows(packs[k], 0.1, 6.5)
        assert np.array_equal(np.sort(parts[k], order=key), np.sort(single["rows"], order=key)), k
        total_in += single["pairs_in_cutoff"]
    assert abs(res["pairs_in_cutoff"] / 1000 - total_in / 5) < 0.01 * total_in / 5  # jitter moves a few pairs across the cutoff


def test_config5_100m_atoms_properties(ctx, bft):
    """BASELINE config 5 at full size on ONE GPU: 6bft tiled 16 x 32 x 24 = 100 515 840 atoms.  The copies sit more than
    dist_cutoff apart, so every counter must be exactly 12 288 times the single-structure value (size-independent
    property; the oracle would need minutes and the table gigabytes)."""
    import psutil
    import torch
    from arpeggia_b200 import synth

    if psutil.virtual_memory().total < 100 << 30 or torch.cuda.get_device_properties(0).total_memory < 100 << 30:
        pytest.skip("needs ~60 GB of host memory and ~30 GB of HBM")
    p1 = A.Packed(bft)
    one = ctx.contacts_rows(p1, 0.1, 6.5)
    n_at, n_rg = p1.structure.n_atoms, p1.structure.n_rings
    copies = 16 * 32 * 24
    p = A.Packed(synth.assembly_100m(bft), "/", flags=2)
    assert p.arrays(copy=False)["n_atoms"] == 100515840
    res = ctx.contacts_rows(p, 0.1, 6.5)
    del p
    assert len(res["rows"]) == copies * len(one["rows"]) == copies * 7236
    assert res["pairs_in_cutoff"] == copies * one["pairs_in_cutoff"] == copies * 396474
    assert np.array_equal(np.bincount(res["rows"]["label"], minlength=19), copies * np.bincount(one["rows"]["label"], minlength=19))
    # every copy's rows are the single structure's rows shifted by the copy's atom / ring offset
    rows = res["rows"]
    ring = (rows["from"] >> 31) == 1
    copy_of = np.where(ring, (rows["from"] & 0x7FFFFFFF) // n_rg, rows["from"] // n_at)
    assert np.array_equal(np.bincount(copy_of, minlength=copies), np.full(copies, 7236))


def test_frame_argument_errors(ctx, ubq, bft):
    """arp_stage_frame refuses what it cannot describe: batch jobs, tables of another structure, rows outside the tables."""
    import ctypes as C
    from arpeggia_b200 import _lib

    L = _lib.lib()
    pu, pb = A.Packed(ubq), A.Packed(bft)
    tu = C.POINTER(_lib.FrameTables)()
    _lib.check(L.arph_packed_frame_tables(pu._h, C.byref(tu)))
    fp = C.POINTER(_lib.Frame)()
    ctx.upload([pu, pb])
    ctx.run(0.1, 6.5)
    assert L.arp_stage_frame(ctx._h, tu, C.byref(fp)) == 8 and not fp  # a frame belongs to one structure
    ctx.upload(pb)
    ctx.run(0.1, 6.5)
    assert L.arp_stage_frame(ctx._h, tu, C.byref(fp)) == 8 and not fp  # 1ubq's tables for 6bft's rows
    ctx.upload(pu)
    ctx.run(0.1, 6.5)
    bad = _lib.FrameTables.from_buffer_copy(tu.contents)
    bad.n_sides = 3  # rows reference sides beyond the table: caught on the device, no out-of-bounds read
    assert L.arp_stage_frame(ctx._h, C.byref(bad), C.byref(fp)) == 8 and not fp
    assert b"outside the frame tables" in L.arp_last_error(ctx._h)
    _lib.check(L.arp_stage_frame(ctx._h, tu, C.byref(fp)), ctx._h)  # and the context is still usable
    assert fp.contents.n_rows == 532
    L.arp_frame_free(ctx._h, fp)


def test_contexts_on_several_host_threads_same_rows(ubq, bft):
    """One arp_ctx per host thread (include/arpeggia_cuda.h): jobs of different contexts overlap on the device and must
    not see each other -- every thread gets the rows a lone context gets, for two different structures in alternation."""
    import threading

    packed = [A.Packed(ubq, "/"), A.Packed(bft, "/")]
    lone = A.Context(0)
    want = [np.sort(lone.contacts_rows(p, 0.1, 6.5, flags=f)["rows"], order=["from", "to", "label"]) for p in packed for f in (0, 1)]
    errors = []

    def worker(k):
        try:
            c = A.Context(0)
            for it in range(12):
                j = (it + k) % 4
                got = np.sort(c.contacts_rows(packed[j // 2], 0.1, 6.5, flags=j % 2)["rows"], order=["from", "to", "label"])
                if not np.array_equal(got, want[j]):
                    errors.append((k, it, j, len(got), len(want[j])))
            c.close()
        except Exception as e:  # noqa: BLE001
            errors.append((k, repr(e)))

    th = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errors, errors


@pytest.mark.parametrize("k", [0, 1, 499, 4242, 9999])
def test_config4_complexes_equal_the_oracle(ctx, bft, k):
    """BASELINE config 4: complex k of the batch (chains A, B, C, G of 6bft, jittered: other ties and labels than 6bft)
    against the ORACLE, full sorted 20-column table, both screen modes -- and its rows inside a batch job are the rows the
    oracle-checked table was built from."""
    from arpeggia_b200 import synth

    atoms = synth.jittered_complex(bft, k)
    want = O.contacts(atoms.view(O.ATOM_DTYPE), "/", 0.1, 6.5, flags=O.FLAG_SKIP_LOAD_FILTER, num_threads=0)
    assert want["n_rows"] > 3000
    for flags in (2, 2 | 4):  # counting mode / emit-radius screen
        assert_tables_equal(ctx.contacts_table(atoms, "/", 0.1, 6.5, flags=flags), want, f"complex {k} flags {flags}")
    # the same complex as one structure of a batch job: its rows, decoded and sorted by the host table, are the oracle's table
    others = [A.Packed(synth.jittered_complex(bft, j), "/", flags=2) for j in (k + 1, k + 2)]
    mine = A.Packed(atoms, "/", flags=2)
    res = ctx.contacts_rows([others[0], mine, others[1]], 0.1, 6.5, 1)
    assert_tables_equal(mine.build_table(A.split_rows(res)[1]), want, f"complex {k} inside a batch")


def test_results_own_their_rows(ctx, bft, ubq):
    """an arp_result keeps its rows while later calls run on the same context (each result owns its page-locked
    buffer; freed buffers are pooled for the next download)."""
    import ctypes as C

    from arpeggia_b200 import _lib as AL

    L = AL.lib()
    prm = AL.Params(0.1, 6.5, 0, 0)
    held = []
    for atoms in (bft, ubq, bft):
        p = A.Packed(atoms)
        rp = C.POINTER(AL.Result)()
        AL.check(L.arp_contacts(ctx._h, p.structure_ptr, C.byref(prm), C.byref(rp)), ctx._h)
        held.append(rp)
    views = [np.frombuffer((C.c_char * (r.contents.n_rows * 16)).from_address(r.contents.rows), dtype=A.ROW_DTYPE) for r in held]
    assert [len(v) for v in views] == [7236, 532, 7236] and len({r.contents.rows for r in held}) == 3
    key = ["from", "to", "label", "distance"]
    assert np.array_equal(np.sort(views[0], order=key), np.sort(views[2], order=key))
    assert views[1]["from"].max() < 660  # still the 1ubq rows, not overwritten by the third call
    for r in held:
        L.arp_result_free(ctx._h, r)


def test_arrays_of_one_host_arena_travel_as_one_copy(ctx, bft, ubq):
    """The upload merges arrays that follow one another in the caller's memory into one host->device copy
    (arp_stats.upload_copies) and moves them to their buffers on the device: the rows are those of the separate copies,
    whatever the order and the padding of the arrays in the arena, with and without the on-lattice coordinates."""
    import ctypes as C

    from arpeggia_b200 import _lib as AL

    L = AL.lib()
    prm = AL.Params(0.1, 6.5, 0, 0)
    key = ["from", "to", "label", "distance"]
    n_of = {"n_atoms": 1, "n_residues": 1, "n_rings": 1, "n_h": 1}

    def rows_of(sp):
        rp = C.POINTER(AL.Result)()
        AL.check(L.arp_contacts(ctx._h, sp, C.byref(prm), C.byref(rp)), ctx._h)
        v = np.frombuffer((C.c_char * (rp.contents.n_rows * 16)).from_address(rp.contents.rows), dtype=A.ROW_DTYPE).copy()
        L.arp_result_free(ctx._h, rp)
        return np.sort(v, order=key), ctx.stats()["upload_copies"]

    for atoms, milli in ((bft, True), (ubq, True), (bft, False), (ubq, "21")):
        p = A.Packed(atoms)
        src = p.structure
        n, nr, ng, nh = src.n_atoms, src.n_residues, src.n_rings, src.n_h
        want, copies_plain = rows_of(p.structure_ptr)
        lens = {"attr": 4 * n, "res_chain_key": 4 * nr, "res_res_ordinal": 4 * nr, "res_atom_begin": 4 * (nr + 1), "res_h_begin": 4 * (nr + 1),
                "h_x": 8 * nh, "h_y": 8 * nh, "h_z": 8 * nh, "res_cb": 4 * nr, "res_sg": 4 * nr, "ring_center": 24 * ng, "ring_normal": 24 * ng,
                "ring_attr": 4 * ng, "ring_chain_key": 4 * ng, "ring_res_ordinal": 4 * ng, "res_tag": 4 * nr}
        if milli == "21":  # the most compact forms: 8-byte lattice words, 16-bit attributes
            del lens["attr"]
            lens.update({"xyz_milli21": 8 * n, "attr16": 2 * n})
        else:
            lens.update({"x_milli": 4 * n, "y_milli": 4 * n, "z_milli": 4 * n} if milli else {"x": 8 * n, "y": 8 * n, "z": 8 * n})
        assert bool(src.x_milli) and bool(src.res_tag)
        names = list(lens)
        np.random.default_rng(5).shuffle(names)  # the arena's order is not the library's upload order
        pads = [4, 16, 256, 1000, 0]
        block = C.c_void_p()
        AL.check(L.arp_host_alloc(C.byref(block), sum(lens.values()) + 1024 * len(lens) + 64))
        try:
            dst = AL.Structure()
            for name, _ in AL.Structure._fields_:
                if name.startswith("n_") or name.startswith("bbox_") or name == "has_bbox" or name == "xyz_milli21_origin":
                    setattr(dst, name, getattr(src, name))
            at = 0
            for k, name in enumerate(names):
                ptr = C.c_void_p(block.value + at)
                if lens[name]:
                    C.memmove(ptr, C.cast(getattr(src, name), C.c_void_p), lens[name])
                setattr(dst, name, C.cast(ptr, dict(AL.Structure._fields_)[name]))
                at += lens[name] + pads[k % len(pads)]
                at = (at + 7) // 8 * 8  # f64 arrays stay 8-byte aligned
            got, copies = rows_of(C.pointer(dst))
            assert np.array_equal(got, want), (milli, len(got), len(want))
            assert copies == 1 and copies_plain >= 1, (copies, copies_plain)  # (the packer's own vectors may happen to be neighbours too)
        finally:
            L.arp_host_free(block)
        # every array in a page-locked allocation of its own: such allocations often lie back to back in the address space,
        # but one copy must not run from one into the next (cudaMemcpyAsync refuses it) -- the upload asks for the extent of
        # the allocation before it merges
        blocks = []
        try:
            dst = AL.Structure()
            for name, _ in AL.Structure._fields_:
                if name.startswith("n_") or name.startswith("bbox_") or name == "has_bbox" or name == "xyz_milli21_origin":
                    setattr(dst, name, getattr(src, name))
            for name in names:
                b = C.c_void_p()
                AL.check(L.arp_host_alloc(C.byref(b), max(lens[name], 1)))
                blocks.append(b)
                if lens[name]:
                    C.memmove(b, C.cast(getattr(src, name), C.c_void_p), lens[name])
                setattr(dst, name, C.cast(b, dict(AL.Structure._fields_)[name]))
            got, copies = rows_of(C.pointer(dst))
            assert np.array_equal(got, want) and copies >= 2, (milli, copies)
        finally:
            for b in blocks:
                L.arp_host_free(b)


def test_device_resident_and_mixed_inputs(ctx, bft):
    """arp_structure arrays may live in device memory (a caller that produced them on the GPU), or some on the host and
    some on the device: never merged into a host copy, same rows."""
    import ctypes as C

    import torch

    from arpeggia_b200 import _lib as AL

    L = AL.lib()
    prm = AL.Params(0.1, 6.5, 0, 0)
    key = ["from", "to", "label", "distance"]
    p = A.Packed(bft)
    src = p.structure

    def rows_of(sp):
        rp = C.POINTER(AL.Result)()
        AL.check(L.arp_contacts(ctx._h, sp, C.byref(prm), C.byref(rp)), ctx._h)
        v = np.frombuffer((C.c_char * (rp.contents.n_rows * 16)).from_address(rp.contents.rows), dtype=A.ROW_DTYPE).copy()
        L.arp_result_free(ctx._h, rp)
        return np.sort(v, order=key)

    want = rows_of(p.structure_ptr)
    n, nr, ng, nh = src.n_atoms, src.n_residues, src.n_rings, src.n_h
    lens = {"x": 8 * n, "y": 8 * n, "z": 8 * n, "attr": 4 * n, "res_chain_key": 4 * nr, "res_res_ordinal": 4 * nr, "res_atom_begin": 4 * (nr + 1),
            "res_h_begin": 4 * (nr + 1), "h_x": 8 * nh, "h_y": 8 * nh, "h_z": 8 * nh, "res_cb": 4 * nr, "res_sg": 4 * nr, "ring_center": 24 * ng,
            "ring_normal": 24 * ng, "ring_attr": 4 * ng, "ring_chain_key": 4 * ng, "ring_res_ordinal": 4 * ng, "res_tag": 4 * nr}
    for on_device in (set(lens), {"x", "z", "attr", "res_tag", "ring_center"}):
        dst = AL.Structure()
        for name, _ in AL.Structure._fields_:
            if name.startswith("n_") or name.startswith("bbox_") or name == "has_bbox":
                setattr(dst, name, getattr(src, name))
        keep = []
        for name, nbytes in lens.items():
            ptr = C.cast(getattr(src, name), C.c_void_p)
            if name in on_device and nbytes:
                host = np.frombuffer((C.c_char * nbytes).from_address(ptr.value), dtype=np.uint8)
                t = torch.from_numpy(host.copy()).to("cuda:0")
                keep.append(t)
                ptr = C.c_void_p(t.data_ptr())
            setattr(dst, name, C.cast(ptr, dict(AL.Structure._fields_)[name]))
        torch.cuda.synchronize()
        got = rows_of(C.pointer(dst))
        assert np.array_equal(got, want), sorted(on_device)[:3]


def test_cli_contacts_writes_the_table(ubq_path, tmp_path, caplog):
    """`python -m arpeggia_b200 contacts` (the flag surface of src/cli/contacts.rs:9-52): the four output formats hold the
    532 rows of 1ubq; an extension in --filename is replaced like Path::with_extension; a missing input is logged, not raised."""
    import json

    import pyarrow.csv as pacsv
    import pyarrow.parquet as pq

    from arpeggia_b200.__main__ import main

    want = O.contacts(A.parse_pdb(ubq_path).view(O.ATOM_DTYPE))
    out = tmp_path / "out"
    assert main(["contacts", "-i", str(ubq_path), "-o", str(out)]) == 0
    t = pacsv.read_csv(out / "contacts.csv")
    assert t.column_names == A.COLUMNS and t.num_rows == want["n_rows"] == 532
    try:
        import polars  # noqa: F401  (with polars the file is polars' own; the golden pins this repository's writer)
    except ImportError:
        assert (out / "contacts.csv").read_text() == (ROOT / "tests" / "golden" / "1ubq_contacts.csv").read_text()
    assert t["from_atomi"].to_pylist() == want["from_atomi"].tolist() and t["interaction"].to_pylist() == want["interaction"].tolist()
    assert main(["contacts", "-i", str(ubq_path), "-o", str(out), "-t", "parquet", "-f", "table.old", "-g", "/", "-d", "6.5", "-c", "0.1", "-j", "2"]) == 0
    t = pq.read_table(out / "table.parquet")  # "table.old" -> "table.parquet"
    assert t.num_rows == 532 and t["sc_dihedral"].null_count == int((~want["sc_valid"]).sum())
    assert np.array_equal(t["distance"].to_numpy().view(np.uint32), want["distance"].view(np.uint32))
    assert main(["contacts", "-i", str(ubq_path), "-o", str(out), "-t", "json", "--ignore-zero-occupancy"]) == 0
    rows = json.loads((out / "contacts.json").read_text())
    assert len(rows) == 532 and rows[0]["from_resn"] == want["from_resn"][0] and list(rows[0]) == A.COLUMNS
    assert main(["contacts", "-i", str(ubq_path), "-o", str(out), "-t", "ndjson"]) == 0
    lines = (out / "contacts.ndjson").read_text().splitlines()
    assert len(lines) == 532 and json.loads(lines[-1])["to_atomi"] == int(want["to_atomi"][-1])
    with caplog.at_level("ERROR"):
        assert main(["contacts", "-i", str(tmp_path / "nope.pdb"), "-o", str(out)]) == 0  # cli/contacts.rs:58-64: logged, no panic
    assert not (out / "nope.csv").exists()
