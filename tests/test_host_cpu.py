"""Host side of the product (parser, packer, plane fit, table builder) against the oracle.  CPU only."""
import os
import collections

import numpy as np
import pytest

import arpeggia_b200 as A
from arpeggia_b200 import _lib
from oracle import oracle as O
from tests.util import assert_tables_equal

ROOT = __import__("pathlib").Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def ubq(ubq_path):
    return A.parse_pdb(ubq_path)


@pytest.fixture(scope="module")
def bft(bft_path):
    return A.parse_pdb(bft_path)


def test_parsers_agree(ubq_path, bft_path):
    for p in (ubq_path, bft_path):
        a, b = A.parse_pdb(p), O.parse_pdb(p)
        assert a.tobytes() == b.tobytes()
    assert len(A.parse_pdb(ubq_path)) == 660 and len(A.parse_pdb(bft_path)) == 8180


def test_g5_g6_g7_through_the_product_host_code(ubq):
    import ctypes as C
    L = _lib.lib()
    ring = ubq[(ubq["resi"] == 4) & np.isin(ubq["name"], [b"CG", b"CD1", b"CD2", b"CE1", b"CE2", b"CZ"])]
    pts = np.ascontiguousarray(np.stack([ring["x"], ring["y"], ring["z"]], 1))
    c, n = np.zeros(3), np.zeros(3)
    assert L.arph_plane_fit(pts.ctypes.data, 6, c.ctypes.data, n.ctypes.data) == 0
    np.testing.assert_allclose(c, [24.96883333, 34.687, 6.16233333], rtol=1e-6)  # residues.rs:355-362
    ref = np.array([0.53253994, -0.82736044, -0.17853828])
    np.testing.assert_allclose(n * np.sign(n @ ref), ref, rtol=1e-6)
    oc, on = O.plane_fit(pts)  # same recipe in both code bases -> same bits
    assert np.array_equal(oc, c) and np.array_equal(on, n)
    o, z, pt, x = np.zeros(3), np.array([0.0, 0.0, 1.0]), np.array([0.0, 1.0, 1.0]), np.array([1.0, 0.0, 0.0])
    ang = lambda cc, nn, p: L.arph_plane_point_angle(cc.ctypes.data, nn.ctypes.data, p.ctypes.data)
    assert abs(ang(o, z, pt) - 45.0) < 1e-6 and abs(ang(pt, x, o) - 90.0) < 1e-6  # residues.rs:306-332
    assert L.arph_plane_dihedral(z.ctypes.data, (-z).ctypes.data) < 1e-6
    lig, rec = C.create_string_buffer(64), C.create_string_buffer(64)
    cases = {"A,B/C,D": ("A,B", "C,D"), "A/C,D": ("A", "C,D"), "/C,D": ("A,B", "C,D"), "C/": ("C", "A,B,D"), "/": ("A,B,C,D", "A,B,C,D")}
    for g, (l, r) in cases.items():  # utils.rs:174-209
        assert L.arph_parse_groups(b"A,B,C,D", g.encode(), lig, rec, 64) == 0
        assert (lig.value.decode(), rec.value.decode()) == (l, r)
    assert L.arph_parse_groups(b"A,B,C,D", b"A,B", lig, rec, 64) == 2
    assert b"Invalid chain groups format! Use '/' for all-to-all comparisons." in L.arph_last_error()
    assert L.arph_parse_groups(b"A,B,C,D", b"A,B,C,D/", lig, rec, 64) == 3 and b"Empty chain groups!" in L.arph_last_error()


def test_packer_1ubq(ubq):
    p = A.Packed(ubq)
    a = p.arrays()
    assert (a["n_atoms"], a["n_residues"], a["n_rings"], a["n_h"]) == (660, 134, 4, 0)
    # chains.rs:21-40: 76 residues then 58 waters in the same chain -> ordinals 0..133, waters last
    assert a["res_ordinal"].max() == 133 and (np.diff(a["res_ordinal"].astype(int)) >= 0).all()
    hoh = ubq["resn"] == b"HOH"
    assert a["res_ordinal"][hoh].min() == 76
    attr = a["attr"]
    el = attr & 0xF
    assert collections.Counter(el.tolist()) == collections.Counter(
        {b"C": 1, b"N": 2, b"O": 3, b"S": 4}[e] for e in ubq["element"])
    assert ((attr & 0x30) == 0x30).all()  # "/" -> every chain is ligand and receptor
    name, resn = ubq["name"], ubq["resn"]
    flag = lambda bit: (attr & bit) != 0
    assert np.array_equal(flag(0x200), (name == b"N") | ((resn == b"ARG") & np.isin(name, [b"NE", b"NH1", b"NH2"]))
                          | ((resn == b"ASN") & (name == b"ND2")) | ((resn == b"GLN") & (name == b"NE2"))
                          | ((resn == b"HIS") & np.isin(name, [b"ND1", b"NE2"])) | ((resn == b"LYS") & (name == b"NZ"))
                          | ((resn == b"SER") & (name == b"OG")) | ((resn == b"THR") & (name == b"OG1"))
                          | ((resn == b"TRP") & (name == b"NE1")) | ((resn == b"TYR") & (name == b"OH")) | ((resn == b"CYS") & (name == b"SG")))
    assert not flag(0x400)[hoh].any()  # water O is no acceptor (hbond.rs:139-142)
    assert flag(0x400)[(name == b"O") & ~hoh].all()
    assert np.array_equal(flag(0x800), (ubq["element"] == b"C") & (name != b"C"))
    assert flag(0x1000)[(name == b"CB") & (resn != b"SER")].all() and not flag(0x1000)[(name == b"CB") & (resn == b"SER")].any()
    assert np.array_equal(flag(0x40), flag(0x100))  # conformer name == residue name in standard files
    # ring planes equal the oracle's, bit for bit
    for k, resi in enumerate([4, 45, 59, 68]):  # PHE4 PHE45 TYR59 HIS68
        n, c, nrm = O.residue_plane(ubq.view(O.ATOM_DTYPE), "A", resi, "ring")
        assert np.array_equal(a["ring_center"][k], c) and np.array_equal(a["ring_normal"][k], nrm)


def test_packer_groups_and_errors(bft):
    a = A.Packed(bft, "A,B/G").arrays()
    chains = bft["chain"]
    assert np.array_equal((a["attr"] & 0x10) != 0, np.isin(chains, [b"A", b"B"]))
    assert np.array_equal((a["attr"] & 0x20) != 0, chains == b"G")
    ranks = {c: i for i, c in enumerate(sorted(set(chains.tolist())))}  # complex.rs:129 string order
    assert np.array_equal(a["chain_key"], np.array([ranks[c] for c in chains.tolist()], dtype=np.uint32))
    cys = A.Packed(bft).arrays()
    sg = np.nonzero((bft["resn"] == b"CYS") & (bft["name"] == b"SG"))[0]
    assert np.array_equal(np.sort(cys["res_sg"][cys["res_sg"] >= 0]), sg) and ((cys["attr"][sg] & 0x2000) != 0).all()
    with pytest.raises(A.ArpeggiaError, match="Invalid chain groups format") as e:
        A.Packed(bft, "A,B")
    assert e.value.code == 2
    with pytest.raises(A.ArpeggiaError, match="Empty chain groups!"):
        A.Packed(bft, "A,B,C,G,H,L/")
    with pytest.raises(A.ArpeggiaError, match="Error building ring positions"):
        A.Packed(bft[np.isin(bft["resn"], [b"GLY", b"ALA", b"LEU"])])
    bad = bft.copy()
    bad["element"][5] = b"XX"
    with pytest.raises(A.ArpeggiaError, match="van der Waals"):
        A.Packed(bad)
    w = A.Packed(bft).warnings  # HIS A230 has no side chain: warning, not an error (mod.rs:64-68)
    assert len(w) == 1 and "230" in w[0]


def rows_from_oracle_table(atoms, packed, t):
    """Invert the oracle's table into device rows (atom / ring indices) to drive arph_build_table alone."""
    a = packed.arrays()
    serial_to_idx = {int(s): i for i, s in enumerate(atoms["serial"])}
    chains_sorted = sorted(set(atoms["chain"].tolist()))
    ring_idx = {}
    res_seq = {}
    for r in atoms:  # ordinal -> resi per chain (fixtures have no insertion codes)
        key = r["chain"]
        seq = res_seq.setdefault(key, [])
        if not seq or seq[-1] != int(r["resi"]):
            seq.append(int(r["resi"]))
    for k in range(a["n_rings"]):
        ch = chains_sorted[int(a["ring_chain_key"][k]) & 0xFFFFFF]
        ring_idx[(ch.decode(), res_seq[ch][int(a["ring_res_ordinal"][k])])] = k
    rows = np.zeros(t["n_rows"], dtype=_lib.ROW_DTYPE)
    for side in ("from", "to"):
        isring = t[f"{side}_atomn"] == "Ring"
        idx = np.array([ring_idx[(c, int(r))] | _lib.RING_FLAG if g else serial_to_idx[int(s)]
                        for c, r, s, g in zip(t[f"{side}_chain"], t[f"{side}_resi"], t[f"{side}_atomi"], isring)], dtype=np.uint32)
        rows[side] = idx
    rows["label"] = t["interaction_code"]
    rows["distance"] = t["distance"]
    return rows


@pytest.mark.parametrize("which", ["ubq", "bft"])
def test_table_builder_equals_oracle(which, ubq, bft):
    atoms = {"ubq": ubq, "bft": bft}[which]
    want = O.contacts(atoms.view(O.ATOM_DTYPE))
    p = A.Packed(atoms)
    rows = rows_from_oracle_table(atoms, p, want)
    rng = np.random.default_rng(7)
    got = p.build_table(rows[rng.permutation(len(rows))])  # device rows arrive in arbitrary order
    # the validity of the sc_* columns also comes as a bitmap in Arrow's layout
    bits = np.unpackbits(got["sc_validity_bits"], bitorder="little")[: got["n_rows"]].astype(bool)
    assert np.array_equal(bits, got["sc_valid"]) and got["sc_null_count"] == int((~got["sc_valid"]).sum())
    assert_tables_equal(got, want, which)
    at = A.to_arrow(got)
    assert at.column_names == A.COLUMNS and at.num_rows == want["n_rows"]
    import pyarrow as pa
    assert at.schema.field("model").type == pa.uint32() and at.schema.field("distance").type == pa.float32()
    assert at.schema.field("from_resi").type == pa.int32() and at.schema.field("interaction").type == pa.large_string()
    assert at["sc_dihedral"].null_count == int((~want["sc_valid"]).sum())


def test_table_builder_rejects_bad_rows(ubq):
    p = A.Packed(ubq)
    rows = np.zeros(1, dtype=_lib.ROW_DTYPE)
    rows["from"] = 100000
    with pytest.raises(A.ArpeggiaError, match="outside the structure"):
        p.build_table(rows)


@pytest.mark.parametrize("which", ["1ubq", "6bft", "altloc", "insertion", "two_models"])
def test_per_residue_keys_expand_to_the_per_atom_arrays(which, ubq, bft):
    """arp_structure.res_chain_key / res_res_ordinal (what the device path uploads) expand to chain_key / res_ordinal."""
    from tests import fixtures as F

    atoms = {"1ubq": ubq, "6bft": bft}.get(which)
    if atoms is None:
        atoms = F.all_cases(ubq.view(O.ATOM_DTYPE))[which].view(A.ATOM_DTYPE)
    a = A.Packed(atoms).arrays()
    assert len(a["res_chain_key"]) == a["n_residues"] == len(a["res_res_ordinal"])
    assert np.array_equal(a["res_chain_key"][a["res_id"]], a["chain_key"])
    assert np.array_equal(a["res_res_ordinal"][a["res_id"]], a["res_ordinal"])
    b = a["res_atom_begin"]  # res_id as a CSR
    assert b is not None and b[0] == 0 and b[-1] == a["n_atoms"]
    assert np.array_equal(np.repeat(np.arange(a["n_residues"], dtype=np.uint32), np.diff(b.astype(np.int64))), a["res_id"])


def test_lattice_form_of_the_coordinates(ubq, bft):
    """arp_structure.x_milli: offered only when (double)v / 1000.0 reproduces every coordinate bit for bit."""
    from arpeggia_b200 import synth

    for atoms in (ubq, bft, synth.tile(bft, 2, 1, 1), synth.jittered_complex(bft, 3)):
        a = A.Packed(atoms, "/", flags=2).arrays()
        for k in "xyz":
            m = a[k + "_milli"]
            assert m is not None and m.dtype == np.int32
            assert np.array_equal((m.astype(np.float64) / 1000.0).view(np.uint64), a[k].view(np.uint64))
        # the 21-bit words (arp_structure.xyz_milli21) hold the same lattice values relative to the box corner, the 16-bit
        # attribute words the same attributes
        w, o = a["xyz_milli21"], a["xyz_milli21_origin"].astype(np.int64)
        assert w is not None and w.dtype == np.uint64
        for j, k in enumerate("xyz"):
            assert np.array_equal(((w >> np.uint64(21 * j)) & np.uint64(0x1FFFFF)).astype(np.int64) + o[j], a[k + "_milli"].astype(np.int64))
            assert o[j] == a[k + "_milli"].min()
        assert (w >> np.uint64(63)).max() == 0
        assert a["attr16"] is not None and np.array_equal(a["attr16"].astype(np.uint32), a["attr"])
    wide = np.concatenate([ubq, ubq])  # 2 200 A across: the 21-bit form does not apply, the 32-bit one does
    wide["x"][len(ubq):] += 2200.0
    wide["serial"][len(ubq):] += 5000
    wide["chain"][len(ubq):] = b"B"
    a = A.Packed(wide).arrays()
    assert a["xyz_milli21"] is None and a["x_milli"] is not None and a["attr16"] is not None
    off = ubq.copy()
    off["x"][17] += 1e-7  # one coordinate off the lattice: the form is withdrawn for the whole structure
    a = A.Packed(off).arrays()
    assert a["x_milli"] is None and a["y_milli"] is None and a["z_milli"] is None and a["xyz_milli21"] is None
    neg0 = ubq.copy()
    neg0["z"][3] = -0.0  # "-0.000" parses to -0.0, which 0 / 1000.0 does not give back
    assert A.Packed(neg0).arrays()["z_milli"] is None


CIF_COLS = ["group_PDB", "id", "type_symbol", "label_atom_id", "label_alt_id", "label_comp_id", "label_asym_id", "label_entity_id",
            "label_seq_id", "pdbx_PDB_ins_code", "Cartn_x", "Cartn_y", "Cartn_z", "occupancy", "B_iso_or_equiv", "pdbx_formal_charge",
            "auth_seq_id", "auth_comp_id", "auth_asym_id", "auth_atom_id", "pdbx_PDB_model_num"]


def label_seq_ids(atoms):
    """label_seq_id the way the archive numbers it: 1, 2, 3 ... along each chain, whatever the author numbers are."""
    out, seq, last = np.zeros(len(atoms), dtype=np.int64), 0, None
    for k, a in enumerate(atoms):
        key = (a["model"], a["chain"], a["resi"], a["icode"])
        if last is None or key[:2] != last[:2]:
            seq = 1
        elif key != last:
            seq += 1
        out[k], last = seq, key
    return out


def write_cif(atoms, path, model=1):
    """A PDBx/mmCIF file the way the PDB writes them: label_* and auth_* identifiers differ (waters get their own label
    chain and no label_seq_id), a quoted value and a text field precede the _atom_site loop."""
    def q(b):
        s = b.decode().strip()
        return "." if not s else (f'"{s}"' if ("'" in s or " " in s) else s)

    out = ["data_TEST", "#", "_entry.id TEST", "_struct.title 'a title with loop_ and _atom_site.fake inside quotes'", "#",
           "loop_", "_citation.id", "_citation.title", "primary", ";A multi-line", "text field with loop_ at its start", ";", "#",
           "loop_"] + [f"_atom_site.{c}" for c in CIF_COLS]
    seqs = label_seq_ids(atoms)
    for k, a in enumerate(atoms):
        hoh = a["resn"] == b"HOH"
        label_asym = "W" if hoh else {"A": "A"}.get(a["chain"].decode(), "L" + a["chain"].decode())
        out.append(" ".join([
            "HETATM" if a["hetero"] else "ATOM", str(int(a["serial"])), q(a["element"]).title() if a["element"] else "?", q(a["name"]),
            q(a["altloc"]), q(a["resn"]), label_asym, "1", "." if hoh else str(int(seqs[k])), q(a["icode"]).replace(".", "?"),
            f"{a['x']:.3f}", f"{a['y']:.3f}", f"{a['z']:.3f}", f"{a['occupancy']:.2f}", "10.00", "?", str(int(a["resi"])), q(a["resn"]),
            q(a["chain"]), q(a["name"]), str(model)]))
    out.append("#")
    out += ["loop_", "_pdbx_after.id", "_pdbx_after.value", "1 x", "2 y", "#"]
    path.write_text("\n".join(out) + "\n")


@pytest.mark.parametrize("which", ["1ubq", "6bft", "altloc", "insertion"])
def test_mmcif_reader_gives_the_pdb_readers_records(which, ubq, bft, tmp_path):
    """arph_parse_cif (pdbtbx reads mmCIF by extension, utils.rs:51-57): the _atom_site loop of a file
    written PDB-style parses to the records the fixed-column reader gives for the same atoms."""
    from tests import fixtures as F

    atoms = {"1ubq": ubq, "6bft": bft}.get(which)
    if atoms is None:
        atoms = F.all_cases(ubq.view(O.ATOM_DTYPE))[which].view(A.ATOM_DTYPE)
    p = tmp_path / "s.cif"
    write_cif(atoms, p)
    got = A.parse_cif(p, auth_ids=True)  # the author identifiers are what the PDB-format file carries
    assert len(got) == len(atoms)
    for f in ("x", "y", "z", "occupancy", "serial", "resi", "chain", "resn", "name", "element", "altloc", "icode", "hetero"):
        assert np.array_equal(got[f], atoms[f]), f
    assert (got["model"] == 1).all()
    assert np.array_equal(A.load_atoms(p), A.parse_cif(p))  # dispatch on the extension; label identifiers by default
    q = tmp_path / "s.pdb1"
    q.write_bytes(open(os.path.join(os.path.dirname(__file__), "data", "1ubq.pdb"), "rb").read())
    assert len(A.load_atoms(q)) == 660  # anything else is read as PDB
    # same packed structure -> same contacts, whatever the format
    a, b = A.Packed(atoms).arrays(), A.Packed(got).arrays()
    for k in ("x", "attr", "chain_key", "res_ordinal", "res_id", "ring_center", "ring_normal"):
        assert np.array_equal(a[k], b[k]), k


def test_mmcif_label_identifiers_are_the_default(bft, tmp_path):
    """pdbtbx builds chain ids and residue numbers from label_asym_id / label_seq_id (as far as its parser is
    known; unpinned), so an entry whose author and label ids differ -- every 6bft chain but A here, all residue numbers,
    the waters' own label chain -- gives other chain / resi columns, another `groups` selection and another chain-rank
    order than its PDB-format file.  The author ids stay available behind an explicit flag."""
    p = tmp_path / "s.cif"
    write_cif(bft, p)
    lab, auth = A.parse_cif(p), A.parse_cif(p, auth_ids=True)
    assert np.array_equal(auth["chain"], bft["chain"]) and np.array_equal(auth["resi"], bft["resi"])
    hoh = bft["resn"] == b"HOH"
    want_chain = np.array([b"W" if h else (b"A" if c == b"A" else b"L" + c) for c, h in zip(bft["chain"], hoh)], dtype="S8")
    assert np.array_equal(lab["chain"], want_chain)
    assert np.array_equal(lab["resi"], np.where(hoh, bft["resi"], label_seq_ids(bft)))  # label_seq_id "." -> auth_seq_id
    assert (lab["resi"] != bft["resi"]).any()
    for f in ("x", "y", "z", "serial", "resn", "name", "element", "altloc", "icode"):
        assert np.array_equal(lab[f], auth[f]), f
    # group selection works on the label chain ids: "LB" exists, "B" does not
    import ctypes as C
    lig, rec = C.create_string_buffer(256), C.create_string_buffer(256)
    chains = ",".join(sorted({c.decode() for c in lab["chain"]}))
    assert _lib.lib().arph_parse_groups(chains.encode(), b"LB,LC/A", lig, rec, 256) == 0
    assert lig.value == b"LB,LC" and rec.value == b"A"
    sel = A.Packed(lab, "LB,LC/A").arrays()
    lig_bit, rec_bit = 0x10, 0x20
    assert int(((sel["attr"] & lig_bit) != 0).sum()) == int(np.isin(lab["chain"], [b"LB", b"LC"]).sum())
    assert int(((sel["attr"] & rec_bit) != 0).sum()) == int((lab["chain"] == b"A").sum())
    # "B" names no chain under the label ids: nothing is selected as ligand, whereas the author ids select chain B
    assert int(((A.Packed(lab, "B/A").arrays()["attr"] & lig_bit) != 0).sum()) == 0
    assert int(((A.Packed(auth, "B/A").arrays()["attr"] & lig_bit) != 0).sum()) == int((bft["chain"] == b"B").sum())


def test_mmcif_tokenizer_edge_cases(tmp_path):
    p = tmp_path / "t.cif"
    p.write_text("""data_X
loop_
_atom_site.group_PDB
_atom_site.id
_atom_site.label_atom_id
_atom_site.label_comp_id
_atom_site.label_asym_id
_atom_site.label_seq_id
_atom_site.Cartn_x
_atom_site.Cartn_y
_atom_site.Cartn_z
ATOM 1 "O5'" ALA A 1 1.0 2.0 3.0 # trailing comment
ATOM 2 'C"x' ALA
A 1 -1.5e0 +2.25 3
HETATM 3 O HOH B . 0 0 0
_other.tag value
""")
    got = A.parse_cif(p)
    assert [n.decode() for n in got["name"]] == ["O5'", 'C"x', "O"]
    assert got["x"].tolist() == [1.0, -1.5, 0.0] and got["y"].tolist() == [2.0, 2.25, 0.0]
    assert got["hetero"].tolist() == [0, 0, 1] and got["resi"].tolist() == [1, 1, 0] and got["chain"].tolist() == [b"A", b"A", b"B"]
    assert got["element"].tolist() == [b"O", b"C", b"O"]  # no type_symbol: from the atom name
    bad = tmp_path / "bad.cif"
    bad.write_text("data_X\nloop_\n_atom_site.id\n_atom_site.Cartn_x\n1 2.0\n")
    with pytest.raises(A.ArpeggiaError):
        A.parse_cif(bad)


def test_mmcif_reader_pieces_on_all_threads(bft, tmp_path, monkeypatch):
    """The one-row-per-line path of arph_parse_cif cuts the loop into pieces (ARPH_CIF_PIECE_BYTES makes them small): the
    same records whatever the cut; a wrapped row or a text-field line anywhere sends the whole loop to the serial tokenizer;
    an unreadable coordinate reports its row number across pieces."""
    p = tmp_path / "s.cif"
    write_cif(bft, p)
    whole = A.parse_cif(p)
    assert len(whole) == len(bft)
    for piece in ("97", "1000", "65536"):
        monkeypatch.setenv("ARPH_CIF_PIECE_BYTES", piece)
        assert np.array_equal(A.parse_cif(p), whole), piece
    lines = p.read_text().split("\n")
    first = next(i for i, l in enumerate(lines) if l.startswith(("ATOM", "HETATM")))
    # row 6000 wrapped over two lines, row 7000 followed by a comment line and a blank line: still the same records
    k = first + 6000
    toks = lines[k].split(" ")
    wrapped = lines[:k] + [" ".join(toks[:5]), " ".join(toks[5:])] + lines[k + 1:first + 7000] + ["# a comment", ""] + lines[first + 7000:]
    q = tmp_path / "w.cif"
    q.write_text("\n".join(wrapped))
    for piece in ("500", "1048576"):
        monkeypatch.setenv("ARPH_CIF_PIECE_BYTES", piece)
        assert np.array_equal(A.parse_cif(q), whole), piece
    # comment and blank lines between rows, every row on its own line: the pieces stay on the parallel path, the gaps
    # they leave in the record array are closed up
    gaps = lines[:first + 3000] + ["# one", "", "# two"] + lines[first + 3000:first + 7000] + ["# a comment", ""] + lines[first + 7000:]
    g = tmp_path / "g.cif"
    g.write_text("\n".join(gaps))
    for piece in ("300", "5000", "1048576"):
        monkeypatch.setenv("ARPH_CIF_PIECE_BYTES", piece)
        assert np.array_equal(A.parse_cif(g), whole), piece
    # not a regular file (a FIFO): read in pieces instead of mapped
    import os
    import threading

    fifo = tmp_path / "pipe.cif"
    os.mkfifo(fifo)
    feeder = threading.Thread(target=lambda: fifo.write_bytes(p.read_bytes()))
    feeder.start()
    try:
        assert np.array_equal(A.parse_cif(fifo), whole)
    finally:
        feeder.join()
    # unreadable coordinate in row 5001
    toks = lines[first + 5000].split(" ")
    toks[10] = "abc"
    bad = tmp_path / "b.cif"
    bad.write_text("\n".join(lines[:first + 5000] + [" ".join(toks)] + lines[first + 5001:]))
    monkeypatch.setenv("ARPH_CIF_PIECE_BYTES", "700")
    with pytest.raises(A.ArpeggiaError, match="row 5001"):
        A.parse_cif(bad)


def _packed_fingerprint(atoms, flags=0, izo=False):
    """Everything the packer and the side-table builder produce for these records, as comparable arrays."""
    import ctypes as C

    from arpeggia_b200 import _lib

    p = A.Packed(atoms, "/", ignore_zero_occupancy=izo, flags=flags)
    out = {k: v for k, v in p.arrays().items() if v is not None}
    tp = C.POINTER(_lib.FrameTables)()
    _lib.check(_lib.lib().arph_packed_frame_tables(p._h, C.byref(tp)))
    t = tp.contents
    E = t.n_atoms + t.n_rings
    out["sides"] = np.frombuffer((C.c_char * (32 * t.n_sides)).from_address(C.addressof(t.sides.contents)), dtype=np.uint32).copy() if t.n_sides else np.zeros(0, np.uint32)
    out["side_sc"] = np.ctypeslib.as_array(t.side_sc, shape=(6 * t.n_sides,)).copy() if t.n_sides else np.zeros(0)
    for k in ("ent_side", "ent_atomn", "ent_atomi", "ent_rank", "ent_tie"):
        out[k] = np.ctypeslib.as_array(getattr(t, k), shape=(E,)).copy()
    out["bounds"] = np.array([t.n_chain_ranks, t.n_ent_ranks, t.max_tie])
    out["warnings"] = np.array(list(p.warnings), dtype=object)
    # the string dictionary, through an (empty) table
    out["strings"] = np.array(p.build_table(np.zeros(0, dtype=A.ROW_DTYPE))["strings"], dtype=object)
    return out, p


@pytest.mark.parametrize("blocks", [2, 7, 64])
def test_packer_blocks_and_table_paths_give_one_result(ubq, bft, blocks, monkeypatch):
    """The packer groups the records block by block on the host's threads and joins the blocks; the side tables rank every
    chain where it lies, or after a counting sort by chain.  Whatever the cut and the path: the same arrays, bit for bit --
    also for records whose chains, residues and conformers are scattered over the file."""
    from tests import fixtures as F

    cases = {"1ubq": ubq, "6bft": bft}
    cases.update({k: v.view(A.ATOM_DTYPE) for k, v in F.all_cases(ubq.view(O.ATOM_DTYPE)).items()})
    rng = np.random.default_rng(11)
    scattered = bft[rng.permutation(len(bft))]  # every chain, residue and conformer continues in later blocks
    cases["scattered"] = scattered
    by_residue = bft[np.argsort(rng.permutation(bft["resi"].max() + 1)[bft["resi"]], kind="stable")]  # whole residues moved about
    cases["residues_moved"] = by_residue
    for name, atoms in cases.items():
        izo = name == "zero_occupancy"
        monkeypatch.setenv("ARPH_PACK_BLOCKS", "1")
        monkeypatch.delenv("ARPH_TABLES_BUCKET", raising=False)
        want, _ = _packed_fingerprint(atoms, izo=izo)
        monkeypatch.setenv("ARPH_PACK_BLOCKS", str(blocks))
        got, _ = _packed_fingerprint(atoms, izo=izo)
        monkeypatch.setenv("ARPH_TABLES_BUCKET", "1")
        got2, _ = _packed_fingerprint(atoms, izo=izo)
        assert set(got) == set(want) == set(got2)
        for k in want:
            for g in (got, got2):
                w = want[k]
                same = (g[k] == w) if np.isscalar(w) else (np.array_equal(g[k].view(np.uint64), w.view(np.uint64)) if getattr(w, "dtype", None) == np.float64 else np.array_equal(g[k], w))
                assert same, (name, k, blocks)


def test_mmcif_reals_are_correctly_rounded(tmp_path):
    """The reader's fast path for reals ([-]digits[.digits], <= 15 digits: one division of two exact doubles) and its
    from_chars path give the correctly rounded double for every spelling: checked against Python's float()."""
    rng = np.random.default_rng(3)
    texts = ["0", "-0", "-0.000", "+1.5", "5.", ".5", "-.25", "000012.3400", "1e3", "-2.5E-3", "123456789012345", "0.123456789012345",
             "1234567.89012345", "12345678901234567890", "0.1234567890123456789", "9007199254740993", "1.7976931348623157e308", "4.9e-324"]
    for _ in range(3000):
        digits = int(rng.integers(1, 19))
        frac = int(rng.integers(0, digits + 1))
        body = "".join(str(int(d)) for d in rng.integers(0, 10, size=digits))
        t = body[: digits - frac] + ("." + body[digits - frac:] if frac or rng.random() < 0.2 else "")
        if t in ("", "."):
            t = "0"
        texts.append(("-" if rng.random() < 0.5 else "") + t)
    head = "data_X\nloop_\n_atom_site.group_PDB\n_atom_site.id\n_atom_site.type_symbol\n_atom_site.label_atom_id\n_atom_site.label_comp_id\n" \
           "_atom_site.label_asym_id\n_atom_site.label_seq_id\n_atom_site.Cartn_x\n_atom_site.Cartn_y\n_atom_site.Cartn_z\n_atom_site.occupancy\n"
    rows = [f"ATOM {k + 1} C CA ALA A {k + 1} {t} 1.0 -2.0 1.00" for k, t in enumerate(texts)]
    p = tmp_path / "reals.cif"
    p.write_text(head + "\n".join(rows) + "\n")
    got = A.parse_cif(p)
    assert len(got) == len(texts)
    want = np.array([float(t) for t in texts])
    assert np.array_equal(got["x"].view(np.uint64), want.view(np.uint64))
    assert (got["y"] == 1.0).all() and (got["z"] == -2.0).all() and (got["occupancy"] == 1.0).all()


def test_golden_csv_of_1ubq(ubq):
    """tests/golden/1ubq_contacts.csv = the oracle's table of 1ubq (pinned by the reference's golden values) through the CLI's
    CSV writer, in the layout of polars' CsvWriter: bare header, quotes only where needed, "2.0" not "2", nulls empty."""
    from arpeggia_b200.frame import csv_text, to_arrow

    golden = (ROOT / "tests" / "golden" / "1ubq_contacts.csv").read_text()
    assert csv_text(to_arrow(O.contacts(ubq.view(O.ATOM_DTYPE)))) == golden
    lines = golden.splitlines()
    assert len(lines) == 533 and lines[0].split(",") == A.COLUMNS and '"' not in golden
    import pyarrow as pa

    t = pa.table({"s": pa.array(["x", "y,z", None, 'q"t'], type=pa.large_string()), "f": pa.array([1.5, None, 2.0, 1e-7], type=pa.float32()),
                  "i": pa.array([1, 2, None, 4], type=pa.int32())})
    assert csv_text(t) == 's,f,i\nx,1.5,1\n"y,z",,2\n,2.0,\n"q""t",1e-7,4\n'
