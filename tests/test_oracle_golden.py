"""Pins the CPU oracle against every golden value the reference's own tests hold for the contacts path
(SURVEY.md 8c, G1-G8).  CPU only."""
import collections

import numpy as np
import pytest

from oracle import oracle as O

COLUMNS = [
    "model", "interaction", "distance", "from_chain", "from_resn", "from_resi", "from_insertion", "from_altloc",
    "from_atomn", "from_atomi", "to_chain", "to_resn", "to_resi", "to_insertion", "to_altloc", "to_atomn", "to_atomi",
    "sc_centroid_dist", "sc_dihedral", "sc_centroid_angle",
]


@pytest.fixture(scope="module")
def ubq(ubq_path):
    return O.parse_pdb(ubq_path)


@pytest.fixture(scope="module")
def bft(bft_path):
    return O.parse_pdb(bft_path)


def test_g1_1ubq_532_rows_20_columns(ubq):
    # python/tests/test_arpeggia.py:28-71
    t = O.contacts(ubq, groups="/", vdw_comp=0.1, dist_cutoff=6.5)
    assert t["n_rows"] == 532
    assert O.COLUMNS == COLUMNS and all(c in t for c in COLUMNS)
    assert t["distance"].dtype == np.float32 and (t["distance"] >= 0).all()
    # label breakdown found by the survey's independent NumPy probe
    assert collections.Counter(t["interaction"]) == {
        "VanDerWaalsContact": 199, "HydrophobicContact": 167, "PolarContact": 107, "WeakPolarContact": 55, "IonicBond": 4}
    assert t["n_atoms"] == 660 and t["pairs_in_cutoff"] == 29028 and t["pairs_classified"] == 9128


def test_g2_zero_occupancy_flag_is_a_noop_on_1ubq(ubq):
    # python/tests/test_arpeggia.py:85-112 and src/utils.rs:230-247
    assert (ubq["occupancy"] != 0.0).all()
    a = O.contacts(ubq, ignore_zero_occupancy=False)
    b = O.contacts(ubq, ignore_zero_occupancy=True)
    assert a["n_rows"] == b["n_rows"] == 532


def _has_row(t, inter, fc, fresi, fatom, tc, tresi, tatom):
    m = ((t["interaction"] == inter) & (t["from_chain"] == fc) & (t["from_resi"] == fresi) & (t["from_atomn"] == fatom)
         & (t["to_chain"] == tc) & (t["to_resi"] == tresi) & (t["to_atomn"] == tatom))
    return bool(m.any())


def test_g3_g4_6bft_cation_pi(bft):
    t = O.contacts(bft)
    # src/contacts/aromatic.rs:72-99: ring of TYR A102 with ARG G82 NE -> CationPi
    assert _has_row(t, "CationPi", "A", 102, "Ring", "G", 82, "NE")
    # src/contacts/aromatic.rs:101-128: ring of TRP A108 with LYS G84 NZ -> None (angle)
    assert not _has_row(t, "CationPi", "A", 108, "Ring", "G", 84, "NZ")
    # survey probe numbers (unpinned by the reference, smoke check only)
    assert t["n_rows"] == 7236 and t["pairs_in_cutoff"] == 396474 and t["pairs_classified"] == 124047


def test_g4_angle_is_the_reason(bft):
    # TRP A108 ring / LYS G84 NZ: within 4.5 A? whatever the distance, theta must exceed 30 or dist exceed 4.5
    n, c, nrm = O.residue_plane(bft, "A", 108, "ring")
    assert n == 9
    nz = bft[(bft["chain"] == b"G") & (bft["resi"] == 84) & (bft["name"] == b"NZ")][0]
    p = np.array([nz["x"], nz["y"], nz["z"]])
    theta, dist = O.plane_point_angle(c, nrm, p), O.plane_point_dist(c, p)
    assert theta > 30.0 or dist > 4.5
    # and the good one
    n, c, nrm = O.residue_plane(bft, "A", 102, "ring")
    ne = bft[(bft["chain"] == b"G") & (bft["resi"] == 82) & (bft["name"] == b"NE")][0]
    p = np.array([ne["x"], ne["y"], ne["z"]])
    assert O.plane_point_angle(c, nrm, p) <= 30.0 and O.plane_point_dist(c, p) <= 4.5


def test_g5_residue_ext_1ubq(ubq):
    # src/contacts/residues.rs:334-395
    n, _, _ = O.residue_plane(ubq, "A", 1, "ring")
    assert n == 0  # MET1 has no ring atoms
    n, c, nrm = O.residue_plane(ubq, "A", 1, "sc")
    assert n == 3 and c is not None
    n, c, nrm = O.residue_plane(ubq, "A", 4, "ring")
    assert n == 6
    np.testing.assert_allclose(c, [24.96883333, 34.687, 6.16233333], rtol=1e-6)
    ref = np.array([0.53253994, -0.82736044, -0.17853828])
    if np.dot(nrm, ref) < 0:  # SVD sign is arbitrary; every consumer folds angles to <= 90 degrees
        nrm = -nrm
    np.testing.assert_allclose(nrm, ref, rtol=1e-6)
    ns, cs, nrms = O.residue_plane(ubq, "A", 4, "sc")  # PHE: sc plane atoms == ring atoms
    assert ns == 6 and np.array_equal(cs, c) and np.array_equal(np.abs(nrms), np.abs(nrm))
    ring = ubq[(ubq["resi"] == 4) & np.isin(ubq["name"], [b"CG", b"CD1", b"CD2", b"CE1", b"CE2", b"CZ"])]
    pts = np.stack([ring["x"], ring["y"], ring["z"]], 1)
    assert np.abs((pts - c) @ nrm).mean() < 0.02


def test_g6_plane_dist_angles():
    # src/contacts/residues.rs:306-332
    o, z = np.zeros(3), np.array([0.0, 0.0, 1.0])
    pt = np.array([0.0, 1.0, 1.0])
    assert abs(O.plane_point_dist(o, pt) - np.sqrt(2.0)) < 1e-6
    assert abs(O.plane_point_angle(o, z, pt) - 45.0) < 1e-6
    assert abs(O.plane_point_angle(pt, -z, o) - 45.0) < 1e-6
    assert O.plane_dihedral(z, -z) < 1e-6
    x = np.array([1.0, 0.0, 0.0])
    assert abs(O.plane_point_angle(pt, x, o) - 90.0) < 1e-6
    assert abs(O.plane_dihedral(z, x) - 90.0) < 1e-6


def test_g7_parse_groups():
    # src/utils.rs:174-228
    chains = {"A", "B", "C", "D"}
    assert O.parse_groups(chains, "A,B/C,D") == ({"A", "B"}, {"C", "D"})
    assert O.parse_groups(chains, "A/C,D") == ({"A"}, {"C", "D"})
    assert O.parse_groups(chains, "/C,D") == ({"A", "B"}, {"C", "D"})
    assert O.parse_groups(chains, "C/") == ({"C"}, {"A", "B", "D"})
    assert O.parse_groups(chains, "/") == (chains, chains)
    with pytest.raises(O.OracleError, match="Invalid chain groups format! Use '/' for all-to-all comparisons."):
        O.parse_groups(chains, "A,B")
    with pytest.raises(O.OracleError, match="Empty chain groups!"):
        O.parse_groups(chains, "A,B,C,D/")


def test_g8_1ubq_chain_layout(ubq):
    # src/contacts/chains.rs:21-40: chain A = 76 amino acids followed by 58 waters in the same chain
    assert set(ubq["chain"]) == {b"A"}
    resis = []
    for r in ubq:
        key = (int(r["resi"]), r["resn"])
        if not resis or resis[-1] != key:
            resis.append(key)
    names = [k[1] for k in resis]
    assert len(names) == 134 and names[76:] == [b"HOH"] * 58 and b"HOH" not in names[:76]
    assert names[0] == b"MET" and names[75] == b"GLY"


def test_group_selection_on_6bft(bft):
    all_rows = O.contacts(bft, groups="/")
    ab = O.contacts(bft, groups="A,B/G")
    assert 0 < ab["n_rows"] < all_rows["n_rows"]
    # every row is ligand -> receptor, except ring-atom rows where the ring is always `from` (complex.rs:334-343)
    ring = ab["from_atomn"] == "Ring"
    assert set(ab["from_chain"][~ring]) <= {"A", "B"} and set(ab["to_chain"][~ring]) == {"G"}
    # "H,L/" means H,L against everything else
    hl = O.contacts(bft, groups="H,L/")
    ring = hl["from_atomn"] == "Ring"
    assert set(hl["to_chain"][~ring]) <= {"A", "B", "C", "G"}
    with pytest.raises(O.OracleError, match="Empty chain groups!"):
        O.contacts(bft, groups="A,B,C,G,H,L/")


def test_rows_are_sorted_by_the_reference_keys(bft):
    t = O.contacts(bft)
    inter = t["interaction"]
    keys = list(zip(t["model"].tolist(), t["from_chain"], t["to_chain"], t["from_resi"].tolist(), t["from_altloc"],
                    t["from_atomi"].tolist(), t["to_resi"].tolist(), t["to_altloc"], t["to_atomi"].tolist(), inter))
    assert keys == sorted(keys)
    # sc_* are null when one side has no side-chain plane: ALA / GLY / HOH, or a truncated residue with
    # fewer than 3 plane atoms (HIS A230 in 6bft has no side chain)
    noplane = np.isin(t["from_resn"], ["ALA", "GLY", "HOH"]) | np.isin(t["to_resn"], ["ALA", "GLY", "HOH"])
    assert not t["sc_valid"][noplane].any()
    extra = (~t["sc_valid"]) & ~noplane
    assert set(zip(t["from_chain"][extra], t["from_resi"][extra].tolist())) == {("A", 230)}
    assert np.isnan(t["sc_dihedral"][~t["sc_valid"]]).all() and np.isfinite(t["sc_dihedral"][t["sc_valid"]]).all()
