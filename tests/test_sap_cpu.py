"""SURVEY 8(f) rank 3, CPU side: the oracle's restatement of the SAP radius query (src/sap.rs:181-215) against a NumPy
brute force, and the residue tables against the known answers of the reference's own tests (src/sap.rs:362-399)."""
import numpy as np
import pytest

import arpeggia_b200 as A
from arpeggia_b200 import sap
from oracle import oracle as O


@pytest.fixture(scope="module")
def ubq(ubq_path):
    return A.parse_pdb(ubq_path)


def brute(p, w, sel, radius):
    d = p[:, None, :] - p[None, :, :]
    d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
    r2 = float(np.float32(radius) * np.float32(radius))  # sap.rs:184
    m = (d2 <= r2) & sel[None, :] & sel[:, None]
    return np.array([np.cumsum(w[m[i]], dtype=np.float32)[-1] if sel[i] else np.float32(0) for i in range(len(p))], dtype=np.float32)


def test_oracle_neighbor_sum_equals_brute_force(ubq):
    rng = np.random.default_rng(7)
    prep = sap.prepare_atoms(ubq)
    p = np.stack([prep["x"], prep["y"], prep["z"]], axis=1)
    sel = sap.sidechain_mask(prep)
    w = rng.normal(size=len(prep)).astype(np.float32)
    for radius in (5.0, 3.3, 7.25):
        got = O.neighbor_sum(p[:, 0], p[:, 1], p[:, 2], w, sel, radius)
        assert np.array_equal(got.view(np.uint32), brute(p, w, sel, radius).view(np.uint32))
    assert (O.neighbor_sum(p[:, 0], p[:, 1], p[:, 2], w, sel, 5.0)[~sel] == 0).all()


def test_inclusive_radius_and_self_term():
    """locate_within_distance is inclusive and returns the query atom itself (sap.rs:193-194)."""
    p = np.array([[0.0, 0.0, 0.0], [3.0, 4.0, 0.0], [3.0, 4.0, 0.001], [100.0, 0.0, 0.0]])  # |p1 - p0| = 5 exactly
    w = np.array([1.0, 2.0, 4.0, 8.0], dtype=np.float32)
    sel = np.ones(4, dtype=bool)
    got = O.neighbor_sum(p[:, 0], p[:, 1], p[:, 2], w, sel, 5.0)
    assert got.tolist() == [3.0, 7.0, 6.0, 8.0]
    sel[1] = False  # an unselected atom neither receives nor contributes
    assert O.neighbor_sum(p[:, 0], p[:, 1], p[:, 2], w, sel, 5.0).tolist() == [1.0, 0.0, 4.0, 8.0]


def test_residue_tables_known_answers():
    """src/sap.rs:362-399 (test_hydrophobicity_values, test_max_asa_values)."""
    assert abs(float(sap.get_hydrophobicity("GLY"))) < 1e-6
    assert float(sap.get_hydrophobicity("PHE")) > 0.4 and float(sap.get_hydrophobicity("ARG")) < -0.4
    assert sap.get_hydrophobicity("XXX") is None and sap.get_hydrophobicity("phe") == sap.get_hydrophobicity("PHE")
    for aa in ("ALA ARG ASN ASP CYS GLN GLU GLY HIS ILE LEU LYS MET PHE PRO SER THR TRP TYR VAL").split():
        assert sap.get_sc_max_asa(aa) is not None and float(sap.get_sc_max_asa(aa)) > 0
    assert sap.get_sc_max_asa("XXX") is None
    assert float(sap.get_sc_max_asa("TRP")) > float(sap.get_sc_max_asa("ALA"))


def test_prepare_and_weights(ubq):
    prep = sap.prepare_atoms(ubq)
    assert len(prep) == 602 and not np.isin(prep["resn"], [b"HOH"]).any()  # 1ubq: 58 waters out (chains.rs:21-40)
    assert len(sap.prepare_atoms(ubq, "B")) == 0 and len(sap.prepare_atoms(ubq, "A, ")) == 602
    sel = sap.sidechain_mask(prep)
    assert sel.sum() == 602 - 4 * 76 + 0  # 76 residues, N / CA / C / O each (OXT of the C-terminus counts as side chain)  # noqa
    w = sap.sap_weights(prep, np.full(len(prep), 1e9))  # clamped at 1: the weight is the hydrophobicity
    phe = prep["resn"] == b"PHE"
    assert np.all(w[phe] == sap.get_hydrophobicity("PHE")) and np.all(sap.sap_weights(prep, np.zeros(len(prep))) == 0)
