"""The synthetic edge-case fixtures really exercise the branches no reference fixture reaches (oracle side),
and the product's packer sees the same hierarchy.  CPU only."""
import collections

import numpy as np
import pytest

import arpeggia_b200 as A
from oracle import oracle as O
from tests import fixtures as F


@pytest.fixture(scope="module")
def cases(ubq_path):
    return F.all_cases(O.parse_pdb(ubq_path))


def test_hydrogen_branches_fire(cases):
    t = O.contacts(cases["protonated"])
    c = collections.Counter(t["interaction"])
    assert c["HydrogenBond"] > 0 and c["WeakHydrogenBond"] > 0 and c["PolarContact"] > 0 and c["WeakPolarContact"] > 0
    s = O.contacts(cases["salt_bridges"])
    c = collections.Counter(s["interaction"][s["from_chain"] == "S"])
    assert c == {"SaltBridge": 1, "IonicBond": 2, "VanDerWaalsContact": 3, "IonicRepulsion": 1}


def test_ties_are_inclusive_and_clash_suppresses(cases):
    t = O.contacts(cases["ties"])
    w = (t["from_chain"] == "W") | (t["to_chain"] == "W")
    # the 0.9 A pair gives a clash and nothing else (complex.rs:233-235); two other tie waters are 1.6 / 2.5 A apart
    assert collections.Counter(t["interaction"][w]) == {"StericClash": 1, "VanDerWaalsContact": 2}
    base = O.contacts(cases["insertion"])
    # exactly-at-cutoff pairs count as within the cutoff (rstar locate_within_distance is inclusive)
    assert t["pairs_in_cutoff"] > base["pairs_in_cutoff"]


def test_altloc_rows_carry_altloc_and_duplicate_shared_atoms(cases):
    t = O.contacts(cases["altloc"])
    assert set(t["from_altloc"]) | set(t["to_altloc"]) == {"", "A", "B"}
    p = A.Packed(cases["altloc"].view(A.ATOM_DTYPE)).arrays()
    # blank-altloc backbone atoms of an altloc residue appear once per conformer (pdbtbx reader post-pass)
    assert p["n_atoms"] == t["n_atoms"] > len(cases["altloc"])


def test_zero_occupancy_and_models(cases):
    a = O.contacts(cases["zero_occupancy"], ignore_zero_occupancy=False)
    b = O.contacts(cases["zero_occupancy"], ignore_zero_occupancy=True)
    assert b["n_rows"] < a["n_rows"] == 532
    m = O.contacts(cases["two_models"])
    assert set(m["model"].tolist()) == {1, 2} and m["n_rows"] > 532
    # the multi-model key overwrite (complex.rs:447-449) only moves the sc_* columns here
    q = O.contacts(cases["two_models"], flags=O.FLAG_NO_MULTIMODEL_QUIRK)
    assert q["n_rows"] == m["n_rows"] and not np.array_equal(q["sc_centroid_angle"], m["sc_centroid_angle"], equal_nan=True)


@pytest.mark.parametrize("case", ["protonated", "altloc", "insertion", "two_models", "ties"])
def test_packer_counts_match_oracle(cases, case):
    p = A.Packed(cases[case].view(A.ATOM_DTYPE)).arrays()
    t = O.contacts(cases[case], flags=O.FLAG_NO_TABLE)
    assert p["n_atoms"] == t["n_atoms"]
    assert p["n_h"] == int((cases[case]["element"] == b"H").sum()) or case == "altloc"
