"""SURVEY 8(f) rank 3 on the GPU: arp_stage_neighbor_sum (the SAP radius query, src/sap.rs:181-215) against the oracle.
The neighbour SET is exact (integer weights make every order of summation give the same f32: bit-exact check); with real
weights the device rounds an f64 sum once, the oracle adds f32 terms in atom order: equal within n * 2^-24 * sum|w|."""
import numpy as np
import pytest

import arpeggia_b200 as A
from arpeggia_b200 import sap, synth
from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    return A.Context(0)


def run_both(ctx, atoms, w, sel, radius):
    packed = A.Packed(atoms, "/", flags=2)
    a = packed.arrays()
    assert np.array_equal(packed.atom_serials(), atoms["serial"])
    ctx.upload(packed)
    got = ctx.neighbor_sum(w, sel, radius)
    want = O.neighbor_sum(a["x"], a["y"], a["z"], w, sel, radius)
    return got, want


@pytest.mark.parametrize("which", ["1ubq", "6bft"])
@pytest.mark.parametrize("radius", [5.0, 3.3, 9.75])
def test_neighbor_set_is_exact(ctx, ubq_path, bft_path, which, radius):
    atoms = sap.prepare_atoms(A.parse_pdb(ubq_path if which == "1ubq" else bft_path))
    sel = sap.sidechain_mask(atoms)
    rng = np.random.default_rng(11)
    w = rng.integers(-8, 9, size=len(atoms)).astype(np.float32)  # exactly representable sums: order cannot matter
    got, want = run_both(ctx, atoms, w, sel, radius)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert (got[~sel] == 0).all() and np.abs(got[sel]).max() > 0
    ones, _ = run_both(ctx, atoms, np.ones(len(atoms), dtype=np.float32), sel, radius)  # = number of side-chain neighbours
    assert ones[sel].min() >= 1  # the atom itself (sap.rs:193-194)


def test_sap_scores_within_f32_summation_tolerance(ctx, bft_path):
    atoms = sap.prepare_atoms(A.parse_pdb(bft_path))
    rng = np.random.default_rng(5)
    sasa = rng.uniform(0.0, 40.0, size=len(atoms)).astype(np.float32)  # stand-in for the SASA step (outside this path)
    w, sel = sap.sap_weights(atoms, sasa), sap.sidechain_mask(atoms)
    got, want = run_both(ctx, atoms, w, sel, 5.0)
    bound, _ = run_both(ctx, atoms, np.abs(w), sel, 5.0)
    counts, _ = run_both(ctx, atoms, np.ones(len(atoms), dtype=np.float32), sel, 5.0)
    assert np.all(np.abs(got - want) <= (counts + 1) * 2.0 ** -24 * bound + 1e-12)
    res = sap.per_atom_sap(atoms, sasa, 5.0, ctx=ctx)
    order = np.argsort(atoms["serial"][sel], kind="stable")
    assert np.array_equal(res["atomi"], atoms["serial"][sel][order]) and np.array_equal(res["sap_score"], got[sel][order])
    assert (res["sap_score"] > 0).any() and (res["sap_score"] < 0).any()  # src/sap.rs:428-449


def test_inclusive_tie_and_errors(ctx, ubq_path):
    ubq = sap.prepare_atoms(A.parse_pdb(ubq_path))
    tie = ubq[:4].copy()  # four real records moved onto a 3-4-5 triangle: |p1 - p0| = 5 exactly
    for k, (x, y, z) in enumerate([(0.0, 0.0, 0.0), (3.0, 4.0, 0.0), (3.0, 4.0, 0.001), (100.0, 0.0, 0.0)]):
        tie["x"][k], tie["y"][k], tie["z"][k] = x, y, z
    atoms = np.concatenate([tie, ubq[4:]])
    atoms["x"][4:] += 500.0  # the rest of the protein far away (the packer wants a ring residue)
    sel = np.zeros(len(atoms), dtype=bool)
    sel[:4] = True
    w = np.zeros(len(atoms), dtype=np.float32)
    w[:4] = [1, 2, 4, 8]
    got, want = run_both(ctx, atoms, w, sel, 5.0)
    assert got[:4].tolist() == want[:4].tolist() == [3.0, 7.0, 6.0, 8.0]
    with pytest.raises(A.ArpeggiaError):
        ctx.neighbor_sum(w, sel, -1.0)


def test_batch_job_sums_stay_inside_each_structure(ctx, ubq_path, bft_path):
    """A batch job (arp_stage_upload of several structures): every structure has its own cell list, the sums are those of
    the structures run one by one -- also for structures that overlap in space (two copies of 1ubq at the same place)."""
    ubq, bft = sap.prepare_atoms(A.parse_pdb(ubq_path)), sap.prepare_atoms(A.parse_pdb(bft_path))
    far = ubq.copy()
    far["x"] += 1000.0
    parts = [ubq, bft, ubq, far]
    rng = np.random.default_rng(17)
    ws = [rng.integers(-8, 9, size=len(p)).astype(np.float32) for p in parts]
    sels = [sap.sidechain_mask(p) for p in parts]
    sels[2] = np.ones(len(ubq), dtype=bool)
    single = [run_both(ctx, p, w, s, 5.0) for p, w, s in zip(parts, ws, sels)]
    for got, want in single:
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    ctx.upload([A.Packed(p, "/", flags=2) for p in parts])
    got = ctx.neighbor_sum(np.concatenate(ws), np.concatenate(sels), 5.0)
    assert np.array_equal(got.view(np.uint32), np.concatenate([g for g, _ in single]).view(np.uint32))


def test_neighbor_sum_1m_atoms(ctx, bft_path):
    """Config-3 size: 6bft tiled 5x5x5, integer weights, bit-exact against the oracle; device time reported."""
    atoms = synth.assembly_1m(sap.prepare_atoms(A.parse_pdb(bft_path)))
    sel = sap.sidechain_mask(atoms)
    w = np.random.default_rng(3).integers(-4, 5, size=len(atoms)).astype(np.float32)
    got, want = run_both(ctx, atoms, w, sel, 5.0)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    st = ctx.neighbor_sum_stats()
    assert st["ms_sum"] > 0 and st["n_cells"] > 0
    print("neighbor_sum 1M atoms:", st)


def test_selection_edge_cases(ctx, ubq_path):
    """The cell list holds the selected atoms only: nothing selected -> all zeros (also right after a run that selected
    many: the output buffer is reused), everything selected -> the oracle's sums over all atoms, one atom -> its own weight."""
    atoms = sap.prepare_atoms(A.parse_pdb(ubq_path))
    n = len(atoms)
    w = np.random.default_rng(2).integers(1, 6, size=n).astype(np.float32)
    every = np.ones(n, dtype=bool)
    got, want = run_both(ctx, atoms, w, every, 5.0)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)) and (got >= w).all()
    none = np.zeros(n, dtype=bool)
    got, want = run_both(ctx, atoms, w, none, 5.0)
    assert not got.any() and not want.any()
    one = none.copy()
    one[n // 2] = True
    got, want = run_both(ctx, atoms, w, one, 5.0)
    assert got[n // 2] == w[n // 2] == want[n // 2] and np.count_nonzero(got) == 1
