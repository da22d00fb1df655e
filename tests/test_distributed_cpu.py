"""World-size-2 gloo tests of the multi-GPU host logic (sharding, slab ownership, halo exchange).  CPU only."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import arpeggia_b200 as A
from arpeggia_b200 import distributed as D

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, case, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle as O
        from tests import fixtures as F

        if case == "bft":
            atoms = A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb"))
        else:
            atoms = F.protonate_crude(O.parse_pdb(os.path.join(ROOT, "tests", "data", "1ubq.pdb"))).view(A.ATOM_DTYPE)
        arr = A.Packed(atoms).arrays()
        whole = D.soa_from_arrays(arr)
        own = D.owned_part(whole, rank, world)
        width = 6.5 * (1 + 1e-6) + 1e-6
        comb, info = D.exchange_halos(own, width)
        n_own = own["x"].numel()
        # ghosts are exactly the appended part
        ghost = (comb["attr"] & D.GHOST) != 0
        assert not ghost[:n_own].any() and ghost[n_own:].all()
        # every atom within `width` of an owned atom is present locally
        from scipy.spatial import cKDTree
        allp = np.stack([arr["x"], arr["y"], arr["z"]], 1)
        ownp = np.stack([own["x"].numpy(), own["y"].numpy(), own["z"].numpy()], 1)
        need = set()
        for lst in cKDTree(allp).query_ball_point(ownp, width):
            need.update(lst)
        have = set(comb["gid"].tolist())
        assert need <= have, f"rank {rank}: {len(need - have)} halo atoms missing"
        # every ring within 6 A of an owned ring is present
        if arr["n_rings"]:
            rc = arr["ring_center"]
            mine = own["ring_gid"].numpy()
            near = set()
            for lst in cKDTree(rc).query_ball_point(rc[mine], 6.0 + 1e-6):
                near.update(lst)
            assert near <= set(comb["ring_gid"].tolist())
        # internal consistency of the combined structure
        rid = comb["res_id"]
        assert (rid[1:] >= rid[:-1]).all() and int(rid.max()) == D.n_res(comb) - 1
        for k in ("res_cb", "res_sg"):
            v = comb[k]
            ok = v >= 0
            assert (rid[v[ok]] == torch.nonzero(ok).flatten()).all()
        hb = comb["h_begin"]
        assert hb[0] == 0 and (hb[1:] >= hb[:-1]).all() and int(hb[-1]) == comb["hx"].numel()
        # hydrogens travelled with their residues: per-residue H count equals the undivided structure's
        g_counts = np.diff(arr["res_h_begin"].astype(np.int64))
        first_atom = torch.zeros(D.n_res(comb), dtype=torch.int64).scatter_reduce(0, rid, comb["gid"], reduce="amin", include_self=False)
        assert np.array_equal((hb[1:] - hb[:-1]).numpy(), g_counts[arr["res_id"][first_atom.numpy()]])
        assert (comb["ring_res"] < D.n_res(comb)).all()
        b = D.BoundStructure(comb)  # pointers into host tensors
        assert b.struct.n_atoms == comb["x"].numel() and b.struct.n_h == comb["hx"].numel()
        # ownership is a partition of the atoms
        gathered = [None] * world
        dist.all_gather_object(gathered, own["gid"].tolist())
        flat = sorted(g for part in gathered for g in part)
        assert flat == list(range(arr["n_atoms"]))
        # round trip of the wire format
        s2 = D.deserialize(*D.serialize(own))
        for k in own:
            assert torch.equal(s2[k].to(own[k].dtype), own[k]), k
        np.save(os.path.join(out_dir, f"ok_{rank}.npy"), np.array([n_own, info["ghost_atoms"], info["sent_atoms"]]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case", ["bft", "protonated_ubq"])
def test_slab_halo_exchange_gloo(case, tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), case, str(tmp_path)), nprocs=world, join=True)
    got = [np.load(tmp_path / f"ok_{r}.npy") for r in range(world)]
    assert got[0][1] == got[1][2] and got[1][1] == got[0][2]  # what one rank sent is what the other holds as ghosts
    assert got[0][0] > 0 and got[1][0] > 0


def test_round_robin_sharding():
    for n, w in [(10, 1), (10, 2), (10, 4), (3, 8), (10000, 8)]:
        parts = [D.shard_round_robin(n, r, w) for r in range(w)]
        assert sorted(k for p in parts for k in p) == list(range(n))
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_slab_bounds_balance():
    x = torch.linspace(0, 100, 1001, dtype=torch.float64)
    b = D.slab_bounds(x, 4)
    counts = [int(((x >= b[r]) & (x < b[r + 1])).sum()) for r in range(4)]
    assert sum(counts) == 1001 and max(counts) - min(counts) <= 1


# ---- the slab path that runs behind the C ABI: what every rank hands to arp_stage_upload (host logic, no GPU) ----------
def _tag_property_holds(a):
    """|tag difference| <= 1  <=>  same chain and residue ordinals at most 1 apart, checked on neighbouring residues and on
    a sample of random residue pairs."""
    t = a["res_tag"].astype(np.int64)
    ck, ro = a["res_chain_key"].astype(np.int64), a["res_res_ordinal"].astype(np.int64)
    rng = np.random.default_rng(0)
    i = np.concatenate([np.arange(len(t) - 1), rng.integers(0, len(t), 4000)])
    j = np.concatenate([np.arange(1, len(t)), rng.integers(0, len(t), 4000)])
    close = (ck[i] == ck[j]) & (np.abs(ro[i] - ro[j]) <= 1)
    return np.array_equal(np.abs(t[i] - t[j]) <= 1, close)


def test_tiled_slab_packed_alone_matches_the_whole_tiling():
    """pack_tiled_slab: a rank tiles and packs ONLY its x-slab; chain ranks, global ids and residue tags are those the whole
    tiling gets when it is packed in one piece."""
    from arpeggia_b200 import synth

    bft = A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb"))
    nx, ny, nz, world = 4, 1, 2, 3  # uneven split of the x-indices: 1 + 1 + 2
    whole = A.Packed(synth.tile(bft, nx, ny, nz), "/", flags=2).arrays()
    assert _tag_property_holds(whole)
    seen_atoms, seen_rings = [], []
    for rank in range(world):
        a = D.pack_tiled_slab(bft, nx, ny, nz, rank, world).arrays()
        g, rg = a["atom_gid"], a["ring_gid"]
        for k in ("x", "y", "z", "attr", "chain_key", "res_ordinal"):
            assert np.array_equal(a[k], whole[k][g]), (rank, k)
        for k in ("ring_center", "ring_normal", "ring_attr", "ring_chain_key", "ring_res_ordinal"):
            assert np.array_equal(a[k], whole[k][rg]), (rank, k)
        # the tags are the whole tiling's up to one constant per copy, which keeps the property across ranks
        assert _tag_property_holds(a)
        first_res = a["res_id"][np.searchsorted(g, g)]  # identity; residues of a slab keep their order
        assert np.array_equal(first_res, a["res_id"])
        seen_atoms.append(g)
        seen_rings.append(rg)
    assert np.array_equal(np.sort(np.concatenate(seen_atoms)), np.arange(whole["n_atoms"]))
    assert np.array_equal(np.sort(np.concatenate(seen_rings)), np.arange(whole["n_rings"]))
    # tags of different ranks never collide into a false neighbourhood: compare across two slabs
    a0, a1 = D.pack_tiled_slab(bft, nx, ny, nz, 0, world).arrays(), D.pack_tiled_slab(bft, nx, ny, nz, 2, world).arrays()
    d = np.abs(a0["res_tag"].astype(np.int64)[:, None][:200] - a1["res_tag"].astype(np.int64)[None, :200])
    assert d.min() > 1


def test_cut_slab_partitions_a_packed_structure():
    """cut_slab: whole residues, every atom / ring owned by exactly one rank, indices rebased, ids and tags global."""
    bft = A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb"))
    whole = A.Packed(bft).arrays()
    for world in (1, 2, 5):
        atoms, rings = [], []
        for rank in range(world):
            o = D.cut_slab(whole, rank, world)
            g = np.asarray(o["atom_gid"])
            assert np.array_equal(o["x"], whole["x"][g]) and np.array_equal(o["attr"], whole["attr"][g])
            rab = np.asarray(o["res_atom_begin"])
            assert rab[0] == 0 and rab[-1] == len(g) and (np.diff(rab) >= 0).all()
            rid = np.repeat(np.arange(len(rab) - 1), np.diff(rab))
            assert np.array_equal(np.asarray(o["res_tag"])[rid], whole["res_tag"][whole["res_id"][g]])  # residues stay whole
            for k in ("res_cb", "res_sg"):
                v = np.asarray(o[k])
                ok = v >= 0
                assert (rid[v[ok]] == np.nonzero(ok)[0]).all()  # the CB / SG of a residue is one of its own atoms
            hb = np.asarray(o["res_h_begin"])
            assert hb[0] == 0 and hb[-1] == len(o["h_x"])
            S = D.ArrayStructure(o)  # the arp_structure handed to arp_stage_upload
            assert (S.struct.n_atoms, S.struct.n_rings) == (len(g), len(o["ring_gid"])) and bool(S.struct.res_tag) and bool(S.struct.atom_gid)
            atoms.append(g)
            rings.append(np.asarray(o["ring_gid"]))
        assert np.array_equal(np.sort(np.concatenate(atoms)), np.arange(whole["n_atoms"]))
        assert np.array_equal(np.sort(np.concatenate(rings)), np.arange(whole["n_rings"]))


def _comm_id_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import ctypes as C

        from arpeggia_b200 import _lib

        # what SlabRank does before arp_ctx_comm_init: rank 0's 128 bytes reach every rank (here a stand-in id: making a
        # real one needs NCCL and a GPU)
        cid = _lib.CommId()
        if rank == 0:
            C.memmove(C.addressof(cid), bytes(range(128)), 128)
        box = [C.string_at(C.addressof(cid), 128) if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        C.memmove(C.addressof(cid), box[0], 128)
        np.save(os.path.join(out_dir, f"id_{rank}.npy"), np.frombuffer(C.string_at(C.addressof(cid), 128), dtype=np.uint8))
        xr = D.slab_x_range(16, rank, world)
        np.save(os.path.join(out_dir, f"xr_{rank}.npy"), np.array(xr))
    finally:
        dist.destroy_process_group()


def test_comm_id_broadcast_and_x_ranges_gloo(tmp_path):
    world = 2
    mp.spawn(_comm_id_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    ids = [np.load(tmp_path / f"id_{r}.npy") for r in range(world)]
    assert np.array_equal(ids[0], np.arange(128, dtype=np.uint8)) and np.array_equal(ids[0], ids[1])  # NUL bytes survive
    xr = [np.load(tmp_path / f"xr_{r}.npy") for r in range(world)]
    assert xr[0][0] == 0 and xr[0][1] == xr[1][0] and xr[1][1] == 16
