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
