"""Multi-rank paths on real GPUs: slab decomposition with halo exchange and batch sharding.  With one GPU the two
ranks exchange halos over gloo (CPU tensors) and both compute on cuda:0; with >= 2 GPUs they use NCCL and CUDA
tensors, one GPU per rank."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import arpeggia_b200 as A
from arpeggia_b200 import distributed as D

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEY = ["from", "to", "label", "distance"]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _atoms(case):
    from oracle import oracle as O
    from tests import fixtures as F

    if case == "bft":
        return A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb"))
    if case == "tiled":
        from arpeggia_b200 import synth
        return synth.tile(A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb")), 3, 1, 1)
    ubq = O.parse_pdb(os.path.join(ROOT, "tests", "data", "1ubq.pdb"))
    return F.all_cases(ubq)[case].view(A.ATOM_DTYPE)


def _slab_worker(rank, world, port, case, nccl, out_dir, flags=0):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = rank if nccl else 0
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl" if nccl else "gloo", rank=rank, world_size=world,
                            **({"device_id": torch.device("cuda", dev)} if nccl else {}))
    try:
        arr = A.Packed(_atoms(case), flags=2 if case == "tiled" else 0).arrays()
        whole = D.soa_from_arrays(arr, device=f"cuda:{dev}" if nccl else "cpu")
        own = D.owned_part(whole, rank, world)
        ctx = A.Context(dev)
        res = D.run_slab(ctx, own, 0.1, 6.5, flags)
        np.save(os.path.join(out_dir, f"rows_{rank}.npy"), res["rows"])
        np.save(os.path.join(out_dir, f"cnt_{rank}.npy"), np.array([res["pairs_in_cutoff"], res["halo"]["ghost_atoms"], own["x"].numel()]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case,world,flags", [("bft", 2, 0), ("protonated", 2, 0), ("altloc_protonated", 2, 0), ("tiled", 3, 0),
                                              ("bft", 2, 1), ("protonated", 2, 1), ("tiled", 3, 1)])
def test_slab_union_equals_single_gpu(case, world, flags, tmp_path):
    """flags = 1: the emit-radius screen (k_screen_atoms) with ghost atoms in the slabs."""
    nccl = torch.cuda.device_count() >= world
    mp.spawn(_slab_worker, args=(world, _free_port(), case, nccl, str(tmp_path), flags), nprocs=world, join=True)
    rows = np.concatenate([np.load(tmp_path / f"rows_{r}.npy") for r in range(world)])
    cnt = np.stack([np.load(tmp_path / f"cnt_{r}.npy") for r in range(world)])
    p = A.Packed(_atoms(case), flags=2 if case == "tiled" else 0)
    single = A.Context(0).contacts_rows(p, 0.1, 6.5)
    assert np.array_equal(np.sort(rows, order=KEY), np.sort(single["rows"], order=KEY))
    if flags == 0:
        assert cnt[:, 0].sum() == single["pairs_in_cutoff"]  # every in-cutoff pair is counted by exactly one owner
    else:
        assert (cnt[:, 0] == -1).all()
    assert cnt[:, 2].sum() == p.arrays()["n_atoms"] and (cnt[:, 1] > 0).all()


def test_batch_sharding_covers_every_structure():
    from arpeggia_b200 import synth

    bft = A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb"))
    packs = [A.Packed(synth.jittered_complex(bft, k), flags=2) for k in range(7)]
    ctx = A.Context(0)
    merged = {}
    for rank in range(3):
        part, _ = D.run_batch_shard(ctx, packs, rank, 3)
        assert set(part) == set(D.shard_round_robin(7, rank, 3))
        merged.update(part)
    assert sorted(merged) == list(range(7))
    for k, p in enumerate(packs):
        single = ctx.contacts_rows(p, 0.1, 6.5)["rows"]
        assert np.array_equal(np.sort(merged[k], order=KEY), np.sort(single, order=KEY))
    # jitter changes the contacts: structures are not copies of each other
    assert len({len(v) for v in merged.values()}) > 1
