"""Multi-rank paths on real GPUs: slab decomposition with the device-side halo exchange (arp_slab_exchange) and batch
sharding.  Two transports are exercised: one process driving several contexts (arp_ctx_create_multi; on a single GPU the
contexts share the device and messages move as peer copies -- the same kernels and byte layout), and one process per GPU
with NCCL send / recv inside the library's communicator (needs >= 2 GPUs)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import arpeggia_b200 as A
from arpeggia_b200 import core, distributed as D

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


def _single(case):
    p = A.Packed(_atoms(case), flags=2 if case == "tiled" else 0)
    return p, A.Context(0).contacts_rows(p, 0.1, 6.5)


def _check_union(rows, cnt, infos, single, p, flags):
    assert np.array_equal(np.sort(rows, order=KEY), np.sort(single["rows"], order=KEY))
    if flags == 0:
        assert sum(cnt) == single["pairs_in_cutoff"]  # every in-cutoff pair is counted by exactly one owner
    else:
        assert all(c == -1 for c in cnt)
    assert sum(i["own_atoms"] for i in infos) == p.arrays()["n_atoms"] and all(i["ghost_atoms"] > 0 for i in infos)
    # what one rank sent is what the others hold as ghosts
    assert sum(i["sent_atoms"] for i in infos) == sum(i["ghost_atoms"] for i in infos)
    assert sum(i["bytes_sent"] for i in infos) == sum(i["bytes_received"] for i in infos)


@pytest.mark.parametrize("case,world,flags", [("bft", 2, 0), ("protonated", 2, 0), ("altloc_protonated", 2, 0), ("tiled", 3, 0),
                                              ("bft", 2, 1), ("protonated", 2, 1), ("tiled", 3, 1), ("bft", 4, 0)])
def test_slab_union_equals_single_gpu_one_process(case, world, flags):
    """arp_ctx_create_multi + arp_slab_exchange_multi: every device-side step of the halo exchange (mark, count, offsets,
    pack, unpack), twice in a row (the second exchange drops the first one's ghosts)."""
    p, single = _single(case)
    whole = p.arrays()
    ndev = torch.cuda.device_count()
    devices = [r % ndev for r in range(world)]
    ctxs = core.create_multi(devices)
    parts = [D.ArrayStructure(D.cut_slab(whole, r, world)) for r in range(world)]
    for cx, part in zip(ctxs, parts):
        cx.upload([part])
    for rep in range(2):
        infos = core.slab_exchange_multi(ctxs, 0.1, 6.5)
        rows, cnt = [], []
        for cx in ctxs:
            cx.run(0.1, 6.5, flags)
            res = cx.download()
            rows.append(res["rows"])
            cnt.append(res["pairs_in_cutoff"])
        _check_union(np.concatenate(rows), cnt, infos, single, p, flags)
        assert all(i["transport"] == (2 if ndev < world else 1) for i in infos)
    for cx in ctxs:
        cx.close()


def test_ghosts_are_the_halo_rule_of_the_host_statement():
    """The device selection against the torch statement of the same rule (distributed.exchange_halos, test infrastructure):
    the same ghost atoms and rings on every rank."""
    p, _ = _single("bft")
    whole = p.arrays()
    world = 2
    ctxs = core.create_multi([0] * world)
    for r, cx in enumerate(ctxs):
        cx.upload([D.ArrayStructure(D.cut_slab(whole, r, world))])
    infos = core.slab_exchange_multi(ctxs, 0.1, 6.5)
    soa = D.soa_from_arrays(whole)
    owned = [D.owned_part(soa, r, world) for r in range(world)]
    width = 6.5 * (1.0 + 1e-6) + 1e-6
    for r in range(world):
        lo, hi = float(owned[r]["x"].min()), float(owned[r]["x"].max())
        want = 0
        for q in range(world):
            if q != r:
                sel = D.halo_selection(owned[q], lo - width, hi + width)
                want += int(sel[owned[q]["res_id"]].sum())
        assert infos[r]["ghost_atoms"] == want
        assert infos[r]["own_atoms"] == owned[r]["x"].numel()
    for cx in ctxs:
        cx.close()


def _nccl_worker(rank, world, port, case, out_dir, flags):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # carries the 128-byte communicator id only
    try:
        whole = A.Packed(_atoms(case), flags=2 if case == "tiled" else 0).arrays()
        sr = D.SlabRank(A.Context(rank), rank, world)
        sr.upload(D.ArrayStructure(D.cut_slab(whole, rank, world)))
        info = sr.exchange(0.1, 6.5)
        res = sr.run(0.1, 6.5, flags)
        np.save(os.path.join(out_dir, f"rows_{rank}.npy"), res["rows"])
        np.save(os.path.join(out_dir, f"cnt_{rank}.npy"), np.array([res["pairs_in_cutoff"]] + [info[k] for k in
                ("own_atoms", "ghost_atoms", "sent_atoms", "bytes_sent", "bytes_received", "transport")]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case,world,flags", [("bft", 2, 0), ("protonated", 2, 1), ("tiled", 2, 1)])
def test_slab_union_equals_single_gpu_nccl(case, world, flags, tmp_path):
    """One process per GPU, ncclSend / ncclRecv on the contexts' streams (arp_ctx_comm_init + arp_slab_exchange)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (the one-process test above covers the kernels on a single GPU)")
    mp.spawn(_nccl_worker, args=(world, _free_port(), case, str(tmp_path), flags), nprocs=world, join=True)
    rows = np.concatenate([np.load(tmp_path / f"rows_{r}.npy") for r in range(world)])
    cnt = np.stack([np.load(tmp_path / f"cnt_{r}.npy") for r in range(world)])
    p, single = _single(case)
    infos = [dict(zip(("own_atoms", "ghost_atoms", "sent_atoms", "bytes_sent", "bytes_received", "transport"), c[1:])) for c in cnt]
    _check_union(rows, cnt[:, 0].tolist(), infos, single, p, flags)
    assert all(i["transport"] == 1 for i in infos)


def test_tiled_slab_packed_alone_equals_the_whole_assembly():
    """BASELINE config 5 in small: every rank tiles and packs ONLY its own x-slab (pack_tiled_slab); the union of the
    ranks' rows is the row set of the whole assembly packed and run on one GPU."""
    from arpeggia_b200 import synth

    bft = A.parse_pdb(os.path.join(ROOT, "tests", "data", "6bft.pdb"))
    nx, ny, nz, world = 4, 2, 1, 4
    whole = A.Packed(synth.tile(bft, nx, ny, nz), flags=2)
    single = A.Context(0).contacts_rows(whole, 0.1, 6.5, 1)
    ndev = torch.cuda.device_count()
    ctxs = core.create_multi([r % ndev for r in range(world)])
    for r, cx in enumerate(ctxs):
        cx.upload([D.pack_tiled_slab(bft, nx, ny, nz, r, world)])
    infos = core.slab_exchange_multi(ctxs, 0.1, 6.5)
    rows = []
    for cx in ctxs:
        cx.run(0.1, 6.5, 1)
        rows.append(cx.download()["rows"])
    rows = np.concatenate(rows)
    assert np.array_equal(np.sort(rows, order=KEY), np.sort(single["rows"], order=KEY))
    assert [i["n_peers"] for i in infos] == [1, 2, 2, 1]
    # ... and against the ORACLE: the union of the ranks' rows (global ids = indices of the whole packed tiling), decoded and
    # sorted by the host table, is the oracle's table of the whole assembly
    from oracle import oracle as O
    from tests.util import assert_tables_equal

    atoms = synth.tile(bft, nx, ny, nz)
    want = O.contacts(atoms.view(O.ATOM_DTYPE), "/", 0.1, 6.5, flags=O.FLAG_SKIP_LOAD_FILTER, num_threads=0)
    assert_tables_equal(whole.build_table(rows), want, "4 slabs vs oracle")
    for cx in ctxs:
        cx.close()


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
