"""Multi-GPU execution of the contacts path: one process per GPU, torch.distributed for the plumbing only.

Two ways the path shards:

* batches of independent structures  -> `shard_round_robin` + `run_batch_shard`: no collective on the data path,
  only an optional all-reduce of the workload counters for the report.
* one giant assembly                  -> x-slabs with a halo.  Every rank packs ONLY the residues it owns
  (`pack_tiled_slab` for the synthetic tilings, `cut_slab` for a packed structure), uploads them and calls
  `arp_slab_exchange` (include/arpeggia_cuda.h): select / pack / unpack run as CUDA kernels on the context's
  stream, the messages move with ncclSend / ncclRecv over NVLink inside the library's own communicator -- no torch
  op and no host-side selection between "owned slab on the device" and `arp_stage_run`.  A row is emitted by the
  owner of its `from` atom / ring and carries global ids, so the union of the ranks' rows is exactly the
  single-GPU row set.  `SlabRank` wraps the calls; torch.distributed only carries the 128-byte communicator id.

The second half of this file (`soa_from_arrays` ... `exchange_halos`) is the same halo rule stated with torch
tensors on the host: test infrastructure (gloo, CPU) that the device path is checked against, not a product path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .core import Context, Packed, comm_get_id

GHOST = 0x4000
ATOM_KEYS = ("x", "y", "z", "attr", "chain_key", "res_ordinal", "res_id", "gid")
RING_KEYS = ("ring_c", "ring_n", "ring_attr", "ring_chain_key", "ring_res_ordinal", "ring_res", "ring_gid")


def shard_round_robin(n: int, rank: int, world: int) -> list[int]:
    """Structure k of a batch goes to GPU k mod world."""
    return list(range(rank, n, world))


# ---- slab jobs through the C ABI ------------------------------------------------------------------------------------
def slab_x_range(nx: int, rank: int, world: int) -> tuple[int, int]:
    """x-indices of the tiling's copies that rank `rank` owns."""
    return (nx * rank) // world, (nx * (rank + 1)) // world


def tiled_chain_ranks(base_chains, n_copies: int) -> dict:
    """Byte-order rank of every chain id "<id><5-digit copy number>" of a tiling (synth.tile): what complex.rs:129
    compares.  Built from the id strings themselves, so it holds for any base ids."""
    ids = sorted(ch + b"%05d" % c for ch in base_chains for c in range(n_copies)) if n_copies > 1 else sorted(base_chains)
    return {cid: k for k, cid in enumerate(ids)}


def pack_tiled_slab(base: np.ndarray, nx: int, ny: int, nz: int, rank: int, world: int, pitch=None, groups: str = "/") -> Packed:
    """Rank `rank`'s x-slab of the tiling of `base` (synth.tile), packed on its own: the rank never sees the rest of the
    assembly.  Chain ranks, atom / ring ids and residue tags are those of the undivided tiling, computed from the copy
    numbers (every copy packs to the same layout; ids are copy-major like the packer's hierarchy order)."""
    from . import synth

    pitch = synth.PITCH_6BFT if pitch is None else pitch
    xr = slab_x_range(nx, rank, world)
    copies = np.asarray(synth.copy_numbers(nx, ny, nz, xr), dtype=np.int64)
    n_copies = nx * ny * nz
    atoms = synth.tile(base, nx, ny, nz, pitch, x_range=xr)
    p = Packed(atoms, groups, flags=_lib.HOST_FLAG_SKIP_LOAD_FILTER)
    st = p.structure
    L = len(copies)
    assert L > 0 and st.n_atoms % L == 0 and st.n_rings % L == 0 and st.n_residues % L == 0
    na, ng, nr = st.n_atoms // L, st.n_rings // L, st.n_residues // L
    # chain ranks in the whole tiling
    ranks = tiled_chain_ranks(sorted(set(np.unique(base["chain"]).tolist())), n_copies)
    p.remap_chain_ranks([ranks[cid] for cid in p.chain_ids()])
    # ids: copy-major
    atom_gid = (copies[:, None] * na + np.arange(na)[None, :]).reshape(-1)
    ring_gid = (copies[:, None] * ng + np.arange(ng)[None, :]).reshape(-1)
    # tags: the packer's running tags restart per copy with a gap of 2 in the whole tiling as well
    tag = p.arrays()["res_tag"].astype(np.int64).reshape(L, nr)
    tag = tag - tag[:, :1]
    span = int(tag.max()) + 2
    res_tag = (copies[:, None] * span + tag).reshape(-1)
    assert atom_gid.max() < 2**31 and res_tag.max() < 2**32
    p.set_global_ids(atom_gid, ring_gid, res_tag)
    return p


class ArrayStructure:
    """arp_structure over numpy arrays in the compact per-residue form (kept alive here)."""

    def __init__(self, a: dict):
        st = _lib.Structure()
        types = dict(_lib.Structure._fields_)
        self.keep = {}

        def put(name, v, dt):
            v = np.ascontiguousarray(v, dtype=dt)
            self.keep[name] = v
            setattr(st, name, C.cast(C.c_void_p(v.ctypes.data if v.size else None), types[name]) if v.size else types[name]())

        st.n_atoms, st.n_residues, st.n_rings, st.n_h = len(a["x"]), len(a["res_cb"]), len(a["ring_attr"]), len(a["h_x"])
        for k in ("x", "y", "z", "h_x", "h_y", "h_z"):
            put(k, a[k], np.float64)
        put("ring_center", np.asarray(a["ring_center"]).reshape(-1), np.float64)
        put("ring_normal", np.asarray(a["ring_normal"]).reshape(-1), np.float64)
        for k in ("attr", "res_h_begin", "res_chain_key", "res_res_ordinal", "res_atom_begin", "ring_attr", "ring_chain_key",
                  "ring_res_ordinal", "atom_gid", "ring_gid", "res_tag"):
            put(k, a[k], np.uint32)
        for k in ("res_cb", "res_sg"):
            put(k, a[k], np.int32)
        self.struct = st

    @property
    def structure_ptr(self):
        return C.pointer(self.struct)


def cut_slab(a: dict, rank: int, world: int) -> dict:
    """The residues rank `rank` owns when the packed structure `a` (Packed.arrays()) is cut into x-slabs of equal residue
    count (by the x of a residue's first atom), as compact arrays with the whole structure's ids, chain ranks and tags."""
    rab = a["res_atom_begin"].astype(np.int64)
    nres = len(rab) - 1
    counts = np.diff(rab)
    anchor = np.where(counts > 0, a["x"][np.minimum(rab[:-1], max(len(a["x"]) - 1, 0))], -np.inf) if len(a["x"]) else np.zeros(nres)
    order = np.sort(anchor)
    inner = [order[min(nres - 1, (nres * r) // world)] for r in range(1, world)]
    bounds = [-np.inf] + inner + [np.inf]
    sel = (anchor >= bounds[rank]) & (anchor < bounds[rank + 1]) if rank < world - 1 else anchor >= bounds[rank]
    ridx = np.nonzero(sel)[0]
    amask = np.repeat(sel, counts)
    aidx = np.nonzero(amask)[0]
    new_atom = np.cumsum(amask) - 1
    out = {k: a[k][aidx] for k in ("x", "y", "z", "attr")}
    out["atom_gid"] = aidx
    out["res_atom_begin"] = np.concatenate([[0], np.cumsum(counts[ridx])])
    for k in ("res_chain_key", "res_res_ordinal", "res_tag"):
        out[k] = a[k][ridx]
    for k in ("res_cb", "res_sg"):
        v = a[k][ridx].astype(np.int64)
        out[k] = np.where(v >= 0, new_atom[np.maximum(v, 0)], -1)
    hb = a["res_h_begin"].astype(np.int64)
    hcounts = np.diff(hb)
    hidx = np.nonzero(np.repeat(sel, hcounts))[0]
    for k in ("h_x", "h_y", "h_z"):
        out[k] = a[k][hidx]
    out["res_h_begin"] = np.concatenate([[0], np.cumsum(hcounts[ridx])])
    # a ring belongs to the residue with its (chain key, ordinal)
    rkey = (a["res_chain_key"].astype(np.int64) << 32) | a["res_res_ordinal"].astype(np.int64)
    o = np.argsort(rkey)
    want = (a["ring_chain_key"].astype(np.int64) << 32) | a["ring_res_ordinal"].astype(np.int64)
    ring_res = o[np.searchsorted(rkey[o], want)] if len(want) else np.zeros(0, dtype=np.int64)
    gidx = np.nonzero(sel[ring_res])[0] if len(want) else np.zeros(0, dtype=np.int64)
    for k in ("ring_center", "ring_normal", "ring_attr", "ring_chain_key", "ring_res_ordinal"):
        out[k] = a[k][gidx]
    out["ring_gid"] = gidx
    return out


def broadcast_comm_id(rank: int, group=None) -> "_lib.CommId":
    """One rank makes the communicator id, torch.distributed carries its 128 bytes to the others (any backend)."""
    cid = comm_get_id() if rank == 0 else _lib.CommId()
    box = [C.string_at(C.addressof(cid), 128) if rank == 0 else None]  # the raw 128 bytes, NULs included
    dist.broadcast_object_list(box, src=0, group=group)
    C.memmove(C.addressof(cid), box[0], 128)
    return cid


class SlabRank:
    """One rank of a slab job: a context with its own NCCL communicator, the owned part resident on the device."""

    def __init__(self, ctx: Context, rank: int, world: int, group=None):
        self.ctx, self.rank, self.world = ctx, rank, world
        if world > 1:
            ctx.comm_init(broadcast_comm_id(rank, group), rank, world)

    def upload(self, own):
        """own: a Packed (pack_tiled_slab) or an ArrayStructure (cut_slab)."""
        self.ctx.upload([own])

    def exchange(self, vdw_comp=0.1, dist_cutoff=6.5) -> dict:
        return self.ctx.slab_exchange(vdw_comp, dist_cutoff)

    def run(self, vdw_comp=0.1, dist_cutoff=6.5, flags=0) -> dict:
        """Kernels on owned + ghosts; rows come back with global ids."""
        self.ctx.run(vdw_comp, dist_cutoff, flags)
        return self.ctx.download()


# ---- SoA dicts ------------------------------------------------------------------------------------------------
def soa_from_arrays(a: dict, device="cpu") -> dict:
    """Packed.arrays() (numpy) -> SoA dict of torch tensors with global ids and the ring -> residue link."""
    t = lambda v, dt: torch.from_numpy(np.ascontiguousarray(v).astype(dt)).to(device)
    n = int(a["n_atoms"])
    s = {
        "x": t(a["x"], np.float64), "y": t(a["y"], np.float64), "z": t(a["z"], np.float64),
        "attr": t(a["attr"].view(np.int32), np.int32), "chain_key": t(a["chain_key"].view(np.int32), np.int32),
        "res_ordinal": t(a["res_ordinal"].view(np.int32), np.int32), "res_id": t(a["res_id"], np.int64),
        "gid": torch.arange(n, dtype=torch.int64, device=device),
        "res_cb": t(a["res_cb"], np.int64), "res_sg": t(a["res_sg"], np.int64), "h_begin": t(a["res_h_begin"], np.int64),
        "hx": t(a.get("h_x", np.zeros(0)), np.float64), "hy": t(a.get("h_y", np.zeros(0)), np.float64),
        "hz": t(a.get("h_z", np.zeros(0)), np.float64),
        "ring_c": t(a["ring_center"].reshape(-1, 3), np.float64), "ring_n": t(a["ring_normal"].reshape(-1, 3), np.float64),
        "ring_attr": t(a["ring_attr"].view(np.int32), np.int32), "ring_chain_key": t(a["ring_chain_key"].view(np.int32), np.int32),
        "ring_res_ordinal": t(a["ring_res_ordinal"].view(np.int32), np.int32),
        "ring_gid": torch.arange(int(a["n_rings"]), dtype=torch.int64, device=device),
    }
    # ring -> residue: the residue with the same (chain key, ordinal)
    key = (a["chain_key"].astype(np.int64) << 32) | a["res_ordinal"].astype(np.int64)
    first = np.unique(a["res_id"], return_index=True)[1]
    rkey, rid = key[first], a["res_id"][first].astype(np.int64)
    order = np.argsort(rkey)
    want = (a["ring_chain_key"].astype(np.int64) << 32) | a["ring_res_ordinal"].astype(np.int64)
    s["ring_res"] = t(rid[order][np.searchsorted(rkey[order], want)] if len(want) else np.zeros(0), np.int64)
    return s


def n_res(s: dict) -> int:
    return int(s["res_cb"].numel())


def extract(s: dict, res_mask: torch.Tensor) -> dict:
    """Sub-structure made of the residues selected by `res_mask` (whole residues: atoms, hydrogens, rings).
    Index lists are built once (four host syncs for their sizes) and every array is gathered with index_select."""
    dev = s["x"].device
    amask = res_mask[s["res_id"]]
    res_new = torch.cumsum(res_mask.to(torch.int64), 0) - 1
    atom_new = torch.cumsum(amask.to(torch.int64), 0) - 1
    aidx = torch.nonzero(amask).squeeze(1)
    ridx = torch.nonzero(res_mask).squeeze(1)
    out = {k: s[k].index_select(0, aidx) for k in ("x", "y", "z", "attr", "chain_key", "res_ordinal", "gid")}
    out["res_id"] = res_new[s["res_id"].index_select(0, aidx)]
    for k in ("res_cb", "res_sg"):
        v = s[k].index_select(0, ridx)
        out[k] = torch.where(v >= 0, atom_new[v.clamp(min=0)], v)
    counts = (s["h_begin"][1:] - s["h_begin"][:-1])
    if s["hx"].numel():
        hidx = torch.nonzero(torch.repeat_interleave(res_mask, counts)).squeeze(1)
        for k in ("hx", "hy", "hz"):
            out[k] = s[k].index_select(0, hidx)
    else:
        for k in ("hx", "hy", "hz"):
            out[k] = s[k][:0]
    out["h_begin"] = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(counts.index_select(0, ridx), 0)])
    if s["ring_res"].numel():
        gidx = torch.nonzero(res_mask[s["ring_res"]]).squeeze(1)
    else:
        gidx = torch.zeros(0, dtype=torch.int64, device=dev)
    for k in ("ring_c", "ring_n", "ring_attr", "ring_chain_key", "ring_res_ordinal", "ring_gid"):
        out[k] = s[k].index_select(0, gidx)
    out["ring_res"] = res_new[s["ring_res"].index_select(0, gidx)]
    return out


def concat(a: dict, b: dict) -> dict:
    """a followed by b (indices of b shifted)."""
    na, nra, nha = a["x"].numel(), n_res(a), a["hx"].numel()
    out = {k: torch.cat([a[k], b[k]]) for k in ("x", "y", "z", "attr", "chain_key", "res_ordinal", "gid", "hx", "hy", "hz",
                                                "ring_c", "ring_n", "ring_attr", "ring_chain_key", "ring_res_ordinal", "ring_gid")}
    out["res_id"] = torch.cat([a["res_id"], b["res_id"] + nra])
    out["ring_res"] = torch.cat([a["ring_res"], b["ring_res"] + nra])
    for k in ("res_cb", "res_sg"):
        out[k] = torch.cat([a[k], torch.where(b[k] >= 0, b[k] + na, b[k])])
    out["h_begin"] = torch.cat([a["h_begin"], b["h_begin"][1:] + nha])
    return out


def mark_ghost(s: dict) -> dict:
    s = dict(s)
    s["attr"] = s["attr"] | GHOST
    s["ring_attr"] = s["ring_attr"] | GHOST
    return s


# ---- wire format: one float64 and one int64 tensor per message -----------------------------------------------------
_F_KEYS = ("x", "y", "z", "hx", "hy", "hz")
_I_KEYS = ("attr", "chain_key", "res_ordinal", "res_id", "gid", "res_cb", "res_sg", "h_begin", "ring_attr", "ring_chain_key",
           "ring_res_ordinal", "ring_res", "ring_gid")


def serialize(s: dict):
    f = torch.cat([s[k] for k in _F_KEYS] + [s["ring_c"].reshape(-1), s["ring_n"].reshape(-1)])
    i = torch.cat([s[k].to(torch.int64) for k in _I_KEYS])
    sizes = torch.tensor([s["x"].numel(), n_res(s), s["hx"].numel(), s["ring_gid"].numel(), f.numel(), i.numel()], dtype=torch.int64)
    return sizes, f, i


def deserialize(sizes, f, i) -> dict:
    na, nr, nh, ng = (int(v) for v in sizes[:4])
    out, p = {}, 0
    for k, n in zip(_F_KEYS, (na, na, na, nh, nh, nh)):
        out[k] = f[p:p + n]
        p += n
    out["ring_c"] = f[p:p + 3 * ng].reshape(-1, 3)
    out["ring_n"] = f[p + 3 * ng:p + 6 * ng].reshape(-1, 3)
    p = 0
    lens = {"attr": na, "chain_key": na, "res_ordinal": na, "res_id": na, "gid": na, "res_cb": nr, "res_sg": nr, "h_begin": nr + 1,
            "ring_attr": ng, "ring_chain_key": ng, "ring_res_ordinal": ng, "ring_res": ng, "ring_gid": ng}
    i32 = {"attr", "chain_key", "res_ordinal", "ring_attr", "ring_chain_key", "ring_res_ordinal"}
    for k in _I_KEYS:
        v = i[p:p + lens[k]]
        out[k] = v.to(torch.int32) if k in i32 else v
        p += lens[k]
    return out


# ---- slab decomposition -------------------------------------------------------------------------------------------------
def slab_bounds(anchor_x: torch.Tensor, world: int) -> list[float]:
    """Faces that split the residues (by the x of their first atom) into `world` slabs of equal residue count."""
    xs = torch.sort(anchor_x.to(torch.float64)).values
    n = xs.numel()
    inner = [float(xs[min(n - 1, (n * r) // world)]) for r in range(1, world)]
    return [-float("inf")] + inner + [float("inf")]


def residue_anchor_x(s: dict) -> torch.Tensor:
    """x of the first atom of every residue (atoms of a residue are contiguous, res_id is non-decreasing)."""
    nr = n_res(s)
    first = torch.full((nr,), s["x"].numel(), dtype=torch.int64, device=s["x"].device)
    idx = torch.arange(s["x"].numel(), dtype=torch.int64, device=s["x"].device)
    first = first.scatter_reduce(0, s["res_id"], idx, reduce="amin", include_self=True)
    return s["x"][first.clamp(max=max(s["x"].numel() - 1, 0))]


def owned_part(s: dict, rank: int, world: int, bounds=None) -> dict:
    """The residues rank `rank` owns when the whole structure `s` is cut into x-slabs."""
    ax = residue_anchor_x(s)
    if bounds is None:
        bounds = slab_bounds(ax, world)
    mask = (ax >= bounds[rank]) & (ax < bounds[rank + 1])
    if rank == world - 1:
        mask = ax >= bounds[rank]
    return extract(s, mask)


def halo_selection(own: dict, lo: float, hi: float) -> torch.Tensor:
    """Residues of `own` with at least one atom whose x lies in [lo, hi]."""
    inside = (own["x"] >= lo) & (own["x"] <= hi)
    flag = torch.zeros(n_res(own), dtype=torch.bool, device=own["x"].device)
    flag[own["res_id"][inside]] = True
    return flag


def exchange_halos(own: dict, width: float, group=None) -> tuple[dict, dict]:
    """Send every neighbour the residues that reach into its range widened by `width`; returns (combined, info).
    combined = owned atoms first, ghosts (ARP_ATTR_GHOST) after.  Works for any process group: NCCL moves the CUDA
    tensors GPU to GPU over NVLink, gloo moves CPU tensors in the tests."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    dev = own["x"].device
    info = {"ghost_atoms": 0, "sent_atoms": 0, "bytes_sent": 0}
    if world == 1:
        return own, info
    # every rank's actual x-range of owned atoms (residues stick out of their slab)
    if own["x"].numel() > 0:
        mine = torch.stack([own["x"].min(), own["x"].max()]).to(torch.float64)  # stays on the device: no host sync here
    else:
        mine = torch.tensor([float("inf"), float("-inf")], dtype=torch.float64, device=dev)
    ranges = [torch.zeros(2, dtype=torch.float64, device=dev) for _ in range(world)]
    dist.all_gather(ranges, mine, group=group)
    ranges = [(r[0], r[1]) for r in torch.stack(ranges).cpu().tolist()]  # one sync for every rank's range
    sends = {}
    for peer in range(world):
        if peer == rank or ranges[peer][0] > ranges[peer][1]:
            continue
        if ranges[peer][0] - width > ranges[rank][1] or ranges[peer][1] + width < ranges[rank][0]:
            continue  # not a neighbour: nothing of mine reaches into its range (ranges are already on the host)
        sel = halo_selection(own, ranges[peer][0] - width, ranges[peer][1] + width)
        sends[peer] = serialize(extract(own, sel))
    # sizes first (every pair of ranks talks, mostly empty), then the payloads
    size_out = [sends[p][0].to(dev) if p in sends else torch.zeros(6, dtype=torch.int64, device=dev) for p in range(world)]
    size_in = [torch.zeros(6, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_to_all(size_in, size_out, group=group) if dev.type == "cuda" else _all_to_all_p2p(size_in, size_out, rank, world, group)
    ops, recv = [], {}
    for peer in range(world):
        if peer == rank:
            continue
        nf, ni = int(size_in[peer][4]), int(size_in[peer][5])
        if int(size_in[peer][0]) > 0:
            recv[peer] = (size_in[peer].cpu(), torch.empty(nf, dtype=torch.float64, device=dev), torch.empty(ni, dtype=torch.int64, device=dev))
            ops += [dist.P2POp(dist.irecv, recv[peer][1], peer, group), dist.P2POp(dist.irecv, recv[peer][2], peer, group)]
        if peer in sends and int(sends[peer][0][0]) > 0:
            ops += [dist.P2POp(dist.isend, sends[peer][1], peer, group), dist.P2POp(dist.isend, sends[peer][2], peer, group)]
            info["sent_atoms"] += int(sends[peer][0][0])
            info["bytes_sent"] += 8 * (sends[peer][1].numel() + sends[peer][2].numel())
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    combined = own
    for peer in sorted(recv):
        ghost = mark_ghost(deserialize(*recv[peer]))
        info["ghost_atoms"] += ghost["x"].numel()
        combined = concat(combined, ghost)
    return combined, info


def _all_to_all_p2p(out_list, in_list, rank, world, group):
    """all_to_all for backends without it (gloo): pairwise isend / irecv."""
    ops = []
    for peer in range(world):
        if peer == rank:
            out_list[peer].copy_(in_list[peer])
            continue
        ops += [dist.P2POp(dist.isend, in_list[peer], peer, group), dist.P2POp(dist.irecv, out_list[peer], peer, group)]
    for req in dist.batch_isend_irecv(ops):
        req.wait()


# ---- SoA dict -> arp_structure ----------------------------------------------------------------------------------------------
class BoundStructure:
    """arp_structure whose pointers are the data pointers of (host or device) torch tensors kept alive here."""

    def __init__(self, s: dict):
        c = lambda v, dt: v.to(dt).contiguous()
        self.t = {
            "x": c(s["x"], torch.float64), "y": c(s["y"], torch.float64), "z": c(s["z"], torch.float64),
            "attr": c(s["attr"], torch.int32), "chain_key": c(s["chain_key"], torch.int32), "res_ordinal": c(s["res_ordinal"], torch.int32),
            "res_id": c(s["res_id"], torch.int32), "res_h_begin": c(s["h_begin"], torch.int32),
            "h_x": c(s["hx"], torch.float64), "h_y": c(s["hy"], torch.float64), "h_z": c(s["hz"], torch.float64),
            "res_cb": c(s["res_cb"], torch.int32), "res_sg": c(s["res_sg"], torch.int32),
            "ring_center": c(s["ring_c"].reshape(-1), torch.float64), "ring_normal": c(s["ring_n"].reshape(-1), torch.float64),
            "ring_attr": c(s["ring_attr"], torch.int32), "ring_chain_key": c(s["ring_chain_key"], torch.int32),
            "ring_res_ordinal": c(s["ring_res_ordinal"], torch.int32),
        }
        st = _lib.Structure()
        st.n_atoms, st.n_residues, st.n_rings, st.n_h = s["x"].numel(), n_res(s), s["ring_gid"].numel(), s["hx"].numel()
        types = dict(_lib.Structure._fields_)
        for k, v in self.t.items():
            setattr(st, k, C.cast(C.c_void_p(v.data_ptr()), types[k]))
        self.struct = st
        self.gid = s["gid"]
        self.ring_gid = s["ring_gid"]

    @property
    def ptr(self):
        return C.pointer(self.struct)


def globalize_rows(rows: np.ndarray, gid: torch.Tensor, ring_gid: torch.Tensor) -> np.ndarray:
    """Local atom / ring indices of device rows -> ids in the undivided structure."""
    g, rg = gid.cpu().numpy(), ring_gid.cpu().numpy()
    out = rows.copy()
    for col in ("from", "to"):
        v = rows[col]
        ring = (v & _lib.RING_FLAG) != 0
        idx = (v & ~np.uint32(_lib.RING_FLAG)).astype(np.int64)
        res = np.empty(len(v), dtype=np.uint32)
        res[ring] = rg[idx[ring]].astype(np.uint32) | np.uint32(_lib.RING_FLAG)
        res[~ring] = g[idx[~ring]].astype(np.uint32)
        out[col] = res
    return out


def run_batch_shard(ctx, packs, rank: int, world: int, vdw_comp=0.1, dist_cutoff=6.5, flags=0):
    """This rank's share of a batch of independent structures (round robin); returns {structure index: rows}."""
    from .core import split_rows

    mine = shard_round_robin(len(packs), rank, world)
    if not mine:
        return {}, {"pairs_in_cutoff": 0, "pairs_screened": 0, "n_rows": 0}
    res = ctx.contacts_rows([packs[k] for k in mine], vdw_comp, dist_cutoff, flags)
    parts = split_rows(res)
    return dict(zip(mine, parts)), {"pairs_in_cutoff": res["pairs_in_cutoff"], "pairs_screened": res["pairs_screened"],
                                    "n_rows": len(res["rows"])}
