#!/usr/bin/env python
"""bench.py -- contacts hot path on B200 (driver contract; see the task prompt and DESIGN.md "Measurement").

A "step" = one pass of the hot path (cell-list build + neighbour/classify kernel + ring kernels) over the
synthetic 1 022 500-atom protein-density assembly of BASELINE.json configs[2] (6bft tiled 5x5x5), inputs already
resident in HBM.  With N > 1 ranks every rank owns one such assembly (sharding by structure, no collective on
the data path): weak scaling.  Metric: atom pairs screened per second = ordered atom pairs within dist_cutoff
(what the reference's R*-tree query returns and its classification loop consumes), a property of the workload
and therefore identical for every implementation; contacts/s, candidate tests/s and atoms/s ride along.

The timed path is the one arpeggia.contacts() runs: the emit-radius screen (arp_params.flags = 1: no label of
complex.rs:217-296 can fire beyond r_emit = 4.5 A at the default vdw_comp, so cells and the screen use r_emit instead
of dist_cutoff = 6.5 A; rows are identical, tests/test_gpu_parity.py).  The pair count of the metric is then the
workload constant, counted by one plain run (flags 0) BEFORE the timed region, as SURVEY.md 8(d) defines it
("implementation-independent pairs-in-cutoff/s so that pruning does not deflate the figure"); the same loop in the
counting mode (flags 0: screen at dist_cutoff, every in-cutoff pair counted on the device) is timed right after and
reported under "counting_mode".  --flags 0 makes the counting mode the headline instead.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (GPU)
  python bench.py --impl reference [...]                         # the CPU restatement of the reference
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

# torchrun exports OMP_NUM_THREADS=1; the host packer (OpenMP) should use this rank's share of the cores instead.
# Must happen before any library that carries libgomp is loaded.
_WORLD = int(os.environ.get("WORLD_SIZE", "1"))
if _WORLD > 1 or "OMP_NUM_THREADS" not in os.environ:
    os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // _WORLD))
os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")  # idle OpenMP workers sleep instead of spinning next to the launching thread


def bind_to_gpu_cores(local_rank: int, world: int):
    """Pin this rank to the cores of the NUMA node its GPU hangs off (sysfs local_cpulist), split among the ranks that
    share the node: pinned host buffers are then allocated next to the GPU and the launching thread stays there.
    Returns a description for the bench line, or None when the topology cannot be read."""
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(local_rank)],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        bus = out[-12:] if len(out) >= 12 else out  # 00000000:1b:00.0 -> 0000:1b:00.0
        cpus = Path(f"/sys/bus/pci/devices/{bus}/local_cpulist").read_text().strip()
        node = Path(f"/sys/bus/pci/devices/{bus}/numa_node").read_text().strip()
        ids = []
        for part in cpus.split(","):
            a, _, b = part.partition("-")
            ids += list(range(int(a), int(b or a) + 1))
        ids = sorted(set(ids) & set(os.sched_getaffinity(0)))
        if not ids:
            return None
        per = max(1, (os.cpu_count() or 1) // max(world, 1))
        k = (local_rank * per) % max(len(ids), 1)
        mine = (ids[k:] + ids[:k])[:per] if world > 1 else ids
        os.sched_setaffinity(0, mine)
        return {"numa_node": int(node), "cores": len(mine), "first_core": mine[0]}
    except Exception:
        return None

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "atom_pairs_screened_per_s"
UNIT = "pairs/s"
WORKLOAD = "6bft tiled 5x5x5 = 1,022,500 atoms (BASELINE.json configs[2]), groups '/', dist_cutoff 6.5, vdw_comp 0.1"


def build_workload(dense: bool = False):
    import arpeggia_b200 as A
    from arpeggia_b200 import synth

    base = A.parse_pdb(ROOT / "tests" / "data" / "6bft.pdb")
    return synth.assembly_1m(base, dense=dense)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe): one streaming
    nvidia-smi process (-lms 50), samples time-stamped on arrival; summary() keeps the ones inside [t0, t1]."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index),
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                f = [x.strip() for x in line.split(",")]
                if len(f) >= 7:
                    self.rows.append((time.perf_counter(), f))
        except Exception:
            pass

    def summary(self, t0: float, t1: float):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=3)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        inside = [f for t, f in self.rows if t0 <= t <= t1]
        note = "inside the timed region"
        if not inside:  # region shorter than the sampling period: the samples that bracket it
            inside = [f for t, f in self.rows if t0 - 0.3 <= t <= t1 + 0.3] or [self.rows[-1][1]]
            note = "timed region shorter than the 50 ms sampling period: samples within 0.3 s of it"
        sm = sorted(float(r[0]) for r in inside)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[3 + k].lower().startswith("active") for r in inside)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(inside[0][1]), "reasons": reasons, "samples": len(sm), "window": note}


def cpu_reference(atoms, sample_copies: int, threads: int):
    """The CPU restatement of the reference (oracle/, "port"): hot-path seconds on a bounded sample."""
    from oracle import oracle as O

    n_per = len(atoms) // 125
    sample = atoms[: n_per * sample_copies]
    r = O.contacts(sample.view(O.ATOM_DTYPE), "/", 0.1, 6.5, flags=O.FLAG_NO_TABLE | O.FLAG_SKIP_LOAD_FILTER, num_threads=threads)
    hot = r["t_search_s"] + r["t_classify_s"] + r["t_rings_s"]
    return {"pairs": r["pairs_in_cutoff"], "rows": r["n_rows"], "atoms": r["n_atoms"], "hot_s": hot,
            "sample": f"first {sample_copies} of the 125 6bft copies = {len(sample):,} atoms; neighbour search + classification + ring loops "
                      f"(hierarchy build and table excluded)"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    atoms = build_workload()
    threads = os.cpu_count() or 1
    copies = 125  # the whole workload: about 1-5 s per step on 8+ cores
    for _ in range(max(args.warmup, 0) and 1):
        cpu_reference(atoms, copies, threads)
    t_tot, pairs, rows = 0.0, 0, 0
    for _ in range(args.steps):
        r = cpu_reference(atoms, copies, threads)
        t_tot += r["hot_s"]
        pairs += r["pairs"]
        rows += r["rows"]
    value = pairs / t_tot
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_tot / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU restatement (oracle/) of the reference's Rust path; the Rust crate cannot be built here"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": r["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "contacts_per_s": rows / t_tot, "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def pinned_structure(A, packed, arena=True):
    """Copy the packed SoA into page-locked memory (arp_host_alloc) and return (Structure, keepalive, bytes).  arena: one
    allocation, the arrays one after the other (256-byte aligned) -- the library then moves them with one copy
    (arp_stats.upload_copies); otherwise one allocation per array."""
    from arpeggia_b200 import _lib

    L = _lib.lib()
    src = packed.structure
    dst = _lib.Structure()
    keep, total = [], 0
    for name, ctype in _lib.Structure._fields_:
        v = getattr(src, name)
        if name.startswith("n_") or name.startswith("bbox_") or name == "has_bbox":
            setattr(dst, name, v)
    # chain_key / res_ordinal travel in their compact per-residue form (arp_structure.res_chain_key); the per-atom
    # arrays stay NULL and are neither copied here nor uploaded
    # ... and res_id as the residues' atom ranges (res_atom_begin)
    # ... and the coordinates on the 0.001 A lattice (x_milli) when the packer offers that form: 12 instead of 24 B / atom
    # ... or as 21-bit lattice words with 16-bit attribute words (xyz_milli21 / attr16): 10 instead of 16 B / atom
    if src.xyz_milli21 and src.attr16:
        sizes = {"xyz_milli21": 8, "attr16": 2}
        for k in range(3):
            dst.xyz_milli21_origin[k] = src.xyz_milli21_origin[k]
    else:
        sizes = {"x_milli": 4, "y_milli": 4, "z_milli": 4, "attr": 4} if src.x_milli else {"x": 8, "y": 8, "z": 8, "attr": 4}
    n, nr, ng, nh = src.n_atoms, src.n_residues, src.n_rings, src.n_h
    lens = {**{k: n * s for k, s in sizes.items()}, "res_chain_key": 4 * nr, "res_res_ordinal": 4 * nr, "res_atom_begin": 4 * (nr + 1), "res_h_begin": 4 * (nr + 1), "h_x": 8 * nh, "h_y": 8 * nh, "h_z": 8 * nh,
            "res_cb": 4 * nr, "res_sg": 4 * nr, "ring_center": 24 * ng, "ring_normal": 24 * ng, "ring_attr": 4 * ng,
            "ring_chain_key": 4 * ng, "ring_res_ordinal": 4 * ng}
    if src.res_tag:
        lens["res_tag"] = 4 * nr  # the packer's residue tags: without them the library derives them on the device
    block, at = None, 0
    if arena:
        block = C.c_void_p()
        _lib.check(L.arp_host_alloc(C.byref(block), sum((b + 255) // 256 * 256 for b in lens.values()) + 256))
        keep.append(block)
    for name, nbytes in lens.items():
        if arena:
            p = C.c_void_p(block.value + at)
            at += (nbytes + 255) // 256 * 256
        else:
            p = C.c_void_p()
            _lib.check(L.arp_host_alloc(C.byref(p), max(nbytes, 1)))
            keep.append(p)
        if nbytes:
            C.memmove(p, C.cast(getattr(src, name), C.c_void_p), nbytes)
        setattr(dst, name, C.cast(p, dict(_lib.Structure._fields_)[name]))
        total += nbytes
    return dst, keep, total


# single-GPU totals of the two multi-GPU configurations (profiles/r1t_big100m.json, r1t_batch.json): what the union of
# the ranks' results must equal; ms_one_gpu = the same job on ONE B200 with the current kernels
# (profiles/r2c_bench_1gpu_multi.json: `bench.py --multi on` at N = 1)
SLAB_EXPECT = {(16, 32, 24): {"rows": 88_915_968, "pairs_in_cutoff": 4_871_872_512, "atoms": 100_515_840, "ms_one_gpu": 21.14}}
BATCH_EXPECT = {10000: {"rows": 42_379_378, "pairs_in_cutoff": 2_332_010_892, "atoms": 48_780_000, "ms_one_gpu": 9.61}}


def bench_slab(ctx_device: int, rank: int, world: int, tiling, reps: int = 3):
    """BASELINE config 5: the 6bft tiling cut into `world` x-slabs.  Every rank tiles and packs ONLY its own slab,
    uploads it, and the halo exchange runs on the device behind the C ABI (arp_slab_exchange: CUDA select / pack /
    unpack kernels + ncclSend / ncclRecv on the context's stream).  Returns this rank's numbers; reduced by the caller."""
    import torch
    import torch.distributed as dist

    import arpeggia_b200 as A
    from arpeggia_b200 import distributed as D

    nx, ny, nz = tiling
    base = A.parse_pdb(ROOT / "tests" / "data" / "6bft.pdb")
    t0 = time.perf_counter()
    own = D.pack_tiled_slab(base, nx, ny, nz, rank, world)
    s_pack = time.perf_counter() - t0
    ctx = A.Context(ctx_device)
    sr = D.SlabRank(ctx, rank, world)
    t0 = time.perf_counter()
    sr.upload(own)
    s_upload = time.perf_counter() - t0
    halo_wall, info = [], None
    for it in range(reps + 1):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        info = sr.exchange(0.1, 6.5)
        halo_wall.append(1e3 * (time.perf_counter() - t0))
    halo_wall = halo_wall[1:]  # the first call sizes the buffers
    ms_run = []
    for it in range(reps + 2):
        ctx.run(0.1, 6.5, 1)
        st = ctx.stats()
        ms_run.append(st["ms_run"])
    st1 = ctx.stats()
    rows = int(st1["n_rows"])
    ctx.run(0.1, 6.5, 0)  # the workload constant: in-cutoff pairs, every one counted by exactly one owner
    r0 = ctx.download()
    st0 = ctx.stats()
    assert len(r0["rows"]) == rows, (len(r0["rows"]), rows)
    out = {"rows": rows, "pairs_in_cutoff": int(r0["pairs_in_cutoff"]), "own_atoms": int(info["own_atoms"]), "ghost_atoms": int(info["ghost_atoms"]),
           "ghost_rings": int(info["ghost_rings"]), "halo_bytes": int(info["bytes_sent"]), "n_peers": int(info["n_peers"]),
           "ms_halo_wall": float(np.median(halo_wall)), "ms_halo_device": float(info["ms_total"]), "ms_halo_select": float(info["ms_select"]),
           "ms_halo_transfer": float(info["ms_transfer"]), "ms_halo_unpack": float(info["ms_unpack"]),
           "ms_device": float(np.median(ms_run[2:])), "ms_device_counting": float(st0["ms_run"]),
           "ms_grid": float(st1["ms_grid"]), "ms_pairs": float(st1["ms_pairs"]), "s_pack": s_pack, "s_upload": s_upload,
           "transport": int(info["transport"]), "local_atoms": int(st1["n_atoms"])}
    ctx.close()
    del own
    return out


def bench_batch(ctx_device: int, rank: int, world: int, n_structures: int, reps: int = 3):
    """BASELINE config 4: complex k of the batch goes to rank k mod world (distributed.run_batch_shard ->
    arp_contacts_batch); no collective on the data path."""
    import arpeggia_b200 as A
    from arpeggia_b200 import distributed as D, synth

    base = A.parse_pdb(ROOT / "tests" / "data" / "6bft.pdb")
    mine = D.shard_round_robin(n_structures, rank, world)
    t0 = time.perf_counter()
    arrs = [synth.jittered_complex(base, k) for k in mine]  # synthesising the workload (numpy), not part of the path
    s_generate = time.perf_counter() - t0
    t0 = time.perf_counter()
    packs = A.pack_many(arrs, "/", flags=2)  # arph_pack_many: one structure per host thread
    s_pack = time.perf_counter() - t0
    ctx = A.Context(ctx_device)
    t0 = time.perf_counter()
    res = ctx.contacts_rows(packs, 0.1, 6.5, 0)  # one-shot through the C ABI: upload, kernels, rows back
    s_call = time.perf_counter() - t0
    rows, pairs = len(res["rows"]), int(res["pairs_in_cutoff"])
    del res
    ctx.upload(packs)
    ms = []
    for it in range(reps + 2):
        ctx.run(0.1, 6.5, 1)
        ms.append(ctx.stats()["ms_run"])
    st = ctx.stats()
    assert int(st["n_rows"]) == rows
    out = {"structures": len(mine), "atoms": int(st["n_atoms"]), "rows": rows, "pairs_in_cutoff": pairs, "ms_device": float(np.median(ms[2:])),
           "ms_device_first": float(ms[0]), "retries": int(st["pair_kernel_retries"]), "cells": int(st["n_cells"]),
           "ms_grid": float(st["ms_grid"]), "ms_pairs": float(st["ms_pairs"]), "s_pack": s_pack, "s_generate": s_generate, "s_one_shot_call": s_call,
           "pack_threads": int(os.environ.get("OMP_NUM_THREADS", "0") or 0)}
    ctx.close()
    return out


def all_ranks_ok(err, world: int, device):
    """Every rank reports whether its leg succeeded before any rank reduces numbers (a failed rank must not leave the
    others inside a collective); returns the first error text, or None."""
    import torch
    import torch.distributed as dist

    if world == 1:
        return err
    box = [None] * world
    dist.all_gather_object(box, err)
    for r, e in enumerate(box):
        if e:
            return f"rank {r}: {e}"
    return None


def guarded_leg(fn, *a):
    try:
        return fn(*a), None
    except Exception as e:  # noqa: BLE001 -- reported in the bench line
        return None, f"{type(e).__name__}: {e}"[:400]


def reduce_records(rec: dict, world: int, device):
    """sum / max of every numeric field over the ranks (NCCL all-gather of one small tensor)."""
    import torch
    import torch.distributed as dist

    keys = sorted(rec)
    t = torch.tensor([float(rec[k]) for k in keys], dtype=torch.float64, device=device)
    if world > 1:
        parts = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(parts, t)
        m = torch.stack(parts).cpu().numpy()
    else:
        m = t.cpu().numpy()[None, :]
    return {k: m[:, i] for i, k in enumerate(keys)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dense", action="store_true", help="overlapping copies (protein-interior density) instead of config 3")
    ap.add_argument("--flags", type=int, default=1, help="arp_params.flags (1 = emit-radius screen, the drop-in's path; 0 = counting mode)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-counting", action="store_true", help="skip the counting-mode (flags 0) side measurement")
    ap.add_argument("--multi", default="auto", choices=["auto", "on", "off"],
                    help="configs 4 and 5 over the ranks (\"batch\" and \"slab\" records); auto = when world > 1")
    ap.add_argument("--slab-tiling", default="16,32,24", help="config 5 = 6bft tiled nx,ny,nz, cut into world x-slabs")
    ap.add_argument("--batch-structures", type=int, default=10000, help="config 4 = this many jittered complexes over the ranks")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist

    import arpeggia_b200 as A

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    # pinning the ranks to cores next to their GPU made no difference on these boxes (one NUMA node; profiles/r2b_e2e_phases_*): opt-in
    binding = bind_to_gpu_cores(local, world) if world > 1 and os.environ.get("ARP_BENCH_BIND") else None
    if world > 1:
        # NCCL announces its version on stdout when the communicator comes up; the contract is ONE json line there
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    atoms = build_workload(args.dense)
    packed = A.Packed(atoms, "/", flags=2)
    ctx = A.Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=f"cuda:{local}")  # > 126 MB of L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ---------------------------------------------------------------------
    ctx.upload(packed)
    for _ in range(args.warmup):
        ctx.run(0.1, 6.5, args.flags)
    res = ctx.download()
    st0 = ctx.stats()
    n_atoms, n_rows = st0["n_atoms"], st0["n_rows"]
    pairs = res["pairs_in_cutoff"] if res["pairs_in_cutoff"] >= 0 else None
    if pairs is None:  # emit-radius screen does not count them: take the workload constant from a plain run
        ctx.run(0.1, 6.5, 0)
        pairs = ctx.download()["pairs_in_cutoff"]
        ctx.run(0.1, 6.5, args.flags)
        ctx.download()
    sampler = ClockSampler(local)
    sampler.start()
    # let nvidia-smi come up before the timed region -- with the GPU kept busy by untimed steps, so that the timed steps see
    # the clocks of a loaded device (an idle 0.4 s lets them fall, and a 20-step run is over before they are back)
    t_busy = time.perf_counter()
    while time.perf_counter() - t_busy < 0.4:
        for _ in range(8):
            ctx.run(0.1, 6.5, args.flags)
        torch.cuda.synchronize()
    barrier()
    t_clock0 = time.perf_counter()
    t_step = t_pairs = t_grid = t_rings = t_screen = t_filter = t_exact = 0.0
    with torch.cuda.stream(stream):
        for _ in range(args.steps):
            flush.zero_()  # evict L2 between timed steps
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            ctx.run(0.1, 6.5, args.flags)
            e1.record(stream)
            e1.synchronize()
            t_step += e0.elapsed_time(e1)
            s = ctx.stats()
            t_pairs += s["ms_pairs"]
            t_grid += s["ms_grid"]
            t_rings += s["ms_rings"]
            t_screen += s["ms_screen"]
            t_filter += s["ms_filter"]
            t_exact += s["ms_exact"]
    barrier()
    clocks = sampler.summary(t_clock0, time.perf_counter())
    ms_step = t_step / args.steps
    times = torch.tensor([ms_step], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_max = float(times.item())
    value = world * pairs / (ms_max * 1e-3)

    # ---- the same loop in the counting mode (screen at dist_cutoff, in-cutoff pairs counted on the device) ----------------
    counting = None
    if args.flags != 0 and rank == 0 and world == 1 and not args.no_counting:
        for _ in range(3):
            ctx.run(0.1, 6.5, 0)
        k = max(20, min(args.steps, 100))
        tc = tp = 0.0
        with torch.cuda.stream(stream):
            for _ in range(k):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                ctx.run(0.1, 6.5, 0)
                e1.record(stream)
                e1.synchronize()
                tc += e0.elapsed_time(e1)
                tp += ctx.stats()["ms_pairs"]
        rc = ctx.download()
        sc = ctx.stats()
        counting = {"flags": 0, "steps": k, "ms_per_step": tc / k, "pair_stage_ms": tp / k, "value": rc["pairs_in_cutoff"] / (tc / k * 1e-3),
                    "pairs_within_cutoff_counted_on_device": int(rc["pairs_in_cutoff"]), "candidate_tests": int(rc["pairs_screened"]),
                    "cells": int(sc["n_cells"]), "rows": int(sc["n_rows"]),
                    "pair_stage": "k_screen -> k_filter -> k_exact"}
        ctx.run(0.1, 6.5, args.flags)
        ctx.download()

    # ---- end to end through the C ABI with pinned HOST buffers (H2D + kernels + D2H inside the timed region) ------
    hs, keep, h2d_bytes = pinned_structure(A, packed)
    from arpeggia_b200 import _lib as AL

    L = AL.lib()
    ptrs = (C.POINTER(AL.Structure) * 1)(C.pointer(hs))
    prm = AL.Params(0.1, 6.5, args.flags, 0)

    def one_call():
        rp = C.POINTER(AL.Result)()
        AL.check(L.arp_contacts_batch(ctx._h, ptrs, 1, C.byref(prm), C.byref(rp)), ctx._h)
        n = rp.contents.n_rows  # rows are in (pinned) host memory when the call returns
        L.arp_result_free(ctx._h, rp)
        return n

    for _ in range(2):
        n_rows_e2e = one_call()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(3, args.steps)
    for _ in range(e2e_steps):
        n_rows_e2e = one_call()
    torch.cuda.synchronize()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / e2e_steps
    e2e_t = torch.tensor([e2e_ms], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = world * pairs / (float(e2e_t.item()) * 1e-3)

    st = ctx.stats()  # of the last end-to-end call
    # the same call with every array in a page-locked allocation of its own: one copy per array (a side figure)
    e2e_sep = None
    if world == 1:
        hs2, keep2, _ = pinned_structure(A, packed, arena=False)
        ptrs_arena, ptrs = ptrs, (C.POINTER(AL.Structure) * 1)(C.pointer(hs2))
        for _ in range(2):
            one_call()
        n_sep = max(3, min(args.steps, 200))
        t0 = time.perf_counter()
        for _ in range(n_sep):
            one_call()
        st2 = ctx.stats()
        e2e_sep = {"ms_per_step": 1e3 * (time.perf_counter() - t0) / n_sep, "ms_h2d": st2["ms_h2d"], "upload_copies": st2["upload_copies"], "calls": n_sep}
        ptrs = ptrs_arena
    # every rank's phases of that call: host wall clock (arp_stats.ms_wall_*) beside the device times inside them
    e2e_keys = ("ms_wall_call", "ms_wall_upload", "ms_h2d", "ms_wall_run", "ms_run", "ms_wall_download", "ms_d2h")
    e2e_ranks = reduce_records({k: st[k] for k in e2e_keys}, world, f"cuda:{local}")

    # ---- the same call from several host threads, one context (= one stream) each: job i+1's upload, job i's kernels and
    # job i-1's row download run on the copy engines and the SMs at the same time (a side figure; `e2e` is the serial call)
    e2e_conc = None
    if rank == 0 and world == 1 and not args.no_counting:
        n_thr, jobs_each = 3, 12
        ctxs = [A.Context(local) for _ in range(n_thr)]
        rows_seen = [0] * n_thr

        def worker(k, n_jobs):
            for _ in range(n_jobs):
                rp = C.POINTER(AL.Result)()
                AL.check(L.arp_contacts_batch(ctxs[k]._h, ptrs, 1, C.byref(prm), C.byref(rp)), ctxs[k]._h)
                rows_seen[k] = rp.contents.n_rows
                L.arp_result_free(ctxs[k]._h, rp)

        for n_jobs in (2, jobs_each):  # first round: buffers sized, graphs captured
            th = [threading.Thread(target=worker, args=(k, n_jobs)) for k in range(n_thr)]
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for t in th:
                t.start()
            for t in th:
                t.join()
            torch.cuda.synchronize()
            conc_ms = 1e3 * (time.perf_counter() - t0) / (n_thr * n_jobs)
        e2e_conc = {"host_threads": n_thr, "contexts": n_thr, "jobs": n_thr * jobs_each, "ms_per_job": conc_ms,
                    "value": pairs / (conc_ms * 1e-3), "unit": UNIT, "rows_each": rows_seen,
                    "api": "arp_contacts from 3 host threads, one arp_ctx each (uploads, kernels and downloads of different jobs overlap)"}
        for cx in ctxs:
            cx.close()

    # ---- rows -> sorted 20-column frame on the device, beside the host comparator sort ---------
    frame_info = None
    if rank == 0 and world == 1:
        ctx.upload(packed)
        ctx.run(0.1, 6.5, args.flags)
        rows_host = ctx.download()["rows"]
        ctx.frame(packed)  # warm-up: builds the side tables, sizes the buffers, fills the pinned pool
        t0 = time.perf_counter()
        fr = ctx.frame(packed)
        t_dev_wall = 1e3 * (time.perf_counter() - t0)
        ctx.frame(packed, device_columns=True)
        fr_full = ctx.frame(packed, device_columns=True)  # every column decoded on the device: 86 bytes per row back
        walls = {}
        for mode in ("dictionary", "string"):  # the same frame straight into Arrow (no Python list): dictionary-encoded / String columns
            ctx.frame(packed, arrow=mode)
            t0 = time.perf_counter()
            at, _ = ctx.frame(packed, arrow=mode)
            walls[mode] = 1e3 * (time.perf_counter() - t0)
            assert at.num_rows == fr["n_rows"]
        del at
        t0 = time.perf_counter()
        host_tab = packed.build_table(rows_host)
        t_host = 1e3 * (time.perf_counter() - t0)
        same = all(np.array_equal(t[c], host_tab[c]) for t in (fr, fr_full)
                   for c in ("from_atomi", "to_atomi", "interaction_code", "from_resi", "to_resi", "from_chain", "to_atomn", "model"))
        frame_info = {"rows": int(fr["n_rows"]), "ms_sort": fr["frame_stats"]["ms_sort"], "ms_columns": fr["frame_stats"]["ms_columns"],
                      "ms_d2h": fr["frame_stats"]["ms_d2h"], "d2h_bytes_per_row": 26,
                      "ms_d2h_device_columns": fr_full["frame_stats"]["ms_d2h"], "ms_columns_device_columns": fr_full["frame_stats"]["ms_columns"],
                      "d2h_bytes_per_row_device_columns": 86, "sort_passes": fr["frame_stats"]["sort_passes"],
                      "key_bits": fr["frame_stats"]["key_bits"], "ms_wall_incl_numpy_copy": t_dev_wall,
                      "ms_wall_to_arrow_dictionary_columns": walls["dictionary"], "ms_wall_to_arrow_string_columns": walls["string"],
                      "host_table_ms_incl_numpy_copy": t_host, "equals_host_table": bool(same),
                      "ms_wall": walls["dictionary"],
                      "note": "ms_wall = ms_wall_to_arrow_dictionary_columns: run's rows on the device -> pyarrow.Table (sort + sc join on the device, "
                              "26 B/row D2H, 15 per-entity columns gathered on the host threads, zero-copy Arrow arrays); "
                              "ms_wall_incl_numpy_copy is the same frame copied into numpy arrays with Python string objects (the test helper)",
                      "api": "arp_stage_frame_compact + arph_table_from_frame (keys + radix sort + sc join on the device; *_device_columns: arp_stage_frame)"}
    d2h_bytes = int(n_rows_e2e * 16)

    # ---- the real drop-in call: file -> sorted frame (arph_contacts_file / arpeggia_b200.contacts), configs 2 and 3 ------------
    dropin = None
    if rank == 0 and world == 1 and not args.no_counting:
        import tempfile

        from arpeggia_b200 import synth

        dropin = {}
        with tempfile.TemporaryDirectory() as td:
            cif3 = os.path.join(td, "config3.cif")  # 6-character chain ids do not fit the PDB format's one-column chain field
            t0 = time.perf_counter()
            synth.write_cif(atoms, cif3)
            dropin["s_write_config3_cif_python"] = time.perf_counter() - t0
            dropin["config3_file_bytes"] = os.path.getsize(cif3)
            for name, path in (("config2_6bft_pdb", str(ROOT / "tests" / "data" / "6bft.pdb")), ("config3_1m_cif", cif3)):
                best = None
                for _ in range(3):
                    t0 = time.perf_counter()
                    at, info = ctx.contacts_file_arrow(path, flags=A.FLAG_NO_COUNTERS, strings="dictionary")
                    w_dict = time.perf_counter() - t0
                    rec = {"s_wall_file_to_arrow": w_dict, "rows": int(info["n_rows"]), "atoms": int(info["n_atoms"]),
                           "s_pack_and_side_tables": info["t_setup_s"], "s_device_upload_run_frame": info["t_classify_s"],
                           "s_table_wrap": info["t_table_s"]}
                    if best is None or w_dict < best["s_wall_file_to_arrow"]:
                        best = rec
                    del at
                t0 = time.perf_counter()
                df = A.contacts(path)
                best["s_wall_python_contacts_string_columns"] = time.perf_counter() - t0
                del df
                dropin[name] = best
        dropin["api"] = ("arph_contacts_file (reader + packer + device + frame, one C call) -> Arrow with dictionary-encoded string "
                         "columns; arpeggia_b200.contacts() = the same + Arrow String columns (the reference's dtype)")

    # ---- the SAP radius query (src/sap.rs:181-215) on the same assembly, beside the oracle ---------
    sap_info = None
    if rank == 0 and world == 1 and not args.no_counting:
        from arpeggia_b200 import sap
        from oracle import oracle as O

        arr = packed.arrays()
        sel = sap.sidechain_mask(atoms) if len(atoms) == arr["n_atoms"] else np.ones(arr["n_atoms"], dtype=bool)
        wts = np.random.default_rng(3).integers(-4, 5, size=arr["n_atoms"]).astype(np.float32)  # exact sums: bit-exact check
        ctx.upload(packed)
        got = ctx.neighbor_sum(wts, sel, 5.0)
        t_ns = [0.0, 0.0]
        for _ in range(5):
            flush.zero_()
            torch.cuda.synchronize()
            ctx.neighbor_sum(wts, sel, 5.0)
            sn = ctx.neighbor_sum_stats()
            t_ns[0] += sn["ms_grid"] / 5
            t_ns[1] += sn["ms_sum"] / 5
        sap_info = {"atoms": int(arr["n_atoms"]), "selected": int(sel.sum()), "radius": 5.0, "ms_grid": t_ns[0], "ms_sum": t_ns[1],
                    "atoms_per_s": arr["n_atoms"] / ((t_ns[0] + t_ns[1]) * 1e-3), "cells": sn["n_cells"],
                    "api": "arp_stage_neighbor_sum (device times: cell-list build, sum kernel; weights / mask H2D and sums D2H excluded)"}
        if not args.no_cpu:
            t0 = time.perf_counter()
            want = O.neighbor_sum(arr["x"], arr["y"], arr["z"], wts, sel, 5.0)
            sap_info["cpu_port_s"] = time.perf_counter() - t0
            sap_info["cpu_cores"] = os.cpu_count() or 1
            sap_info["equals_oracle"] = bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))

    # ---- BASELINE configs 4 and 5 over the ranks (every rank takes part; rank 0 reports) --------------------------------
    slab_rec = batch_rec = None
    if args.multi == "on" or (args.multi == "auto" and world > 1):
        dev = f"cuda:{local}"
        ctx.close()  # the config-3 buffers are not needed any more
        del flush
        torch.cuda.empty_cache()
        tiling = tuple(int(v) for v in args.slab_tiling.split(","))
        try:
            t0 = time.perf_counter()
            mine, err = guarded_leg(bench_slab, local, rank, world, tiling)
            err = all_ranks_ok(err, world, dev)
            if err:
                raise RuntimeError(err)
            m = reduce_records(mine, world, dev)
            exp = SLAB_EXPECT.get(tiling)
            slab_rec = {
                "workload": f"6bft tiled {tiling[0]}x{tiling[1]}x{tiling[2]} (BASELINE.json configs[4]) cut into {world} x-slabs; every rank tiles and packs only its own slab",
                "atoms": int(m["own_atoms"].sum()), "rows_union": int(m["rows"].sum()), "pairs_in_cutoff_sum": int(m["pairs_in_cutoff"].sum()),
                "ms_device_max": float(m["ms_device"].max()), "ms_device_counting_max": float(m["ms_device_counting"].max()),
                "ms_device_each": [round(float(v), 4) for v in m["ms_device"]],
                "ms_halo": float(m["ms_halo_wall"].max()), "ms_halo_device_max": float(m["ms_halo_device"].max()),
                "ms_halo_select_max": float(m["ms_halo_select"].max()), "ms_halo_transfer_max": float(m["ms_halo_transfer"].max()),
                "ms_halo_unpack_max": float(m["ms_halo_unpack"].max()),
                "halo_bytes": [int(v) for v in m["halo_bytes"]], "ghost_atoms": [int(v) for v in m["ghost_atoms"]],
                "ghost_rings": [int(v) for v in m["ghost_rings"]], "own_atoms": [int(v) for v in m["own_atoms"]],
                "n_peers": [int(v) for v in m["n_peers"]], "transport": {0: "none", 1: "nccl send/recv", 2: "peer copies"}[int(m["transport"].max())],
                "ms_grid_max": float(m["ms_grid"].max()), "ms_pairs_max": float(m["ms_pairs"].max()),
                "s_tile_and_pack_max": float(m["s_pack"].max()), "s_upload_max": float(m["s_upload"].max()),
                "pairs_per_s": float(m["pairs_in_cutoff"].sum() / (m["ms_device"].max() * 1e-3)),
                "pairs_per_s_incl_halo": float(m["pairs_in_cutoff"].sum() / ((m["ms_device"].max() + m["ms_halo_wall"].max()) * 1e-3)),
                "api": "arp_stage_upload (owned slab) -> arp_slab_exchange (CUDA select/pack/unpack + ncclSend/ncclRecv) -> arp_stage_run (flags 1)",
                "s_total": time.perf_counter() - t0,
            }
            if exp:
                slab_rec["expected"] = {"rows": exp["rows"], "pairs_in_cutoff": exp["pairs_in_cutoff"], "source": "profiles/r1t_big100m.json (one GPU)"}
                slab_rec["verified"] = bool(slab_rec["rows_union"] == exp["rows"] and slab_rec["pairs_in_cutoff_sum"] == exp["pairs_in_cutoff"]
                                            and slab_rec["atoms"] == exp["atoms"])
                slab_rec["speedup_vs_one_gpu"] = exp["ms_one_gpu"] / slab_rec["ms_device_max"]
                slab_rec["one_gpu_ms"] = exp["ms_one_gpu"]
                assert slab_rec["verified"], ("slab totals differ from the single-GPU run", slab_rec["rows_union"], slab_rec["pairs_in_cutoff_sum"])
        except Exception as e:  # keep the headline line even if this leg fails; the failure is in the record
            slab_rec = {**(slab_rec or {}), "error": f"{type(e).__name__}: {e}"[:400]}
            if world > 1:
                print(f"[bench] slab leg failed on rank {rank}: {slab_rec['error']}", file=sys.stderr)
        torch.cuda.empty_cache()
        try:
            t0 = time.perf_counter()
            mine, err = guarded_leg(bench_batch, local, rank, world, args.batch_structures)
            err = all_ranks_ok(err, world, dev)
            if err:
                raise RuntimeError(err)
            m = reduce_records(mine, world, dev)
            exp = BATCH_EXPECT.get(args.batch_structures)
            batch_rec = {
                "workload": f"{args.batch_structures} jittered ~5k-atom complexes (BASELINE.json configs[3]), complex k on rank k mod {world}",
                "structures": int(m["structures"].sum()), "atoms": int(m["atoms"].sum()), "rows": int(m["rows"].sum()),
                "pairs_in_cutoff": int(m["pairs_in_cutoff"].sum()), "ms_device_max": float(m["ms_device"].max()),
                "ms_device_each": [round(float(v), 4) for v in m["ms_device"]], "ms_device_first_each": [round(float(v), 4) for v in m["ms_device_first"]],
                "ms_grid_each": [round(float(v), 4) for v in m["ms_grid"]], "ms_pairs_each": [round(float(v), 4) for v in m["ms_pairs"]],
                "cells_each": [int(v) for v in m["cells"]], "retries_each": [int(v) for v in m["retries"]],
                "structures_per_s": float(m["structures"].sum() / (m["ms_device"].max() * 1e-3)),
                "pairs_per_s": float(m["pairs_in_cutoff"].sum() / (m["ms_device"].max() * 1e-3)),
                "s_pack_max": float(m["s_pack"].max()), "pack_threads_per_rank": int(m["pack_threads"].max()),
                "s_generate_synthetic_input_max": float(m["s_generate"].max()), "s_one_shot_call_max": float(m["s_one_shot_call"].max()),
                "structures_per_s_one_shot_call": float(m["structures"].sum() / m["s_one_shot_call"].max()),
                "api": "arph_pack_many -> arp_contacts_batch per rank, complex k on rank k mod world (no data-path collective); ms_device = arp_stage_run, flags 1",
                "s_total": time.perf_counter() - t0,
            }
            if exp:
                batch_rec["expected"] = {"rows": exp["rows"], "pairs_in_cutoff": exp["pairs_in_cutoff"], "source": "profiles/r1t_batch.json (one GPU)"}
                batch_rec["verified"] = bool(batch_rec["rows"] == exp["rows"] and batch_rec["pairs_in_cutoff"] == exp["pairs_in_cutoff"]
                                             and batch_rec["atoms"] == exp["atoms"])
                batch_rec["speedup_vs_one_gpu"] = exp["ms_one_gpu"] / batch_rec["ms_device_max"]
                assert batch_rec["verified"], ("batch totals differ from the single-GPU run", batch_rec["rows"], batch_rec["pairs_in_cutoff"])
        except Exception as e:
            batch_rec = {**(batch_rec or {}), "error": f"{type(e).__name__}: {e}"[:400]}
            if world > 1:
                print(f"[bench] batch leg failed on rank {rank}: {batch_rec['error']}", file=sys.stderr)

    if rank == 0:
        peaks = {}
        pk = ROOT / "MEASURED_PEAKS.json"
        if pk.exists():
            peaks = json.loads(pk.read_text())
        peak = float(peaks.get("hbm_gbs", 6650.0))
        ms_pairs = t_pairs / args.steps
        alg_bytes = 32 * n_atoms + 8 * st["n_cells"] + 16 * n_rows  # B_alg = 32 N + 8 C + 16 R
        achieved = alg_bytes / (ms_pairs * 1e-3) / 1e9
        # DRAM bytes per launch of the three pair-stage kernels, from the committed ncu --set full capture
        traffic, traffic_src = None, None
        tj = ROOT / "profiles" / "pair_stage_traffic.json"
        tight = bool(args.flags & 1)
        stage_kernels = ("k_screen_atoms", "k_exact") if tight else ("k_screen", "k_filter", "k_exact")
        if tj.exists() and not args.dense:
            t = json.loads(tj.read_text())
            ks = ({**t.get("counting_mode_kernels", {}), **t.get("kernels", {})} if tight else
                  {**t.get("kernels", {}), **t.get("counting_mode_kernels", {})})
            if all(k in ks for k in stage_kernels):
                traffic = int(sum(ks[k]["dram_read_bytes"] + ks[k]["dram_write_bytes"] for k in stage_kernels))
                traffic_src = t["source"] if tight else t.get("counting_mode_source", t["source"])
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "warmup_note": "plus untimed steps for 0.4 s while the clock sampler starts (the device stays loaded up to the timed region)",
            "ms_per_step": ms_max, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": WORKLOAD + (" [dense pitch]" if args.dense else ""), "per_gpu": "one assembly per rank (sharded by structure)",
                       "l2": "512 MiB buffer written between timed steps (L2 flushed)", "flags": args.flags,
                       "screen": ("emit radius (cells and screen at r_emit = 4.5 A, rows identical; pairs_within_cutoff is the workload "
                                  "constant, counted by a flags=0 run before the timed region: SURVEY 8(d))") if args.flags & 1 else
                                 "dist_cutoff (every in-cutoff pair counted on the device inside the timed region)",
                       "atoms": n_atoms, "cells": st["n_cells"], "nonempty_cells": st["n_nonempty_cells"], "rows": n_rows,
                       "pairs_within_cutoff": pairs, "candidate_tests": res["pairs_screened"],
                       "pair_queue_entries": st["n_screened_pairs"], "filtered_pairs": st["n_filtered_pairs"]},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "kernel": ("pair stage = k_screen_atoms -> k_exact (neighbour search + pair filter | classification, 2 launches per step)"
                                    if tight else
                                    "pair stage = k_screen -> k_filter -> k_exact (neighbour search + classification, 3 launches per step)"),
                         "kernel_ms": ms_pairs,
                         "kernel_ms_each": ({"k_screen_atoms": t_screen / args.steps, "k_exact": t_exact / args.steps} if tight else
                                            {"k_screen": t_screen / args.steps, "k_filter": t_filter / args.steps, "k_exact": t_exact / args.steps}),
                         "algorithmic_bytes": alg_bytes, "traffic_source": traffic_src,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "ms_per_step": float(e2e_t.item()), "ms_h2d": st["ms_h2d"], "ms_run": st["ms_run"], "ms_d2h": st["ms_d2h"],
                    "ms_host_other_rank0": e2e_ms - st["ms_h2d"] - st["ms_run"] - st["ms_d2h"], "calls": e2e_steps, "cpu_binding": binding,
                    "last_call_per_rank": {k: [round(float(v), 4) for v in e2e_ranks[k]] for k in e2e_keys},
                    "host_layout": "one page-locked arena, arrays 256-byte aligned one after the other", "upload_copies": st["upload_copies"],
                    "host_forms": ("xyz_milli21 (8 B/atom) + attr16 (2 B/atom)" if packed.structure.xyz_milli21 and packed.structure.attr16 else
                                   "x_milli, y_milli, z_milli (12 B/atom) + attr (4 B/atom)" if packed.structure.x_milli else "x, y, z f64 + attr")
                                  + " + the packer's per-residue arrays (chain key, ordinal, atom range, H range, CB, SG, tag)",
                    "separate_allocations": e2e_sep,
                    "api": "arp_contacts (C ABI, pinned host SoA in, host rows out)"},
            "e2e_concurrent": e2e_conc,
            "frame": frame_info,
            "counting_mode": counting,
            "dropin_e2e": dropin,
            "sap_neighbor_sum": sap_info,
            "slab": slab_rec,
            "batch": batch_rec,
            "gpu_launches": int(st0["kernel_launches"]) * args.steps,
            "clocks": clocks,
            "contacts_per_s": world * n_rows / (ms_max * 1e-3),
            "pairs_in_cutoff_per_s": value,  # what `value` is: the workload's in-cutoff pairs (a constant, see config.screen) / step time
            "candidate_tests_per_s": world * res["pairs_screened"] / (ms_max * 1e-3),
            "atoms_per_s": world * n_atoms / (ms_max * 1e-3),
            "stage_ms": {"grid": t_grid / args.steps, "pairs": ms_pairs, "rings_overlapped": t_rings / args.steps},
        }
        if not args.no_cpu and world == 1:
            threads = os.cpu_count() or 1
            c = cpu_reference(atoms, 125, threads)
            line["cpu_baseline"] = {"value": c["pairs"] / c["hot_s"], "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": c["sample"], "seconds": c["hot_s"]}
            # the reference's own shape (shape A): serial neighbour search + classification at the default -j 1
            # (cli/contacts.rs:46-47, complex.rs:197-213), on a bounded sample
            c1 = cpu_reference(atoms, 25, 1)
            line["cpu_baseline_ref_shaped"] = {"value": c1["pairs"] / c1["hot_s"], "unit": UNIT, "cores": 1, "kind": "port",
                                               "sample": c1["sample"], "seconds": c1["hot_s"],
                                               "shape": "serial neighbour search, 1 thread (the reference's default num_threads = 1)"}
            # the whole get_contacts on the CPU: hierarchy + index build, search, classification, ring loops, frame (join + sort)
            from oracle import oracle as O
            t0 = time.perf_counter()
            full = O.contacts(atoms.view(O.ATOM_DTYPE), "/", 0.1, 6.5, flags=O.FLAG_SKIP_LOAD_FILTER, num_threads=threads)
            line["cpu_full"] = {"s_wall": time.perf_counter() - t0, "cores": threads, "kind": "port", "rows": int(full["n_rows"]),
                                "s_setup": full["t_setup_s"], "s_search": full["t_search_s"], "s_classify": full["t_classify_s"],
                                "s_rings": full["t_rings_s"], "s_table": full["t_table_s"],
                                "what": "oracle get_contacts on atom records of config 3 (no file parsing), all host threads"}
            del full
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
