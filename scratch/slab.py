"""x-slab decomposition of ONE assembly over the ranks of a torchrun job (SURVEY 8(e), config 5 style): halo exchange over
NCCL, unchanged kernels, union of the ranks' rows = the single-GPU row set.  usage: torchrun ... scratch/slab.py nx ny nz"""
import os, sys, time, json
import numpy as np
import torch
import torch.distributed as dist
import arpeggia_b200 as A
from arpeggia_b200 import synth, distributed as D

nx, ny, nz = (int(v) for v in sys.argv[1:4])
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
bft = A.parse_pdb("tests/data/6bft.pdb")
t0 = time.perf_counter()
arr = A.Packed(synth.tile(bft, nx, ny, nz), "/", flags=2).arrays()   # every rank packs the assembly, then keeps its slab
t_pack = time.perf_counter() - t0
whole = D.soa_from_arrays(arr, device=f"cuda:{local}")
own = D.owned_part(whole, rank, world)
n_total = int(whole["x"].numel())
del whole, arr
torch.cuda.empty_cache()
ctx = A.Context(local)
width = 6.5 * (1.0 + 1e-6) + 1e-6
times = []
for it in range(3):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    combined, info = D.exchange_halos(own, width)
    torch.cuda.synchronize()
    t_x = time.perf_counter() - t0
    bound = D.BoundStructure(combined)
    t1 = time.perf_counter()
    res = ctx.contacts_rows([bound.ptr], 0.1, 6.5, 0)
    t_c = time.perf_counter() - t1
    st = ctx.stats()
    times.append((t_x, t_c, st["ms_run"]))
own_rows = int(len(res["rows"]))
tot = torch.tensor([own_rows, res["pairs_in_cutoff"], own["x"].numel(), info["ghost_atoms"], info["bytes_sent"]], dtype=torch.int64, device=f"cuda:{local}")
per = [torch.zeros_like(tot) for _ in range(world)]
dist.all_gather(per, tot)
ms = torch.tensor([times[-1][0] * 1e3, times[-1][1] * 1e3, times[-1][2]], dtype=torch.float64, device=f"cuda:{local}")
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    per = torch.stack(per).cpu().numpy()
    print(json.dumps({"world": world, "atoms": n_total, "rows_union": int(per[:, 0].sum()), "pairs_in_cutoff_sum": int(per[:, 1].sum()),
                      "owned_atoms": per[:, 2].tolist(), "ghost_atoms": per[:, 3].tolist(), "halo_bytes_sent": per[:, 4].tolist(),
                      "ms_halo_exchange_max": float(ms[0]), "ms_contacts_call_max": float(ms[1]), "ms_device_run_max": float(ms[2]),
                      "s_pack": t_pack}))
dist.destroy_process_group()
