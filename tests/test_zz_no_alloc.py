"""arp_neighbor_sum_static -- the neighbour sum for callers that cannot allocate (include/arpeggia_cuda.h; no counterpart
in the reference) -- through the C ABI, against the existing entry point
arp_stage_neighbor_sum.  Both launch the same kernels; the static one keeps all scratch in a file-scope __device__ array.
Integer weights give exactly representable sums (every order of summation gives the same f32: bit-exact check, like
tests/test_gpu_sap.py); real weights agree within that file's bound n * 2^-24 * sum|w|.  Two calls in a row share the
array: the second must not see the first one's cell list.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import arpeggia_b200 as A
from arpeggia_b200 import _lib, sap

pytestmark = pytest.mark.gpu


def static_sum(dev, a, w, sel, radius, stream=None):
    """arp_neighbor_sum_static on device copies of the arrays (the TEST allocates them with torch; the call must not)."""
    import torch

    t = lambda v, dt: torch.from_numpy(np.ascontiguousarray(v, dtype=dt)).to(f"cuda:{dev}")
    x, y, z = (t(a[k], np.float64) for k in ("x", "y", "z"))
    wt, st = t(w, np.float32), t(np.asarray(sel, dtype=np.uint8), np.uint8)
    out = torch.full((len(w),), 123.0, dtype=torch.float32, device=f"cuda:{dev}")  # stale values must be overwritten
    torch.cuda.synchronize(dev)
    before = torch.cuda.memory_allocated(dev)
    free0, _ = torch.cuda.mem_get_info(dev)
    rc = _lib.lib().arp_neighbor_sum_static(dev, stream, x.data_ptr(), y.data_ptr(), z.data_ptr(), wt.data_ptr(), st.data_ptr(), len(w),
                                            C.c_float(radius), out.data_ptr())
    _lib.check(rc)
    free1, _ = torch.cuda.mem_get_info(dev)
    assert torch.cuda.memory_allocated(dev) == before
    return out.cpu().numpy(), free0 - free1


def test_static_neighbor_sum_equals_the_allocating_entry_point(ubq_path, bft_path):
    ctx = A.Context(0)
    rng = np.random.default_rng(23)
    first = True
    # two calls in a row on different structures and radii (the larger one first: its cell list must not leak into the
    # second call), then the first again
    for path, radius in ((bft_path, 5.0), (ubq_path, 3.3), (bft_path, 9.75)):
        atoms = sap.prepare_atoms(A.parse_pdb(path))
        packed = A.Packed(atoms, "/", flags=2)
        a = packed.arrays()
        sel = sap.sidechain_mask(atoms)
        w = rng.integers(-8, 9, size=len(atoms)).astype(np.float32)
        ctx.upload(packed)
        want = ctx.neighbor_sum(w, sel, radius)
        got, grew = static_sum(0, a, w, sel, radius)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        assert (got[~sel] == 0).all() and np.abs(got[sel]).max() > 0
        if not first:  # the very first call may fault in the module's lazily loaded code; later calls take no device memory
            assert grew < (1 << 20), f"free device memory shrank by {grew} bytes across the call (cudaMalloc granularity is 2 MiB)"
        first = False
        # real weights: the same neighbour set, f64 sums in a different order -> within the summation bound
        wr = rng.normal(size=len(atoms)).astype(np.float32)
        want_r = ctx.neighbor_sum(wr, sel, radius)
        got_r, _ = static_sum(0, a, wr, sel, radius)
        bound = ctx.neighbor_sum(np.abs(wr), sel, radius)
        count = ctx.neighbor_sum(np.ones(len(atoms), dtype=np.float32), sel, radius)
        assert np.all(np.abs(got_r - want_r) <= (count + 1) * 2.0 ** -24 * bound + 1e-12)
    ctx.close()


def test_static_neighbor_sum_on_a_caller_stream_and_without_a_context(bft_path):
    """No arp_ctx anywhere: the entry point stands on its own; a caller-owned stream is honoured."""
    import torch

    from oracle import oracle as O

    atoms = sap.prepare_atoms(A.parse_pdb(bft_path))
    a = A.Packed(atoms, "/", flags=2).arrays()
    sel = sap.sidechain_mask(atoms)
    w = np.random.default_rng(4).integers(-3, 4, size=len(atoms)).astype(np.float32)
    s = torch.cuda.Stream(device=0)
    got, _ = static_sum(0, a, w, sel, 5.0, stream=s.cuda_stream)
    want = O.neighbor_sum(a["x"], a["y"], a["z"], w, sel, 5.0)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_static_neighbor_sum_rejects_what_it_cannot_hold():
    L = _lib.lib()
    n = _lib.STATIC_MAX_ATOMS + 1
    rc = L.arp_neighbor_sum_static(0, None, 1, 1, 1, 1, 1, n, C.c_float(5.0), 1)  # the size check comes before any access
    assert rc == 11  # ARP_ERR_LIMIT
    msg = L.arp_last_error(None).decode()
    assert "ARP_STATIC_MAX_ATOMS" in msg and str(n) in msg
    assert L.arp_neighbor_sum_static(0, None, None, None, None, None, None, 0, C.c_float(5.0), None) == 0  # nothing to do
    assert L.arp_neighbor_sum_static(0, None, 1, 1, 1, 1, 1, 8, C.c_float(-1.0), 1) == 8  # ARP_ERR_ARG
    assert L.arp_neighbor_sum_static(99, None, 1, 1, 1, 1, 1, 8, C.c_float(5.0), 1) == 8
