// Halo exchange of a slab-decomposed job (sm_100a): select / count / pack on the sender, unpack on the receiver.
// Partitions get_atomic_contacts / get_ring_atom_contacts / get_ring_ring_contacts (src/contacts/complex.rs:197-213,
// 306-321, 356-370) by the owner of the row's `from` entity; see arp_slab_exchange in include/arpeggia_cuda.h.
//
// A message (one per peer) is four sections of fixed-size records, each section 16-byte aligned:
//   HaloRes  x n_res    chain key, ordinal, tag, first atom / first hydrogen (message-relative), CB / SG (message-relative)
//   HaloAtom x n_atoms  f64 position, attribute word, global id            (the atoms of a residue are contiguous)
//   HaloH    x n_h      f64 position                                        (the hydrogens of a residue are contiguous)
//   HaloRing x n_rings  centre, normal, attribute word, chain key, ordinal, global id
// Offsets inside a message are deterministic (residue order = the sender's residue order): block sums of the per-peer
// (residue, atom, hydrogen) counts, one small scan over the blocks, a block-local scan again when packing.
#pragma once
#include "grid_kernels.cuh"

namespace arpd {

constexpr int HALO_MAX_PEERS = 8;    // peers within reach of one slab (a slab thinner than the halo width has more: refused)
constexpr int HALO_BLOCK = 256;      // residues per block of the count / pack kernels

struct __align__(16) HaloRes {
  uint32_t chain_key, res_ordinal, tag, atom_begin, h_begin;
  int32_t cb, sg;
  uint32_t n_atoms;
};
struct __align__(16) HaloAtom {
  double x, y, z;
  uint32_t attr, gid;
};
struct __align__(8) HaloH {
  double x, y, z;
};
struct __align__(16) HaloRing {
  double c[3], n[3];
  uint32_t attr, chain_key, res_ordinal, gid;
};

// per peer: residues, atoms, hydrogens, rings selected for it
struct HaloCounts {
  unsigned long long n[HALO_MAX_PEERS][4];
};

struct HaloSelect {
  int n_peers;
  double lo[HALO_MAX_PEERS], hi[HALO_MAX_PEERS];  // x-interval whose residues peer k needs (its owned range widened by the halo width)
};

// the owned part of the job as the halo kernels read it
struct SlabOwn {
  const double *x, *y, *z, *hx, *hy, *hz, *ring_c, *ring_n;
  const uint32_t *attr, *res_id, *res_atom_begin, *res_h_begin, *res_chain_key, *res_res_ordinal, *res_tag, *atom_gid;
  const int32_t *res_cb, *res_sg;
  const uint32_t *ring_attr, *ring_chain_key, *ring_res_ordinal, *ring_gid;
  uint32_t n_atoms, n_res, n_rings;
};

// x-range of the owned atoms, largest x-extent of a residue (both feed the peers' conservative bounding boxes)
__global__ void __launch_bounds__(256) k_slab_extent(SlabOwn S, unsigned long long* __restrict__ out /* [8]: min xyz, max xyz, max residue extent */) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  double lo[3] = {1e308, 1e308, 1e308}, hi[3] = {-1e308, -1e308, -1e308}, ext = 0.0;
  if (r < S.n_res) {
    const uint32_t b = S.res_atom_begin[r], e = S.res_atom_begin[r + 1];
    for (uint32_t a = b; a < e; ++a) {
      const double p[3] = {S.x[a], S.y[a], S.z[a]};
#pragma unroll
      for (int c = 0; c < 3; ++c) lo[c] = fmin(lo[c], p[c]), hi[c] = fmax(hi[c], p[c]);
    }
    if (e > b) ext = hi[0] - lo[0];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      lo[c] = fmin(lo[c], __shfl_xor_sync(0xffffffffu, lo[c], o));
      hi[c] = fmax(hi[c], __shfl_xor_sync(0xffffffffu, hi[c], o));
    }
    ext = fmax(ext, __shfl_xor_sync(0xffffffffu, ext, o));
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      if (lo[c] < 1e307) atomicMin(&out[c], enc_f64(lo[c]));
      if (hi[c] > -1e307) atomicMax(&out[3 + c], enc_f64(hi[c]));
    }
    atomicMax(&out[6], enc_f64(ext));
  }
}

// bit k of res_mask[r]: residue r goes to peer k (an atom of it lies in the peer's interval); the same for rings by centroid
__global__ void __launch_bounds__(256) k_halo_mark(SlabOwn S, HaloSelect Q, uint32_t* __restrict__ res_mask, uint32_t* __restrict__ ring_mask,
                                                   HaloCounts* __restrict__ counts) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < S.n_atoms) {
    const double x = S.x[i];
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < HALO_MAX_PEERS; ++k)
      if (k < Q.n_peers && x >= Q.lo[k] && x <= Q.hi[k]) m |= 1u << k;
    if (m) atomicOr(&res_mask[S.res_id[i]], m);
  }
  if (i < S.n_rings) {
    const double x = S.ring_c[3 * (size_t)i];
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < HALO_MAX_PEERS; ++k)
      if (k < Q.n_peers && x >= Q.lo[k] && x <= Q.hi[k]) {
        m |= 1u << k;
        atomicAdd(&counts->n[k][3], 1ull);
      }
    ring_mask[i] = m;
  }
}

// block-wide exclusive scan of three counters at once (residues, atoms, hydrogens), HALO_BLOCK threads
struct Tri {
  uint32_t r, a, h;
};
__device__ __forceinline__ Tri tri_add(Tri p, Tri q) { return Tri{p.r + q.r, p.a + q.a, p.h + q.h}; }
__device__ __forceinline__ Tri block_scan_tri(Tri v, Tri* total, Tri* wsum /* [HALO_BLOCK / 32] shared */) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  Tri inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    Tri t;
    t.r = __shfl_up_sync(0xffffffffu, inc.r, o);
    t.a = __shfl_up_sync(0xffffffffu, inc.a, o);
    t.h = __shfl_up_sync(0xffffffffu, inc.h, o);
    if (lane >= o) inc = tri_add(inc, t);
  }
  __syncthreads();  // wsum may still be read by the previous call
  if (lane == 31) wsum[w] = inc;
  __syncthreads();
  Tri before{0u, 0u, 0u}, tot{0u, 0u, 0u};
#pragma unroll
  for (int k = 0; k < HALO_BLOCK / 32; ++k) {
    if (k < w) before = tri_add(before, wsum[k]);
    tot = tri_add(tot, wsum[k]);
  }
  *total = tot;
  return Tri{before.r + inc.r - v.r, before.a + inc.a - v.a, before.h + inc.h - v.h};
}

__device__ __forceinline__ Tri residue_tri(const SlabOwn& S, uint32_t r, bool sel) {
  if (!sel) return Tri{0u, 0u, 0u};
  return Tri{1u, S.res_atom_begin[r + 1] - S.res_atom_begin[r], S.res_h_begin[r + 1] - S.res_h_begin[r]};
}

// per (block of residues, peer): how many residues / atoms / hydrogens the block sends
__global__ void __launch_bounds__(HALO_BLOCK) k_halo_count(SlabOwn S, int n_peers, const uint32_t* __restrict__ res_mask, Tri* __restrict__ block_tri /* [n_peers][n_blocks] */) {
  __shared__ Tri wsum[HALO_BLOCK / 32];
  const uint32_t r = blockIdx.x * HALO_BLOCK + threadIdx.x;
  const uint32_t m = r < S.n_res ? res_mask[r] : 0u;
  for (int k = 0; k < n_peers; ++k) {
    Tri tot;
    block_scan_tri(residue_tri(S, r, (m >> k) & 1u), &tot, wsum);
    if (threadIdx.x == 0) block_tri[(size_t)k * gridDim.x + blockIdx.x] = tot;
  }
}

// exclusive scan over the blocks (one CTA per peer walks its row of block sums); totals -> counts
__global__ void __launch_bounds__(HALO_BLOCK) k_halo_offsets(Tri* __restrict__ block_tri, uint32_t n_blocks, HaloCounts* __restrict__ counts) {
  __shared__ Tri wsum[HALO_BLOCK / 32];
  Tri* row = block_tri + (size_t)blockIdx.x * n_blocks;
  Tri carry{0u, 0u, 0u};
  for (uint32_t b0 = 0; b0 < n_blocks; b0 += HALO_BLOCK) {
    const uint32_t b = b0 + threadIdx.x;
    const Tri v = b < n_blocks ? row[b] : Tri{0u, 0u, 0u};
    Tri tot;
    const Tri ex = block_scan_tri(v, &tot, wsum);
    if (b < n_blocks) row[b] = tri_add(carry, ex);
    carry = tri_add(carry, tot);
  }
  if (threadIdx.x == 0) {
    counts->n[blockIdx.x][0] = carry.r;
    counts->n[blockIdx.x][1] = carry.a;
    counts->n[blockIdx.x][2] = carry.h;
  }
}

// where the four sections of peer k's message start (device pointers into the send buffer of that peer)
struct HaloMsg {
  HaloRes* res[HALO_MAX_PEERS];
  HaloAtom* atoms[HALO_MAX_PEERS];
  HaloH* h[HALO_MAX_PEERS];
  HaloRing* rings[HALO_MAX_PEERS];
};

__global__ void __launch_bounds__(HALO_BLOCK) k_halo_pack(SlabOwn S, int n_peers, const uint32_t* __restrict__ res_mask, const Tri* __restrict__ block_tri,
                                                          HaloMsg M) {
  __shared__ Tri wsum[HALO_BLOCK / 32];
  const uint32_t r = blockIdx.x * HALO_BLOCK + threadIdx.x;
  const uint32_t m = r < S.n_res ? res_mask[r] : 0u;
  for (int k = 0; k < n_peers; ++k) {
    const bool sel = (m >> k) & 1u;
    Tri tot;
    const Tri ex = block_scan_tri(residue_tri(S, r, sel), &tot, wsum);
    if (!sel) continue;
    const Tri base = block_tri[(size_t)k * gridDim.x + blockIdx.x];
    const uint32_t j = base.r + ex.r, a0 = base.a + ex.a, h0 = base.h + ex.h;
    const uint32_t ab = S.res_atom_begin[r], ae = S.res_atom_begin[r + 1], hb = S.res_h_begin[r], he = S.res_h_begin[r + 1];
    HaloRes q;
    q.chain_key = S.res_chain_key[r];
    q.res_ordinal = S.res_res_ordinal[r];
    q.tag = S.res_tag[r];
    q.atom_begin = a0;
    q.h_begin = h0;
    const int32_t cb = S.res_cb[r], sg = S.res_sg[r];
    q.cb = cb < 0 ? -1 : (int32_t)(a0 + ((uint32_t)cb - ab));  // the first CB / SG of a residue is one of its own atoms
    q.sg = sg < 0 ? -1 : (int32_t)(a0 + ((uint32_t)sg - ab));
    q.n_atoms = ae - ab;
    M.res[k][j] = q;
    for (uint32_t a = ab; a < ae; ++a) M.atoms[k][a0 + (a - ab)] = HaloAtom{S.x[a], S.y[a], S.z[a], S.attr[a], S.atom_gid[a]};
    for (uint32_t h = hb; h < he; ++h) M.h[k][h0 + (h - hb)] = HaloH{S.hx[h], S.hy[h], S.hz[h]};
  }
}

// rings: slot order inside a message is arrival order (ring records are independent of each other)
__global__ void __launch_bounds__(256) k_halo_pack_rings(SlabOwn S, int n_peers, const uint32_t* __restrict__ ring_mask, unsigned int* __restrict__ cursor /* [peers] */,
                                                         HaloMsg M) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= S.n_rings) return;
  const uint32_t m = ring_mask[g];
  if (!m) return;
  HaloRing q;
#pragma unroll
  for (int c = 0; c < 3; ++c) q.c[c] = S.ring_c[3 * (size_t)g + c], q.n[c] = S.ring_n[3 * (size_t)g + c];
  q.attr = S.ring_attr[g];
  q.chain_key = S.ring_chain_key[g];
  q.res_ordinal = S.ring_res_ordinal[g];
  q.gid = S.ring_gid[g];
  for (int k = 0; k < n_peers; ++k)
    if ((m >> k) & 1u) M.rings[k][atomicAdd(&cursor[k], 1u)] = q;
}

// the resident job's arrays, written at and beyond the owned part
struct SlabJob {
  double *x, *y, *z, *hx, *hy, *hz, *ring_c, *ring_n;
  uint32_t *attr, *res_id, *res_atom_begin, *res_h_begin, *res_chain_key, *res_res_ordinal, *res_tag, *atom_gid;
  int32_t *res_cb, *res_sg;
  uint32_t *ring_attr, *ring_chain_key, *ring_res_ordinal, *ring_gid;
};

// one received message -> the job, appended at (res0, atom0, h0, ring0); everything it holds becomes a ghost
__global__ void __launch_bounds__(256) k_halo_unpack(SlabJob J, const HaloRes* __restrict__ res, const HaloAtom* __restrict__ atoms, const HaloH* __restrict__ hs,
                                                     const HaloRing* __restrict__ rings, uint32_t n_res, uint32_t n_atoms, uint32_t n_h, uint32_t n_rings,
                                                     uint32_t res0, uint32_t atom0, uint32_t h0, uint32_t ring0) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_res) {
    const HaloRes q = res[i];
    const uint32_t r = res0 + i;
    J.res_chain_key[r] = q.chain_key;
    J.res_res_ordinal[r] = q.res_ordinal;
    J.res_tag[r] = q.tag;
    J.res_atom_begin[r] = atom0 + q.atom_begin;
    J.res_h_begin[r] = h0 + q.h_begin;
    J.res_cb[r] = q.cb < 0 ? -1 : (int32_t)(atom0 + (uint32_t)q.cb);
    J.res_sg[r] = q.sg < 0 ? -1 : (int32_t)(atom0 + (uint32_t)q.sg);
    for (uint32_t a = 0; a < q.n_atoms; ++a) J.res_id[atom0 + q.atom_begin + a] = r;
  }
  if (i < n_atoms) {
    const HaloAtom q = atoms[i];
    const uint32_t a = atom0 + i;
    J.x[a] = q.x, J.y[a] = q.y, J.z[a] = q.z;
    J.attr[a] = q.attr | ARP_ATTR_GHOST;
    J.atom_gid[a] = q.gid;
  }
  if (i < n_h) {
    const HaloH q = hs[i];
    J.hx[h0 + i] = q.x, J.hy[h0 + i] = q.y, J.hz[h0 + i] = q.z;
  }
  if (i < n_rings) {
    const HaloRing q = rings[i];
    const uint32_t g = ring0 + i;
#pragma unroll
    for (int c = 0; c < 3; ++c) J.ring_c[3 * (size_t)g + c] = q.c[c], J.ring_n[3 * (size_t)g + c] = q.n[c];
    J.ring_attr[g] = q.attr | ARP_ATTR_GHOST;
    J.ring_chain_key[g] = q.chain_key;
    J.ring_res_ordinal[g] = q.res_ordinal;
    J.ring_gid[g] = q.gid;
  }
}

// closing entries of the two residue CSRs after the last message
__global__ void k_halo_close(uint32_t* res_atom_begin, uint32_t* res_h_begin, uint32_t n_res, uint32_t n_atoms, uint32_t n_h) {
  res_atom_begin[n_res] = n_atoms;
  res_h_begin[n_res] = n_h;
}

// rows of a job with global ids: job-local entity index -> the caller's id (arp_structure.atom_gid / ring_gid)
__global__ void __launch_bounds__(256) k_rows_to_gid(arp_row* __restrict__ rows, unsigned long long n_rows, const uint32_t* __restrict__ atom_gid,
                                                     const uint32_t* __restrict__ ring_gid) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  arp_row r = rows[i];
  r.from = (r.from & ARP_RING_FLAG) ? (ring_gid[r.from & ~ARP_RING_FLAG] | ARP_RING_FLAG) : atom_gid[r.from];
  r.to = (r.to & ARP_RING_FLAG) ? (ring_gid[r.to & ~ARP_RING_FLAG] | ARP_RING_FLAG) : atom_gid[r.to];
  rows[i] = r;
}

// residue tags from (chain key, ordinal) of neighbouring residues (arp_structure.res_tag == NULL; atoms_with_hierarchy()
// order): step[r] = 1 when residue r + 1 continues the chain of r with the next ordinal, else 2; tag = exclusive scan.
// A residue that follows one of the same chain with an ordinal that is not larger breaks the assumed order: error bit 8.
__global__ void __launch_bounds__(256) k_res_tag_steps(const uint32_t* __restrict__ res_chain_key, const uint32_t* __restrict__ res_res_ordinal,
                                                       const uint32_t* __restrict__ struct_res_begin /* [n_struct + 1] */, uint32_t n_struct, uint32_t n_res,
                                                       uint32_t* __restrict__ step, unsigned int* __restrict__ error) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n_res) return;
  uint32_t s = 2u;
  if (r + 1 < n_res && res_chain_key[r] == res_chain_key[r + 1]) {
    bool same_struct = true;
    if (n_struct > 1u) {  // does a structure start at residue r + 1?
      uint32_t lo = 0, hi = n_struct;  // largest k with struct_res_begin[k] <= r + 1
      while (hi - lo > 1u) {
        const uint32_t mid = (lo + hi) >> 1;
        if (struct_res_begin[mid] <= r + 1u) lo = mid; else hi = mid;
      }
      same_struct = struct_res_begin[lo] != r + 1u;
    }
    if (same_struct) {
      const uint32_t o0 = res_res_ordinal[r], o1 = res_res_ordinal[r + 1];
      if (o1 <= o0) atomicOr(error, 8u);
      if (o1 == o0 + 1u) s = 1u;
    }
  }
  step[r] = s;
}

// per-residue (chain key, ordinal) from the per-atom arrays: every atom of a residue carries the same pair (benign race)
__global__ void __launch_bounds__(256) k_res_keys_from_atoms(const uint32_t* __restrict__ chain_key, const uint32_t* __restrict__ res_ordinal,
                                                             const uint32_t* __restrict__ res_id, uint32_t n_atoms,
                                                             uint32_t* __restrict__ res_chain_key, uint32_t* __restrict__ res_res_ordinal) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n_atoms) return;
  const uint32_t r = res_id[a];
  res_chain_key[r] = chain_key[a];
  res_res_ordinal[r] = res_ordinal[a];
}

// one row of the count matrix: what this rank sends to every rank (4 counts each), from the per-peer counters
__global__ void k_halo_row(const HaloCounts* __restrict__ counts, const int* __restrict__ peer_rank, int n_peers, int world,
                           unsigned long long* __restrict__ row /* [world * 4] */) {
  for (int i = threadIdx.x; i < world * 4; i += blockDim.x) row[i] = 0ull;
  __syncthreads();
  if ((int)threadIdx.x < n_peers * 4) row[peer_rank[threadIdx.x >> 2] * 4 + (threadIdx.x & 3)] = counts->n[threadIdx.x >> 2][threadIdx.x & 3];
}

}  // namespace arpd
