// Ring kernel (sm_100a): cation-pi and pi-pi contacts, both against the ring-centroid grid.
//   k_rings  <->  get_ring_atom_contacts + find_cation_pi   src/contacts/complex.rs:301-352, aromatic.rs:14-29
//                 get_ring_ring_contacts + find_pi_pi       src/contacts/complex.rs:354-405, aromatic.rs:33-64
// The reference's ring-ring loop is O(R^2) over all ring pairs; find_pi_pi rejects centroid distances above
// 6.0 A first, so a centroid grid with edge >= 6.0 visits exactly the pairs that can produce a row.
#pragma once
#include "classify.cuh"

namespace arpd {

__device__ __forceinline__ void push_row(arp_row* rows, unsigned long long cap, Counters* ctr, uint32_t from, uint32_t to, uint32_t label,
                                         double dist) {
  const unsigned long long w = atomicAdd(&ctr->rows, 1ull);
  if (w < cap) rows[w] = arp_row{from, to, label, __double2float_rn(dist)};
}

// The 27 cells around a point are 9 rows (y, z) of up to 3 x-adjacent cells = 9 contiguous runs of sorted slots.
// A query is handled by a HALF WARP (16 lanes, `hm` = its shuffle mask, `l` = lane within it): lane k < 9 fetches the
// bounds of run k (all nine in one memory round trip); the half warp then walks the concatenation of the runs, one entry
// per lane.  Returns this lane's run (begin, length) and the exclusive prefix / total over the runs.
struct Runs9 {
  uint32_t begin, len, off, total;
};
__device__ __forceinline__ Runs9 fetch_runs9(const uint32_t* __restrict__ cell_start, uint32_t cell_base, int nx, int ny, int nz, int ix,
                                             int iy, int iz, int l, uint32_t hm) {
  Runs9 r{0u, 0u, 0u, 0u};
  if (l < 9) {
    const int y = iy + l % 3 - 1, z = iz + l / 3 - 1;
    if (y >= 0 && y < ny && z >= 0 && z < nz) {
      const uint32_t row = cell_base + (uint32_t)((z * ny + y) * nx);
      r.begin = cell_start[row + (uint32_t)max(ix - 1, 0)];
      r.len = cell_start[row + (uint32_t)min(ix + 1, nx - 1) + 1u] - r.begin;
    }
  }
  uint32_t inc = r.len;
#pragma unroll
  for (int o = 1; o < 16; o <<= 1) {
    const uint32_t t = __shfl_up_sync(hm, inc, o, 16);
    if (l >= o) inc += t;
  }
  r.off = inc - r.len;
  r.total = __shfl_sync(hm, inc, 8, 16);
  return r;
}
// entry j of the concatenation -> sorted slot (every lane asks for its own j)
__device__ __forceinline__ uint32_t runs9_slot(const Runs9& r, uint32_t j, uint32_t hm) {
  uint32_t slot = 0;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const uint32_t off = __shfl_sync(hm, r.off, k, 16), len = __shfl_sync(hm, r.len, k, 16), begin = __shfl_sync(hm, r.begin, k, 16);
    if (j >= off && j < off + len) slot = begin + (j - off);
  }
  return slot;
}

// k_rings: one half warp per QUERY -- a ring (pi-pi, complex.rs:354-405) or a cation (cation-pi, complex.rs:301-352) -- against
// the rings of the 27 ring-grid cells around it.  The reference searches the atoms around every ring centroid; only
// atoms that are positively ionizable by residue name can give a row (aromatic.rs:18) and a row needs
// dist <= 4.5 < ring-grid edge, so walking the few cations and looking for rings around them visits exactly the pairs
// that can produce a row -- about 20x fewer tests than sweeping all atoms around all rings.
#ifndef RINGS_MIN_BLOCKS
#define RINGS_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__(256, RINGS_MIN_BLOCKS) k_rings(const RingArgs R, const DevParams P) {
  const int lane = threadIdx.x & 15;
  const uint32_t hm = (threadIdx.x & 16) ? 0xffff0000u : 0x0000ffffu;  // the half warps of a warp run independently
  const uint32_t groups = (gridDim.x * blockDim.x) >> 4;
  const uint32_t n_queries = R.n_rings + R.ctr->n_cations;
  for (uint32_t q = (blockIdx.x * blockDim.x + threadIdx.x) >> 4; q < n_queries; q += groups) {
    const bool is_ring = q < R.n_rings;
    // query position, attributes, structure
    double c1[3], n1[3] = {0.0, 0.0, 0.0};
    uint32_t w1, k1, o1, sid, id1;
    if (is_ring) {
      id1 = q;
      w1 = R.attr[q];
      if (!(w1 & ARP_ATTR_LIGAND) || (w1 & ARP_ATTR_GHOST)) continue;  // complex.rs:363
      c1[0] = R.center[3 * q], c1[1] = R.center[3 * q + 1], c1[2] = R.center[3 * q + 2];
      n1[0] = R.normal[3 * q], n1[1] = R.normal[3 * q + 1], n1[2] = R.normal[3 * q + 2];
      k1 = R.chain_key[q], o1 = R.res_ordinal[q];
      sid = R.ring_struct ? R.ring_struct[q] : 0u;
    } else {
      const uint32_t t = R.cations[q - R.n_rings];
      const RecK a = R.recK[t];
      const RecB b = R.recB[t];
      id1 = a.perm;
      w1 = a.attr, k1 = a.chain_key, o1 = a.res_ordinal;
      c1[0] = b.x, c1[1] = b.y, c1[2] = b.z;
      sid = R.atom_struct ? R.atom_struct[a.perm] : 0u;
    }
    const StructDesc& D = R.descs[sid];
    const int ix = min(max((int)floor((c1[0] - D.ox) * P.inv_redge), 0), D.rnx - 1);
    const int iy = min(max((int)floor((c1[1] - D.oy) * P.inv_redge), 0), D.rny - 1);
    const int iz = min(max((int)floor((c1[2] - D.oz) * P.inv_redge), 0), D.rnz - 1);
    const Runs9 runs = fetch_runs9(R.rcell_start, D.rcell_base, D.rnx, D.rny, D.rnz, ix, iy, iz, lane, hm);
    for (uint32_t j0 = 0; j0 < runs.total; j0 += 16) {
      const uint32_t j = j0 + lane;
      const uint32_t t = runs9_slot(runs, j, hm);
      if (j >= runs.total) continue;
      const uint32_t r2 = R.rperm[t];
      const uint32_t w2 = R.attr[r2];
      const uint32_t k2 = R.chain_key[r2], o2 = R.res_ordinal[r2];
      const double c2[3] = {R.center[3 * r2], R.center[3 * r2 + 1], R.center[3 * r2 + 2]};
      const double n2[3] = {R.normal[3 * r2], R.normal[3 * r2 + 1], R.normal[3 * r2 + 2]};
      if (is_ring) {
        if (!(w2 & ARP_ATTR_RECEPTOR)) continue;  // complex.rs:364
        // should_compare_residues(k1, k2, true), complex.rs:94-131
        if ((k1 ^ k2) >> 24) continue;
        if (k1 == k2) {
          if (!((o2 > 1u) && (o1 < o2 - 1u))) continue;
        } else {
          const bool both = (w1 & w2 & ARP_ATTR_LIGAND) && (w1 & w2 & ARP_ATTR_RECEPTOR);
          if (both && k1 > k2) continue;
        }
        double dist;
        const int label = find_pi_pi(P, c1, n1, c2, n2, &dist);
        if (label >= 0) push_row(R.rows, R.row_cap, R.ctr, id1 | ARP_RING_FLAG, r2 | ARP_RING_FLAG, (uint32_t)label, dist);
      } else {
        // ring r2 (the `from` side, complex.rs:334-343) against cation atom id1
        const double vx = dsub(c1[0], c2[0]), vy = dsub(c1[1], c2[1]), vz = dsub(c1[2], c2[2]);
        const double d2 = dot3(vx, vy, vz, vx, vy, vz);
        if (!(d2 <= P.rc2)) continue;  // tree.locate_within_distance(center, rc^2), complex.rs:310-313
        if (!ring_atom_filter(w2, k2, o2, w1, k1, o1)) continue;
        const double dist = __dsqrt_rn(d2);  // ring.point_dist
        if (!(dist <= 4.5)) continue;
        const double c = __ddiv_rn(dot3(n2[0], n2[1], n2[2], vx, vy, vz), dmul(norm3(n2[0], n2[1], n2[2]), dist));
        if (!fold_pred(c, P.ang.f_le_30_pos, P.ang.f_le_30_neg)) continue;  // theta <= 30
        push_row(R.rows, R.row_cap, R.ctr, r2 | ARP_RING_FLAG, id1, ARP_CationPi, dist);
      }
    }
  }
}

}  // namespace arpd
