// k_pairs: neighbour search + classification + compaction for atom-atom contacts (sm_100a).
// Replaces get_atomic_contacts, src/contacts/complex.rs:189-299.
//
// Work decomposition: one warp per non-empty cell ("from" cell), persistent CTAs pulling cells from an atomic
// cursor.  A warp visits the half shell of its cell -- the cell itself plus the 13 forward neighbours, which
// in the x-fastest cell order are 5 contiguous runs of sorted atoms -- so every unordered atom pair is screened
// once; both orientations are then filtered and classified.  Lanes own CANDIDATE atoms (coalesced 16-byte
// RecA loads); the from-cell's atoms sit in shared memory and are broadcast to all lanes.  The screen is an f32
// distance test with a rigorous guard band; each lane collects its hits in a 32-bit mask, the masks are
// compacted by a warp prefix sum into a per-warp shared-memory queue, and full batches of 32 queued pairs are
// drained through the exact FP64 test + filter + classification with all lanes busy.  Rows leave through one
// atomicAdd per drained batch.
#pragma once
#include "classify.cuh"

namespace arpd {

constexpr int PAIR_WARPS = 8;
constexpr int PAIR_THREADS = PAIR_WARPS * 32;
constexpr int FROM_CHUNK = 32;                       // from-atoms per pass (bits of the hit mask)
constexpr int QUEUE_CAP = 32 + 32 * FROM_CHUNK;      // carry-over + worst case of one candidate pass

struct PairSmem {
  RecA from[PAIR_WARPS][FROM_CHUNK];
  uint2 queue[PAIR_WARPS][QUEUE_CAP];
};

// Exact stage for up to 32 queued pairs (one per lane): FP64 distance, both orientations, labels, rows.
__device__ __forceinline__ void drain_batch(const PairArgs& G, const DevParams& P, const uint2* q, int count, int lane,
                                            unsigned long long& n_in_cutoff, unsigned long long& n_classified) {
  uint32_t m_ab = 0, m_ba = 0, sa = 0, sb = 0;
  double dist = 0.0;
  if (lane < count) {
    const uint2 e = q[lane];
    sa = e.x;
    sb = e.y;
    const RecB a = G.recB[sa], b = G.recB[sb];
    const uint32_t wa = G.recA[sa].attr, wb = G.recA[sb].attr;
    // rstar distance_2 / pdbtbx Atom::distance: ((dx*dx + dy*dy) + dz*dz), every op rounded
    const double dx = dsub(b.x, a.x), dy = dsub(b.y, a.y), dz = dsub(b.z, a.z);
    const double d2 = dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz));
    if (d2 <= P.rc2) {  // locate_within_distance is inclusive
      if (P.count_in_cutoff) {
        n_in_cutoff += ((wa & ARP_ATTR_LIGAND) && (wa & ARP_ATTR_ELEM_MASK) && !(wa & ARP_ATTR_GHOST)) ? 1u : 0u;
        n_in_cutoff += ((wb & ARP_ATTR_LIGAND) && (wb & ARP_ATTR_ELEM_MASK) && !(wb & ARP_ATTR_GHOST)) ? 1u : 0u;
      }
      const bool f_ab = pair_filter(wa, a.chain_key, a.res_ordinal, wb, b.chain_key, b.res_ordinal);
      const bool f_ba = pair_filter(wb, b.chain_key, b.res_ordinal, wa, a.chain_key, a.res_ordinal);
      n_classified += (f_ab ? 1u : 0u) + (f_ba ? 1u : 0u);
      if ((f_ab || f_ba) && d2 <= P.emit_d2_max) {
        dist = __dsqrt_rn(d2);
        if (f_ab) m_ab = classify_pair(G, P, a, wa, sa, b, wb, sb, dist);
        if (f_ba) m_ba = classify_pair(G, P, b, wb, sb, a, wa, sa, dist);
      }
    }
  }
  const uint32_t mine = __popc(m_ab) + __popc(m_ba);
  if (!__any_sync(0xffffffffu, mine != 0u)) return;
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
  unsigned long long base = 0;
  if (lane == 31) base = atomicAdd(&G.ctr->rows, (unsigned long long)total);
  base = __shfl_sync(0xffffffffu, base, 31);
  unsigned long long w = base + (inc - mine);
  if (mine) {
    const uint32_t ia = G.perm[sa], ib = G.perm[sb];
    const float fd = __double2float_rn(dist);  // Rust `as f32`, mod.rs:148
    while (m_ab) {
      const uint32_t l = __ffs(m_ab) - 1;
      m_ab &= m_ab - 1;
      if (w < G.row_cap) G.rows[w] = arp_row{ia, ib, l, fd};
      ++w;
    }
    while (m_ba) {
      const uint32_t l = __ffs(m_ba) - 1;
      m_ba &= m_ba - 1;
      if (w < G.row_cap) G.rows[w] = arp_row{ib, ia, l, fd};
      ++w;
    }
  }
}

__global__ void __launch_bounds__(PAIR_THREADS) k_pairs(const PairArgs G, const DevParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PairSmem& S = *reinterpret_cast<PairSmem*>(smem_raw);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  RecA* sfrom = S.from[wid];
  uint2* queue = S.queue[wid];
  int qlen = 0;
  unsigned long long n_screened = 0, n_in = 0, n_cls = 0;
  const uint32_t n_items = G.ctr->n_items;

  while (true) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(&G.ctr->work, 1u);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= n_items) break;
    const uint2 item = G.items[w];
    const StructDesc& D = G.descs[item.y];
    const int nx = D.nx, ny = D.ny, nz = D.nz;
    const uint32_t local = item.x - D.cell_base;
    const int ix = (int)(local % (uint32_t)nx), iy = (int)((local / (uint32_t)nx) % (uint32_t)ny), iz = (int)(local / ((uint32_t)nx * (uint32_t)ny));
    const uint32_t cb = G.cell_start[item.x], ce = G.cell_start[item.x + 1];

    // the 5 contiguous runs of the half shell: (dy, dz, x-from, x-to)
    uint32_t run_b[5], run_e[5];
#pragma unroll
    for (int r = 0; r < 5; ++r) {
      const int dy = (r == 0) ? 0 : (r == 1) ? 1 : (r - 3);  // r: 0 -> (0,0)  1 -> (+1,0)  2,3,4 -> (-1,0,+1 ; +1)
      const int dz = (r < 2) ? 0 : 1;
      const int y = iy + dy, z = iz + dz;
      run_b[r] = run_e[r] = 0;
      if (y < 0 || y >= ny || z >= nz) continue;
      const int x0 = (r == 0) ? ix : max(ix - 1, 0), x1 = min(ix + 1, nx - 1);
      const uint32_t row = D.cell_base + (uint32_t)((z * ny + y) * nx);
      run_b[r] = G.cell_start[row + x0];
      run_e[r] = G.cell_start[row + x1 + 1];
    }

    for (uint32_t fb = cb; fb < ce; fb += FROM_CHUNK) {
      const int nf = (int)min((uint32_t)FROM_CHUNK, ce - fb);
      const uint32_t full = nf == 32 ? 0xffffffffu : ((1u << nf) - 1u);
      __syncwarp();
      if (lane < nf) sfrom[lane] = G.recA[fb + lane];
      __syncwarp();
#pragma unroll 1
      for (int r = 0; r < 5; ++r) {
        // run 0 starts inside the from-cell: candidates of this chunk are the atoms after fb (pairs i < t once)
        const uint32_t sb = (r == 0) ? fb + 1 : run_b[r], se = run_e[r];
        for (uint32_t t0 = sb; t0 < se; t0 += 32) {
          const uint32_t t = t0 + lane;
          const bool valid = t < se;
          RecA c;
          if (valid) c = G.recA[t];
          else c.x = c.y = c.z = 3.0e18f, c.attr = 0;
          uint32_t mask = 0;
#pragma unroll 4
          for (int ii = 0; ii < nf; ++ii) {
            const RecA f = sfrom[ii];
            const float dx = c.x - f.x, dy = c.y - f.y, dz = c.z - f.z;
            const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
            mask |= (d2 <= P.rc2_screen ? 1u : 0u) << ii;
          }
          // inside the from-cell only from-atoms that precede the candidate count
          uint32_t allow = full;
          if (r == 0 && t < ce) allow = (t - fb < 32u) ? ((1u << (t - fb)) - 1u) & full : full;
          if (!valid) allow = 0;
          mask &= allow;
          n_screened += __popc(allow);
          if (__any_sync(0xffffffffu, mask != 0u)) {
            const uint32_t cnt = __popc(mask);
            uint32_t inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
              const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
              if (lane >= o) inc += v;
            }
            const int total = (int)__shfl_sync(0xffffffffu, inc, 31);
            int pos = qlen + (int)(inc - cnt);
            while (mask) {
              const uint32_t ii = __ffs(mask) - 1;
              mask &= mask - 1;
              queue[pos++] = make_uint2(fb + ii, t);
            }
            qlen += total;
            __syncwarp();
            while (qlen >= 32) {
              drain_batch(G, P, queue + (qlen - 32), 32, lane, n_in, n_cls);
              qlen -= 32;
              __syncwarp();
            }
          }
        }
      }
    }
  }
  if (qlen > 0) drain_batch(G, P, queue, qlen, lane, n_in, n_cls);
  // counters: warp reduce, one atomic per warp
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_screened += __shfl_xor_sync(0xffffffffu, n_screened, o);
    n_in += __shfl_xor_sync(0xffffffffu, n_in, o);
    n_cls += __shfl_xor_sync(0xffffffffu, n_cls, o);
  }
  if (lane == 0) {
    atomicAdd(&G.ctr->screened, n_screened);
    atomicAdd(&G.ctr->in_cutoff, n_in);
    atomicAdd(&G.ctr->classified, n_cls);
  }
}

}  // namespace arpd
