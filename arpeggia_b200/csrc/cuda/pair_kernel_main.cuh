// k_pairs main loop (stage 1 of the funnel described in pair_kernel.cuh).
#pragma once

namespace arpd {

__global__ void __launch_bounds__(PAIR_THREADS, 3) k_pairs(const PairArgs G, const DevParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PairSmem& S = *reinterpret_cast<PairSmem*>(smem_raw);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  RecA* sfrom = S.from[wid];
  uint2* q1 = S.q1[wid];
  uint2* q2 = S.q2[wid];
  int q1len = 0, q2len = 0;
  unsigned long long n_screened = 0, n_in = 0, n_cls = 0;
  const uint32_t n_items = G.ctr->n_items;
  const float thr_hit = P.thr_hit, thr_cert = P.thr_cert, thr_near = P.thr_near;
  const bool counting = P.count_in_cutoff != 0;

  while (true) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(&G.ctr->work, 1u);
    w = __shfl_sync(0xffffffffu, w, 0);
    if (w >= n_items) break;
    const uint2 item = G.items[w];
    const StructDesc& D = G.descs[item.y];
    const int nx = D.nx, ny = D.ny, nz = D.nz;
    const uint32_t local = item.x - D.cell_base;
    const int ix = (int)(local % (uint32_t)nx), iy = (int)((local / (uint32_t)nx) % (uint32_t)ny),
              iz = (int)(local / ((uint32_t)nx * (uint32_t)ny));
    const uint32_t cb = G.cell_start[item.x], ce = G.cell_start[item.x + 1];

    // The 5 contiguous runs of the half shell, fetched by lanes 0-4 in parallel: run r covers
    // (dy, dz) = (0,0) (+1,0) (-1,+1) (0,+1) (+1,+1); x from ix (run 0) or ix-1 to ix+1.
    uint32_t my_b = 0, my_e = 0;
    if (lane < 5) {
      const int dy = (lane == 0) ? 0 : (lane == 1) ? 1 : (lane - 3);
      const int dz = (lane < 2) ? 0 : 1;
      const int y = iy + dy, z = iz + dz;
      if (y >= 0 && y < ny && z < nz) {
        const int x0 = (lane == 0) ? ix : max(ix - 1, 0), x1 = min(ix + 1, nx - 1);
        const uint32_t row = D.cell_base + (uint32_t)((z * ny + y) * nx);
        my_b = G.cell_start[row + x0];
        my_e = G.cell_start[row + x1 + 1];
      }
    }

    for (uint32_t fb = cb; fb < ce; fb += FROM_CHUNK) {
      const int nf = (int)min((uint32_t)FROM_CHUNK, ce - fb);
      const uint32_t full = nf == 32 ? 0xffffffffu : ((1u << nf) - 1u);
      __syncwarp();
      RecA mine_from;
      mine_from.attr = 0;
      if (lane < nf) {
        mine_from = G.recA[fb + lane];
        sfrom[lane] = mine_from;
      }
      // from-atoms that may start a row / be counted as `from`: ligand, not hydrogen, not a ghost
      const uint32_t elig_from = __ballot_sync(
          0xffffffffu, lane < nf && (mine_from.attr & ARP_ATTR_LIGAND) && (mine_from.attr & ARP_ATTR_ELEM_MASK) && !(mine_from.attr & ARP_ATTR_GHOST));
      __syncwarp();
#pragma unroll 1
      for (int r = 0; r < 5; ++r) {
        // run 0 starts inside the from-cell: candidates of this chunk are the atoms after fb (each pair i < t once)
        const uint32_t rb = __shfl_sync(0xffffffffu, my_b, r), se = __shfl_sync(0xffffffffu, my_e, r);
        const uint32_t sb = (r == 0) ? fb + 1 : rb;
        if (sb >= se) continue;
        RecA nxt;
        nxt.x = nxt.y = nxt.z = 3.0e18f;
        nxt.attr = 0;
        if (sb + lane < se) nxt = G.recA[sb + lane];
        for (uint32_t t0 = sb; t0 < se; t0 += 32) {
          const uint32_t t = t0 + lane;
          const bool valid = t < se;
          const RecA c = nxt;
          nxt.x = nxt.y = nxt.z = 3.0e18f;
          nxt.attr = 0;
          if (t + 32 < se) nxt = G.recA[t + 32];  // prefetch the next batch while this one is screened
          uint32_t m_hit = 0, m_cert = 0, m_near = 0, bit = 1u;
#pragma unroll 4
          for (int ii = 0; ii < nf; ++ii) {
            const RecA f = sfrom[ii];
            const float dx = c.x - f.x, dy = c.y - f.y, dz = c.z - f.z;
            const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
            if (d2 <= thr_hit) m_hit |= bit;
            if (d2 <= thr_cert) m_cert |= bit;
            if (d2 <= thr_near) m_near |= bit;
            bit <<= 1;
          }
          // inside the from-cell only from-atoms that precede the candidate count
          uint32_t allow = full;
          if (r == 0 && t < ce) allow = (t - fb < 32u) ? ((1u << (t - fb)) - 1u) & full : full;
          if (!valid) allow = 0;
          n_screened += __popc(allow);
          m_hit &= allow;
          m_cert &= m_hit;
          uint32_t push = m_near & m_hit;
          if (counting) {
            // certain pairs are counted here: (from -> candidate) for eligible from-atoms, (candidate -> from) if the
            // candidate is eligible
            const bool elig_c = (c.attr & ARP_ATTR_LIGAND) && (c.attr & ARP_ATTR_ELEM_MASK) && !(c.attr & ARP_ATTR_GHOST);
            n_in += __popc(m_cert & elig_from) + (elig_c ? __popc(m_cert) : 0);
            push |= m_hit & ~m_cert;  // ambiguous shell: the exact stage decides and counts
          }
          if (!__any_sync(0xffffffffu, push != 0u)) continue;
          const uint32_t uncounted = counting ? ~m_cert : 0u;
#pragma unroll 1
          for (int sbit = 0; sbit < nf; sbit += SUB_BITS) {
            uint32_t part = (push >> sbit) & ((1u << SUB_BITS) - 1u);
            if (!__any_sync(0xffffffffu, part != 0u)) continue;
            const uint32_t cnt = __popc(part);
            const uint32_t inc = warp_inclusive_sum(cnt, lane);
            int pos = q1len + (int)(inc - cnt);
            while (part) {
              const uint32_t ii = sbit + __ffs(part) - 1;
              part &= part - 1;
              q1[pos++] = make_uint2(fb + ii, t | (((uncounted >> ii) & 1u) ? Q_FLAG : 0u));
            }
            q1len += (int)__shfl_sync(0xffffffffu, inc, 31);
            __syncwarp();
            while (q1len >= 32) {
              exact_batch(G, P, q1 + (q1len - 32), 32, lane, q2, q2len, n_in, n_cls);
              q1len -= 32;
              __syncwarp();
            }
          }
        }
      }
    }
  }
  if (q1len > 0) exact_batch(G, P, q1, q1len, lane, q2, q2len, n_in, n_cls);
  __syncwarp();
  if (q2len > 0) classify_batch(G, P, q2, q2len, lane);
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
