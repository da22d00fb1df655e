// k_pairs main loop (stage 1 of the funnel described in pair_kernel.cuh).
//
// Staging.  The half shell of a cell is 5 contiguous runs of cell-sorted RecA records.  Each warp copies them,
// concatenated, into its own shared-memory buffer with TMA bulk copies (cp.async.bulk global -> shared, completion
// on an mbarrier), CHUNK records at a time, double buffered: while chunk k is screened the copies of chunk k+1
// (usually the next cell's) are in flight.  The chain of dependent global loads that describes a work item
// (cursor -> item -> cell_start values) is software-pipelined three items ahead, so none of it is exposed either.
#pragma once

namespace arpd {

// ---- PTX helpers: mbarrier + 1-D bulk copy ----------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}

// A work item as the warp holds it: lanes 0-4 carry one run each (begin slot, length, offset in the concatenation).
struct ItemMeta {
  uint32_t cb, n_own;       // own cell: first slot, atoms
  uint32_t run_b, run_len;  // lane r < 5: run r
  uint32_t run_off;         // lane r < 5: entries before run r
  uint32_t total;           // all candidates (own cell included)
  bool valid;
};

__device__ __forceinline__ uint32_t fetch_cursor(const PairArgs& G, int lane) {
  uint32_t w = 0;
  if (lane == 0) w = atomicAdd(&G.ctr->work, 1u);
  return __shfl_sync(0xffffffffu, w, 0);
}
__device__ __forceinline__ uint4 fetch_item(const PairArgs& G, uint32_t w, uint32_t n_items) {
  return w < n_items ? G.items[w] : make_uint4(0xffffffffu, 0u, 0u, 0u);
}
// cell_start lookups of one item; run r = (dy, dz) in (0,0) (+1,0) (-1,+1) (0,+1) (+1,+1), x from ix (run 0) or ix-1 to ix+1
__device__ __forceinline__ ItemMeta fetch_meta(const PairArgs& G, uint4 item, int lane) {
  ItemMeta m;
  m.valid = item.x != 0xffffffffu;
  m.cb = m.n_own = m.run_b = m.run_len = m.run_off = m.total = 0;
  if (!m.valid) return m;
  const StructDesc& D = G.descs[item.y];
  const int nx = D.nx, ny = D.ny, nz = D.nz;
  // cell coordinates travel with the item: ix | iy << 21 | iz << 42 in (z, w)
  const unsigned long long key = (unsigned long long)item.z | ((unsigned long long)item.w << 32);
  const int ix = (int)(key & 0x1fffffu), iy = (int)((key >> 21) & 0x1fffffu), iz = (int)(key >> 42);
  m.cb = G.cell_start[item.x];
  m.n_own = G.cell_start[item.x + 1] - m.cb;
  uint32_t b = 0, e = 0;
  if (lane < 5) {
    const int dy = (lane == 0) ? 0 : (lane == 1) ? 1 : (lane - 3);
    const int dz = (lane < 2) ? 0 : 1;
    const int y = iy + dy, z = iz + dz;
    if (y >= 0 && y < ny && z < nz) {
      const int x0 = (lane == 0) ? ix : max(ix - 1, 0), x1 = min(ix + 1, nx - 1);
      const uint32_t row = D.cell_base + (uint32_t)((z * ny + y) * nx);
      b = G.cell_start[row + x0];
      e = G.cell_start[row + x1 + 1];
    }
  }
  m.run_b = b;
  m.run_len = e - b;
  const uint32_t inc = warp_inclusive_sum(m.run_len, lane);
  m.run_off = inc - m.run_len;
  m.total = __shfl_sync(0xffffffffu, inc, 31);
  return m;
}

#ifndef PAIR_MIN_BLOCKS
#define PAIR_MIN_BLOCKS 3
#endif
constexpr int CHUNK = 128;  // records per staged chunk (2 KB)

struct PairSmemV3 {
  RecA buf[PAIR_WARPS][2][CHUNK];
  RecA from[PAIR_WARPS][FROM_CHUNK];
  uint2 q1[PAIR_WARPS][Q1_CAP];
  uint2 q2[PAIR_WARPS][Q2_CAP];
  uint64_t bar[PAIR_WARPS][2];
  uint32_t runs[PAIR_WARPS][16];  // current item: run_off[0..4], run_b[8..12]
};

// issue the bulk copies of entries [e0, e0 + CHUNK) of item m into dst
__device__ __forceinline__ void stage_chunk(const PairArgs& G, const ItemMeta& m, uint32_t e0, RecA* dst, uint64_t* bar, int lane) {
  const uint32_t e1 = min(e0 + (uint32_t)CHUNK, m.total);
  __syncwarp();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic reads of dst before the async writes
  if (lane == 0) mbar_expect_tx(bar, (e1 - e0) * (uint32_t)sizeof(RecA));
  __syncwarp();
  if (lane < 5) {
    const uint32_t lo = max(m.run_off, e0), hi = min(m.run_off + m.run_len, e1);
    if (lo < hi) bulk_g2s(dst + (lo - e0), G.recA + m.run_b + (lo - m.run_off), (hi - lo) * (uint32_t)sizeof(RecA), bar);
  }
}

__global__ void __launch_bounds__(PAIR_THREADS, PAIR_MIN_BLOCKS) k_pairs(const PairArgs G, const DevParams P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  PairSmemV3& S = *reinterpret_cast<PairSmemV3*>(smem_raw);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  RecA* sfrom = S.from[wid];
  uint2* q1 = S.q1[wid];
  uint2* q2 = S.q2[wid];
  uint32_t* sruns = S.runs[wid];
  int q1len = 0, q2len = 0;
  uint32_t n_in32 = 0;  // certain pairs counted by the screen, per lane; summed in 64 bit at the end
  unsigned long long n_screened = 0, n_in = 0, n_cls = 0;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t n_items = G.ctr->n_items;
  const float thr_hit = P.thr_hit, thr_cert = P.thr_cert, thr_near = P.thr_near;
  const bool counting = P.count_in_cutoff != 0;

  if (lane == 0) {
    mbar_init(&S.bar[wid][0], 1);
    mbar_init(&S.bar[wid][1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  uint32_t parity = 0u;  // bit b = phase parity to wait for on buffer b

  // ---- pipeline prologue: three cursors, two items, one meta in flight ----------------------------------------------
  uint32_t w_c = fetch_cursor(G, lane), w_d = fetch_cursor(G, lane), w_e = fetch_cursor(G, lane);
  uint4 it_b = fetch_item(G, w_c, n_items), it_c = fetch_item(G, w_d, n_items);
  ItemMeta cur = fetch_meta(G, it_b, lane);
  int pb = 0;  // buffer holding the current chunk
  if (cur.valid) stage_chunk(G, cur, 0, S.buf[wid][0], &S.bar[wid][0], lane);
  ItemMeta nxt = fetch_meta(G, it_c, lane);
  uint4 it_n = fetch_item(G, w_e, n_items);  // item after nxt
  uint32_t w_n = fetch_cursor(G, lane);       // cursor after that
  uint32_t chunk0 = 0;                        // first entry of the current chunk

  while (cur.valid) {
    // ---- what comes after the current chunk: the next chunk of this item, or chunk 0 of the next item -------------
    const bool more = chunk0 + CHUNK < cur.total;
    ItemMeta after_next;
    after_next.valid = false;
    uint4 it_nn = it_n;
    uint32_t w_nn = w_n;
    if (more) {
      stage_chunk(G, cur, chunk0 + CHUNK, S.buf[wid][pb ^ 1], &S.bar[wid][pb ^ 1], lane);
    } else {
      if (nxt.valid) stage_chunk(G, nxt, 0, S.buf[wid][pb ^ 1], &S.bar[wid][pb ^ 1], lane);
      // advance the item pipeline; these loads are consumed one iteration later
      after_next = fetch_meta(G, it_n, lane);
      it_nn = fetch_item(G, w_n, n_items);
      w_nn = fetch_cursor(G, lane);
    }

    // ---- current chunk -------------------------------------------------------------------------------------------------
    const RecA* buf = S.buf[wid][pb];
    mbar_wait(&S.bar[wid][pb], (parity >> pb) & 1u);
    parity ^= 1u << pb;
    const uint32_t n_own = cur.n_own;
    if (chunk0 == 0) {
      __syncwarp();
      if (lane < 5) {
        sruns[lane] = cur.run_off;
        sruns[8 + lane] = cur.run_b;
      }
      // distance tests this item stands for: own-cell pairs once, every other candidate against every own atom
      if (lane == 0) n_screened += (unsigned long long)n_own * (cur.total - n_own) + (unsigned long long)n_own * (n_own - 1) / 2;
      __syncwarp();
    }
    const uint32_t n_here = min((uint32_t)CHUNK, cur.total - chunk0);
    // from-atoms in passes of 32 (one pass for every realistic cell)
    for (uint32_t fbo = 0; fbo < n_own; fbo += FROM_CHUNK) {
      const int nf = (int)min((uint32_t)FROM_CHUNK, n_own - fbo);
      __syncwarp();
      RecA mine_from;
      mine_from.x = mine_from.y = mine_from.z = 0.f;
      mine_from.attr = 0;
      if (lane < nf) {
        // own-cell atoms are the first entries of the concatenation; chunk 0 holds them when fbo + nf <= CHUNK
        mine_from = (chunk0 == 0 && fbo + nf <= (uint32_t)CHUNK) ? buf[fbo + lane] : G.recA[cur.cb + fbo + lane];
        sfrom[lane] = mine_from;
      }
      // from-atoms that may start a row / be counted as `from`: ligand, not hydrogen, not a ghost
      const uint32_t elig_from = __ballot_sync(
          0xffffffffu, lane < nf && (mine_from.attr & ARP_ATTR_LIGAND) && (mine_from.attr & ARP_ATTR_ELEM_MASK) && !(mine_from.attr & ARP_ATTR_GHOST));
      __syncwarp();
      const uint32_t from_slot0 = cur.cb + fbo;
      for (uint32_t j0 = 0; j0 < n_here; j0 += 32) {
        const uint32_t j = j0 + lane;
        const uint32_t e = chunk0 + j;  // index in the concatenation
        RecA c;
        if (j < n_here) c = buf[j];
        else c.x = c.y = c.z = 3.0e18f, c.attr = 0;  // far away: fails every threshold
        const uint32_t elig_c = ((c.attr & ARP_ATTR_LIGAND) && (c.attr & ARP_ATTR_ELEM_MASK) && !(c.attr & ARP_ATTR_GHOST)) ? 1u : 0u;
        // global slot of this lane's candidate
        const uint32_t r = (e >= sruns[1]) + (e >= sruns[2]) + (e >= sruns[3]) + (e >= sruns[4]);
        const uint32_t t = sruns[8 + r] + (e - sruns[r]);
        // candidates inside the own cell pair only with from-atoms that precede them (each pair once)
        const bool own_group = chunk0 + j0 < n_own;
        const uint32_t first_ok = own_group ? (e >= n_own ? 0u : (e > fbo ? e - fbo : 0u)) : 0u;  // ii < first_ok allowed when e < n_own
        const bool in_own = own_group && e < n_own;
        for (int ii0 = 0; ii0 < nf; ii0 += 4) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int ii = ii0 + u;
            if (ii < nf) {
              const RecA f = sfrom[ii];
              const float dx = c.x - f.x, dy = c.y - f.y, dz = c.z - f.z;
              const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
              bool hit = d2 <= thr_hit;
              if (in_own) hit = hit && (uint32_t)ii < first_ok;
              const bool cert = hit && d2 <= thr_cert;
              bool push = hit && d2 <= thr_near;
              if (counting) {
                // certainly-inside pairs are counted here: (from -> candidate) if the from-atom is eligible, and back
                n_in32 += cert ? (((elig_from >> ii) & 1u) + elig_c) : 0u;
                push = push || (hit && !cert);  // ambiguous shell: the exact stage decides and counts
              }
              const uint32_t bal = __ballot_sync(0xffffffffu, push);
              if (bal) {
                if (push) q1[q1len + __popc(bal & lt_mask)] = make_uint2(from_slot0 + ii, t | ((counting && !cert) ? Q_FLAG : 0u));
                q1len += __popc(bal);
              }
            }
          }
          if (q1len >= 32) {
            __syncwarp();
            do {
              exact_batch(G, P, q1 + (q1len - 32), 32, lane, q2, q2len, n_in, n_cls);
              q1len -= 32;
              __syncwarp();
            } while (q1len >= 32);
          }
        }
      }
    }

    // ---- advance -----------------------------------------------------------------------------------------------------------
    pb ^= 1;
    if (more) {
      chunk0 += CHUNK;
    } else {
      chunk0 = 0;
      cur = nxt;
      nxt = after_next;
      it_n = it_nn;
      w_n = w_nn;
    }
  }
  if (q1len > 0) exact_batch(G, P, q1, q1len, lane, q2, q2len, n_in, n_cls);
  __syncwarp();
  if (q2len > 0) classify_batch(G, P, q2, q2len, lane);
  n_in += n_in32;
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
