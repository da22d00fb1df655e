// k_pairs: neighbour search + classification + compaction for atom-atom contacts (sm_100a).
// Replaces get_atomic_contacts, src/contacts/complex.rs:189-299.
//
// Work decomposition: one warp per non-empty cell ("from" cell), persistent CTAs pulling cells from an atomic
// cursor.  A warp visits the half shell of its cell -- the cell itself plus the 13 forward neighbours, which
// in the x-fastest cell order are 5 contiguous runs of sorted atoms -- so every unordered atom pair is screened
// once; both orientations are then filtered and classified.  Lanes own CANDIDATE atoms (coalesced 16-byte
// RecA loads, next batch prefetched into registers); the from-cell's atoms sit in shared memory and are
// broadcast to all lanes.
//
// Three-stage funnel, each stage running with (nearly) all lanes busy:
//   1. screen   f32 squared distance against three thresholds (rigorous guard band g, see api_run.inc):
//               hit  = d2 <= rc^2 + g          may be within the cutoff
//               cert = d2 <= rc^2 - g          certainly within the cutoff: counted here, no FP64 needed
//               near = d2 <= r_emit^2 + g      may be close enough for some label to fire
//               Only near pairs and the ambiguous shell (hit & !cert) go on; they are compacted by a warp
//               prefix sum into a shared-memory queue.
//   2. exact    (32 queued pairs at a time) FP64 d^2 exactly as the reference computes it, inclusive cutoff,
//               pair filter for both orientations; survivors are queued again.
//   3. classify (32 survivors at a time) sqrt, label ladder, rows out through one atomicAdd per batch.
#pragma once
#include "classify.cuh"

namespace arpd {

constexpr int PAIR_WARPS = 8;
constexpr int PAIR_THREADS = PAIR_WARPS * 32;
constexpr int FROM_CHUNK = 32;                 // from-atoms per pass (bits of the hit mask)
constexpr int Q1_CAP = 32 + 4 * 32;            // carry-over (< 32) + four ballots
constexpr int Q2_CAP = 32 + 64;                // carry-over + two orientations of one exact batch
constexpr uint32_t Q_FLAG = 0x80000000u;       // queue 1: pair still has to be counted; queue 2: orientation b -> a

struct PairSmem {
  RecA from[PAIR_WARPS][FROM_CHUNK];
  uint2 q1[PAIR_WARPS][Q1_CAP];
  uint2 q2[PAIR_WARPS][Q2_CAP];
};

__device__ __forceinline__ uint32_t warp_inclusive_sum(uint32_t v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// Stage 3: up to 32 (from, to) survivors, one per lane.
__device__ __forceinline__ void classify_batch(const PairArgs& G, const DevParams& P, const uint2* q, int count, int lane) {
  uint32_t m = 0, sp = 0, sq = 0;
  double dist = 0.0;
  if (lane < count) {
    const uint2 e = q[lane];
    // orientation flag: the row runs from e.y to e.x
    sp = (e.x & Q_FLAG) ? e.y : (e.x & ~Q_FLAG);
    sq = (e.x & Q_FLAG) ? (e.x & ~Q_FLAG) : e.y;
    const RecB p = G.recB[sp], r = G.recB[sq];
    const uint32_t wp = G.recA[sp].attr, wr = G.recA[sq].attr;
    // pdbtbx Atom::distance: sqrt((dx*dx + dy*dy) + dz*dz), every op rounded once
    const double dx = dsub(r.x, p.x), dy = dsub(r.y, p.y), dz = dsub(r.z, p.z);
    dist = __dsqrt_rn(dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz)));
    m = classify_pair(G, P, p, wp, sp, r, wr, sq, dist);
  }
  const uint32_t mine = __popc(m);
  if (!__any_sync(0xffffffffu, mine != 0u)) return;
  const uint32_t inc = warp_inclusive_sum(mine, lane);
  const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
  unsigned long long base = 0;
  if (lane == 31) base = atomicAdd(&G.ctr->rows, (unsigned long long)total);
  base = __shfl_sync(0xffffffffu, base, 31);
  unsigned long long w = base + (inc - mine);
  if (mine) {
    const uint32_t ip = G.perm[sp], iq = G.perm[sq];
    const float fd = __double2float_rn(dist);  // Rust `as f32`, mod.rs:148
    while (m) {
      const uint32_t l = __ffs(m) - 1;
      m &= m - 1;
      if (w < G.row_cap) G.rows[w] = arp_row{ip, iq, l, fd};
      ++w;
    }
  }
}

// Stage 2: up to 32 candidate pairs, one per lane: exact cutoff test, counters, pair filter; survivors -> q2.
__device__ __forceinline__ void exact_batch(const PairArgs& G, const DevParams& P, const uint2* q, int count, int lane, uint2* q2,
                                            int& q2len, unsigned long long& n_in_cutoff, unsigned long long& n_classified) {
  bool f_ab = false, f_ba = false;
  uint32_t sa = 0, sb = 0;
  if (lane < count) {
    const uint2 e = q[lane];
    sa = e.x;
    sb = e.y & ~Q_FLAG;
    const RecB a = G.recB[sa], b = G.recB[sb];
    const uint32_t wa = G.recA[sa].attr, wb = G.recA[sb].attr;
    // rstar distance_2: ((dx*dx + dy*dy) + dz*dz); locate_within_distance is inclusive
    const double dx = dsub(b.x, a.x), dy = dsub(b.y, a.y), dz = dsub(b.z, a.z);
    const double d2 = dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz));
    if (d2 <= P.rc2) {
      if (e.y & Q_FLAG) {  // not yet counted by the screen
        n_in_cutoff += ((wa & ARP_ATTR_LIGAND) && (wa & ARP_ATTR_ELEM_MASK) && !(wa & ARP_ATTR_GHOST)) ? 1u : 0u;
        n_in_cutoff += ((wb & ARP_ATTR_LIGAND) && (wb & ARP_ATTR_ELEM_MASK) && !(wb & ARP_ATTR_GHOST)) ? 1u : 0u;
      }
      f_ab = pair_filter(wa, a.chain_key, a.res_ordinal, wb, b.chain_key, b.res_ordinal);
      f_ba = pair_filter(wb, b.chain_key, b.res_ordinal, wa, a.chain_key, a.res_ordinal);
      n_classified += (f_ab ? 1u : 0u) + (f_ba ? 1u : 0u);
      if (!(d2 <= P.emit_d2_max)) f_ab = f_ba = false;  // no label can fire beyond the emit radius
    }
  }
  const uint32_t mine = (f_ab ? 1u : 0u) + (f_ba ? 1u : 0u);
  if (__any_sync(0xffffffffu, mine != 0u)) {
    const uint32_t inc = warp_inclusive_sum(mine, lane);
    int pos = q2len + (int)(inc - mine);
    if (f_ab) q2[pos++] = make_uint2(sa, sb);
    if (f_ba) q2[pos++] = make_uint2(sa | Q_FLAG, sb);
    q2len += (int)__shfl_sync(0xffffffffu, inc, 31);
    __syncwarp();
    while (q2len >= 32) {
      classify_batch(G, P, q2 + (q2len - 32), 32, lane);
      q2len -= 32;
      __syncwarp();
    }
  }
}

}  // namespace arpd

#include "pair_kernel_main.cuh"
