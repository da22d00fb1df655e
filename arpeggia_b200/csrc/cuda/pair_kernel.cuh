// Neighbour search + classification + compaction for atom-atom contacts (sm_100a), three kernels.
// Replaces get_atomic_contacts, src/contacts/complex.rs:189-299.
//
// k_screen (stage A: all the volume, ~116 distance tests per atom at protein density).
//   One warp per non-empty cell ("from" cell), persistent CTAs pulling work items (Item, dev_common.cuh) from an
//   atomic cursor.  A warp sweeps the HALF SHELL of its cell -- the cell itself plus its 13 forward neighbours,
//   5 contiguous runs of cell-sorted atoms -- so every unordered atom pair is screened once.  Lanes own CANDIDATES:
//   up to 6 per lane, held in registers in a cell-local f32 frame (cx, cy, cz, h = -|c|^2/2); the cell's own atoms
//   are broadcast from shared memory one after the other as (fx, fy, fz, k).  With s = f.c + h the test d^2 <= T
//   is s >= |f|^2/2 - T/2: three FMAs and compares against per-from-atom constants, no subtraction, no branch:
//       inside = T_near < d^2 <= T_cert   certainly within dist_cutoff and too far for any label: COUNTED here
//       push   = d^2 <= T_hit and not inside   (close enough for a label, or too close to the cutoff to call in f32)
//   `inside` pairs add a packed weight to a per-lane counter, `push` pairs set bit i of a per-lane mask; the masks
//   are expanded through a warp prefix sum into the job's PAIR QUEUE in HBM as (from slot, candidate slot), one
//   atomicAdd per warp and pass.  The guard band gA (api_run.inc) makes both decisions rigorous against the exact
//   FP64 distances.  T_near is per from-atom: no label can fire for an atom beyond its own largest limit.
//   Latency: a warp always has three items in flight -- cursor of item k+3 (atomic), record of item k+2 (one
//   coalesced 64-byte load), candidates of item k+1 (cp.async into a double-buffered shared-memory stage).
//
// k_filter (stage B1: what survives the screen, ~6 pairs per atom).  Grid-stride over the pair queue, one pair per lane:
//   f32 re-test on the stored positions, in-cutoff counting, per-pair label limit (no label of complex.rs:217-296 can
//   fire beyond max(vdW sum + comp, 4.5 | 4.0 | 3.5 by flags)), pair filter (at most one orientation of a pair can pass
//   should_compare_entities, complex.rs:76-131).  About one pair in nine goes on to the second queue.
// k_exact (stage B2, ~0.7 pairs per atom).  Exact FP64 d^2 in the reference's operation order, inclusive cutoff (rstar),
//   sqrt, label ladder, rows out through a warp prefix sum and ONE atomicAdd per 32 pairs.
//
// Why three kernels: fused into one persistent kernel the stages made a 70-160 KB instruction footprint with every
// warp in a different place; ncu showed 27-36 % of the stall samples as "no instruction" (the L1.5 I-cache is 32 KB)
// and the stage-B gathers stalled behind the screen's registers.  Split, each kernel is a tight loop, k_screen keeps no
// stage-B state, k_filter runs at full occupancy to hide its gathers and k_exact carries the heavy FP64 ladder for the few
// pairs that need it.  The price is the queues' trip through HBM: 8 bytes written and read per surviving pair.
#pragma once
#include "classify.cuh"

namespace arpd {

constexpr int PAIR_WARPS = 4;
constexpr int PAIR_THREADS = PAIR_WARPS * 32;
#ifndef PAIR_MIN_BLOCKS
#define PAIR_MIN_BLOCKS 5
#endif
constexpr int SLOTS = 6;                 // candidates per lane per chunk
constexpr int CHUNK = 32 * SLOTS;        // candidates per chunk
constexpr int FROM_PASS = 32;            // from-atoms per pass (bits of a mask)
constexpr int CLS_WARPS = 8;
constexpr int CLS_THREADS = CLS_WARPS * 32;
constexpr uint32_t CHUNKQ = 512;         // pair-queue entries a k_screen warp reserves at a time
constexpr uint32_t Q_SENTINEL = 0xffffffffu;  // pair-queue entry that holds no pair (unused tail of a reserved window)
constexpr uint32_t SLOT_MASK = 0x3fffffffu;
constexpr uint32_t QF_A = 0x80000000u;   // q2.x: in-cutoff still undecided (count in B2)   q2.y: row runs candidate -> from
constexpr uint32_t QF_B = 0x40000000u;   // q2.x: the pair filter passed                     q2.y: a label may fire

// One from-atom as the screen reads it (two LDS.128 broadcasts).
#ifndef SCREEN_F32X2
#define SCREEN_F32X2 0  // 1: the screen's dot products run as packed fma.rn.f32x2 (two candidate slots per instruction)
#endif
#if SCREEN_F32X2
struct __align__(16) FromRec {  // 48 bytes: coordinates duplicated so that a 64-bit shared load yields the (f, f) operand
  float x, x2, y, y2;        // cell-local position
  float z, z2, k_hit, k_cert;  // s >= k_hit <=> d^2 <= T_hit; s >= k_cert <=> d^2 <= T_cert
  float k_near;              // s < k_near <=> d^2 > T_near
  uint32_t w;                // (eligible as `from` ? 1 << 16 : 0) | 1
  uint32_t pad[2];
};
#else
struct __align__(16) FromRec {
  float x, y, z, k_hit;      // cell-local position; s >= k_hit <=> d^2 <= T_hit
  float k_cert, k_near;      // s >= k_cert <=> d^2 <= T_cert; s < k_near <=> d^2 > T_near
  uint32_t w;                // (eligible as `from` ? 1 << 16 : 0) | 1
  uint32_t pad;
};
#endif

struct ScreenSmem {
  uint4 buf[PAIR_WARPS][2][CHUNK];      // staged candidate records (RecA bits), double buffered
  uint32_t tbuf[PAIR_WARPS][2][CHUNK];  // their sorted slots
  FromRec from[PAIR_WARPS][FROM_PASS + 1];
  unsigned int* cursor[32];             // &ctr->work, one copy per lane (see draw_cursor)
  float lim1[16];                       // per element class: largest (max(cov sum, vdW sum) + |comp|)^2 + g over partners
};

struct PairCounts {
  unsigned long long n_in, n_cls;
};

__device__ __forceinline__ uint32_t warp_inclusive_sum(uint32_t v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// atoms that can start a row / are counted as `from`: in the ligand set, not hydrogen, not a ghost (complex.rs:200-202);
// k_scatter_atoms caches the predicate in bit 31 of the sorted attribute word
__device__ __forceinline__ uint32_t eligible(uint32_t w) { return w >> 31; }

// per element-class pair: no clash / covalent / vdW label beyond (max(cov sum, vdW sum) + |comp|)^2 (vdw.rs:33-41)
__device__ __forceinline__ float class_pair_limit2(const DevParams& P, int ea, int eb) {
  const double r = fmax(P.cov[ea] + P.cov[eb], P.vdw[ea] + P.vdw[eb]) + fabs(P.comp);
  const double r2 = r > 0.0 ? r * r * (1.0 + 1e-9) : 0.0;
  return __double2float_ru(r2) + P.guard;
}

// ---- B2: up to 32 filtered pairs, one per lane: exact distance, counters, label ladder, rows ----------------------
// The rows of a trip are written one trip later: the atomicAdd that reserves their slots is issued at the end of a trip
// and its result is first needed after the next trip's gathers have been issued, so neither latency is exposed.
struct PendingRows {
  unsigned long long base;  // lane 31: result of the reservation
  uint32_t m, ip, iq, off;  // this lane's label bits, row ends, offset within the reservation
  float fd;
};
__device__ __forceinline__ void flush_rows(const PairArgs& G, PendingRows& R) {
  const unsigned long long base = __shfl_sync(0xffffffffu, R.base, 31);
  unsigned long long w = base + R.off;
  uint32_t m = R.m;
  while (m) {
    const uint32_t l = __ffs(m) - 1;
    m &= m - 1;
    if (w < G.row_cap) G.rows[w] = arp_row{R.ip, R.iq, l, R.fd};
    ++w;
  }
  R.m = 0u;
}
// the records of one queue entry: ONE 32-byte sector per atom (f64 position, attribute word, residue id); the input
// indices the rows carry ride in the queue entry itself
struct ExactOperands {
  RecB a, b;
};
__device__ __forceinline__ void exact_gather(const PairArgs& G, const uint4 e, bool have, ExactOperands& X) {
  if (!have) return;
  ARP_CHK(G.ctr, (e.x & SLOT_MASK) < G.n_atoms && (e.y & SLOT_MASK) < G.n_atoms, 0);  // queue entries decode to sorted slots
  X.a = G.recB[e.x & SLOT_MASK];
  X.b = G.recB[e.y & SLOT_MASK];
}
__device__ __forceinline__ void exact_batch(const PairArgs& G, const DevParams& P, const uint4 e, bool have, const ExactOperands& X, int lane,
                                            PairCounts& K, PendingRows& R) {
  uint32_t m = 0, ip = 0, iq = 0;
  double dist = 0.0;
  if (have) {
    const RecB &a = X.a, &b = X.b;
    const uint32_t wa = a.attr, wb = b.attr;
    // rstar distance_2 / pdbtbx Atom::distance: ((dx*dx + dy*dy) + dz*dz), every operation rounded once
    const double dx = dsub(b.x, a.x), dy = dsub(b.y, a.y), dz = dsub(b.z, a.z);
    const double d2 = dadd(dadd(dmul(dx, dx), dmul(dy, dy)), dmul(dz, dz));
    if (d2 <= P.rc2) {  // locate_within_distance is inclusive
      if (e.x & QF_A) {
        K.n_in += eligible(wa) + eligible(wb);
        if (P.count_classified && (e.x & QF_B)) K.n_cls += 1u;
      }
      if ((e.x & QF_B) && (e.y & QF_B) && d2 <= P.emit_d2_max) {
        dist = __dsqrt_rn(d2);
        const bool swap = (e.y & QF_A) != 0u;
        ip = swap ? e.w : e.z;
        iq = swap ? e.z : e.w;
        // one instance of the ladder (instruction-cache footprint): operands selected, not two inlined calls
        RecB p, r;
        p.x = swap ? b.x : a.x, p.y = swap ? b.y : a.y, p.z = swap ? b.z : a.z;
        r.x = swap ? a.x : b.x, r.y = swap ? a.y : b.y, r.z = swap ? a.z : b.z;
        p.attr = r.attr = 0u;  // classify_pair reads positions only
        p.res_id = r.res_id = 0u;
        m = classify_pair(G, P, p, swap ? wb : wa, swap ? b.res_id : a.res_id, r, swap ? wa : wb, swap ? a.res_id : b.res_id, dist);
      }
    }
  }
  const uint32_t mine = __popc(m);
  if (!__any_sync(0xffffffffu, mine != 0u)) return;
  const uint32_t inc = warp_inclusive_sum(mine, lane);
  R.base = 0ull;
  if (lane == 31) R.base = atomicAdd(&G.ctr->rows, (unsigned long long)inc);
  R.m = m, R.ip = ip, R.iq = iq, R.off = inc - mine;
  R.fd = __double2float_rn(dist);  // Rust `as f32`, mod.rs:148
}

// largest squared distance (f32, guard included) at which any label of complex.rs:217-296 can fire for this pair
__device__ __forceinline__ float label_limit2(const DevParams& P, const float* lim2, uint32_t wa, uint32_t wb) {
  float l = lim2[((wa & ARP_ATTR_ELEM_MASK) << 4) | (wb & ARP_ATTR_ELEM_MASK)];  // clash / covalent / vdW ladder
  const uint32_t both = wa & wb;
  // ionic bond (opposite charges) and ionic repulsion (like charges): <= 4.0
  const bool charged = ((wa & (ARP_ATTR_POS_CONF | ARP_ATTR_NEG_CONF)) && (wb & (ARP_ATTR_POS_CONF | ARP_ATTR_NEG_CONF)));
  if (charged) l = fmaxf(l, 16.0f + P.guard);
  // (weak) hydrogen bond / polar contact: an acceptor on one side, a (weak) donor on the other
  const bool polar = ((wa & ARP_ATTR_ACCEPTOR) && (wb & (ARP_ATTR_DONOR | ARP_ATTR_WEAK_DONOR))) ||
                     ((wb & ARP_ATTR_ACCEPTOR) && (wa & (ARP_ATTR_DONOR | ARP_ATTR_WEAK_DONOR)));
  if (polar) l = fmaxf(l, P.polar_d2);
  if (both & ARP_ATTR_HYDROPHOBIC) l = fmaxf(l, 20.25f + P.guard);
  return l;
}

// ---- k_filter (B1) ----------------------------------------------------------------------------------------------
// Grid-stride over the pair queue, one pair per lane, in two compacted phases so that the expensive half only runs on
// full warps of pairs that need it:
//   phase 1 (every pair)      one 32-byte gather per atom, f32 re-test on the stored positions, in-cutoff counting, and
//                             the cheapest strong filter: same chain and residue ordinals at most 1 apart can pass
//                             neither orientation (complex.rs:112-113) -- two pairs in three end here;
//   phase 2 (32 at a time)    per-pair label limit, pair filter with its orientation; about one queue entry in nine goes
//                             on to the second queue, 32 at a time.
struct FilterOut {
  bool go;
  uint4 out;
};
__device__ __forceinline__ FilterOut filter_phase2(const PairArgs& G, const DevParams& P, const float* lim2, uint2 e, bool have,
                                                   uint32_t& n_cls32) {
  FilterOut r{false, make_uint4(0u, 0u, 0u, 0u)};
  if (have) {
    const uint4* pa = reinterpret_cast<const uint4*>(G.recA + e.x);
    const uint4* pb = reinterpret_cast<const uint4*>(G.recA + e.y);
    const uint4 a0 = pa[0], a1 = pa[1], b0 = pb[0], b1 = pb[1];  // just read by phase 1: L1 hits
    const float dx = __uint_as_float(b0.x) - __uint_as_float(a0.x), dy = __uint_as_float(b0.y) - __uint_as_float(a0.y),
                dz = __uint_as_float(b0.z) - __uint_as_float(a0.z);
    const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));  // the same operations as phase 1: the same value
    const bool cert = d2 <= P.thr_cert;
    const bool amb = P.count_in_cutoff != 0 && !cert;  // the exact stage decides and counts
    const bool pot = d2 <= P.thr_near && d2 <= label_limit2(P, lim2, a0.w, b0.w);
    const int dir = pair_direction(a0.w, a1.x, a1.y, b0.w, b1.x, b1.y);
    const bool valid = dir != 0;
    if (P.count_classified && cert) n_cls32 += valid ? 1u : 0u;
    r.go = amb || (pot && valid);
    r.out.x = e.x | (amb ? QF_A : 0u) | (valid ? QF_B : 0u);
    r.out.y = e.y | (dir == 2 ? QF_A : 0u) | (pot ? QF_B : 0u);
    r.out.z = a1.w;  // input indices of the two atoms: the exact stage writes them into the rows
    r.out.w = b1.w;
  }
  return r;
}

// what phase 1 reads of a queue entry: the first half of both 32-byte records and (chain key, ordinal) of the second
struct FilterGather {
  uint4 a0, b0;
  uint2 a1, b1;
};
__device__ __forceinline__ void filter_gather(const PairArgs& G, const uint2 e, FilterGather& g) {
  if (e.x == Q_SENTINEL) return;
  const uint4* pa = reinterpret_cast<const uint4*>(G.recA + e.x);
  const uint4* pb = reinterpret_cast<const uint4*>(G.recA + e.y);
  g.a0 = pa[0];
  g.b0 = pb[0];
  g.a1 = *reinterpret_cast<const uint2*>(pa + 1);
  g.b1 = *reinterpret_cast<const uint2*>(pb + 1);
}

__global__ void __launch_bounds__(CLS_THREADS, 4) k_filter(const PairArgs G, const DevParams P) {
  __shared__ float lim2[256];  // per element-class pair: (max(cov sum, vdW sum) + |comp|)^2 + g, rounded up
  __shared__ uint2 stash1[CLS_WARPS][64];  // phase-1 survivors waiting for a full warp (carry-over < 32 + one trip)
  __shared__ uint4 stash2[CLS_WARPS][64];  // phase-2 survivors waiting for a full 32-entry flush
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int k = threadIdx.x; k < 256; k += CLS_THREADS) lim2[k] = class_pair_limit2(P, k >> 4, k & 15);
  __syncthreads();
  uint2* s1 = stash1[wid];
  uint4* s2 = stash2[wid];
  int n1 = 0, n2 = 0;
  uint32_t n_in32 = 0, n_cls32 = 0;
  const unsigned long long n_pairs = min(G.ctr->q_len, G.pairq_cap);
  const unsigned long long stride = (unsigned long long)gridDim.x * CLS_WARPS * 32ull;
  const bool counting = P.count_in_cutoff != 0;
  const uint32_t lt = (1u << lane) - 1u;
  unsigned long long base = ((unsigned long long)blockIdx.x * CLS_WARPS + wid) * 32ull;
  // No dependent memory round trip inside a trip: the queue entry is read two trips ahead, the records of its two atoms
  // are gathered one trip ahead (the loads fly while the current trip computes), and the output atomic is paid once per
  // 32 survivors.  The trip after the last one flushes the stashes.
  FilterGather g_next{};
  uint2 e_gath = make_uint2(Q_SENTINEL, Q_SENTINEL), e_next = make_uint2(Q_SENTINEL, Q_SENTINEL);
  if (base + lane < n_pairs) e_gath = __ldcs(G.pairq + base + lane);  // streamed once: keep the atom records in L2 instead
  if (base + stride + lane < n_pairs) e_next = __ldcs(G.pairq + base + stride + lane);
  filter_gather(G, e_gath, g_next);
  while (true) {
    const bool last = base >= n_pairs;
    const uint2 e = e_gath;
    const FilterGather g = g_next;
    e_gath = e_next;
    e_next = make_uint2(Q_SENTINEL, Q_SENTINEL);
    if (base + 2ull * stride + lane < n_pairs) e_next = __ldcs(G.pairq + base + 2ull * stride + lane);
    filter_gather(G, e_gath, g_next);
    // ---- phase 1 ----------------------------------------------------------------------------------------------------
    bool keep = false;
    if (!last && e.x != Q_SENTINEL) {  // sentinel: past the end, or the unused tail of a k_screen queue window
      const uint4 a0 = g.a0, b0 = g.b0;  // position + attr
      const uint2 a1 = g.a1, b1 = g.b1;  // chain key, ordinal
      const float dx = __uint_as_float(b0.x) - __uint_as_float(a0.x), dy = __uint_as_float(b0.y) - __uint_as_float(a0.y),
                  dz = __uint_as_float(b0.z) - __uint_as_float(a0.z);
      const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
      if (d2 <= P.thr_hit) {
        const bool cert = d2 <= P.thr_cert;
        if (counting && cert) n_in32 += eligible(a0.w) + eligible(b0.w);
        const bool close_in_chain = a1.x == b1.x && (a1.y - b1.y + 1u) <= 2u;  // same chain, |ordinal difference| <= 1
        keep = (counting && !cert) || P.count_classified || (d2 <= P.thr_near && !close_in_chain);
      }
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, keep);
    if (keep) s1[n1 + __popc(bal & lt)] = e;
    n1 += __popc(bal);
    __syncwarp();
    // ---- phase 2 on full warps (on whatever is left after the last trip) ------------------------------------------------
    while (n1 >= 32 || (last && n1 > 0)) {
      const int m = min(n1, 32);
      n1 -= m;
      const FilterOut r = filter_phase2(G, P, lim2, s1[n1 + (lane < m ? lane : 0)], lane < m, n_cls32);
      const uint32_t b2 = __ballot_sync(0xffffffffu, r.go);
      if (r.go) s2[n2 + __popc(b2 & lt)] = r.out;
      n2 += __popc(b2);
      __syncwarp();
      if (n2 >= 32 || (last && n1 == 0 && n2 > 0)) {
        const int k = min(n2, 32);
        unsigned long long w = 0;
        if (lane == 0) w = atomicAdd(&G.ctr->q2_len, (unsigned long long)k);
        w = __shfl_sync(0xffffffffu, w, 0) + lane;
        n2 -= k;
        if (lane < k && w < G.pairq2_cap) __stcs(G.pairq2 + w, s2[n2 + lane]);
        __syncwarp();
      }
    }
    if (last) {
      if (n2 > 0) {  // phase 2 did not run on the last trip
        unsigned long long w = 0;
        if (lane == 0) w = atomicAdd(&G.ctr->q2_len, (unsigned long long)n2);
        w = __shfl_sync(0xffffffffu, w, 0) + lane;
        if (lane < n2 && w < G.pairq2_cap) __stcs(G.pairq2 + w, s2[lane]);
      }
      break;
    }
    base += stride;
  }
  unsigned long long n_in = n_in32, n_cls = n_cls32;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_in += __shfl_xor_sync(0xffffffffu, n_in, o);
    n_cls += __shfl_xor_sync(0xffffffffu, n_cls, o);
  }
  if (lane == 0) {
    if (n_in) atomicAdd(&G.ctr->in_cutoff, n_in);
    if (n_cls) atomicAdd(&G.ctr->classified, n_cls);
  }
}

// ---- k_exact (B2) -----------------------------------------------------------------------------------------------
#ifndef EXACT_MIN_BLOCKS
#define EXACT_MIN_BLOCKS 3
#endif
__global__ void __launch_bounds__(CLS_THREADS, EXACT_MIN_BLOCKS) k_exact(const PairArgs G, const DevParams P) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  PairCounts K{0ull, 0ull};
  const unsigned long long n_pairs = min(G.ctr->q2_len, G.pairq2_cap);
  const unsigned long long stride = (unsigned long long)gridDim.x * CLS_WARPS * 32ull;
  unsigned long long base = ((unsigned long long)blockIdx.x * CLS_WARPS + wid) * 32ull;
#ifndef EXACT_PREFETCH
#define EXACT_PREFETCH 0  // r2b: 0.035 -> 0.038 ms with it (80 registers spill; 120 registers at 2 CTAs / SM: 0.039): off.  1: the records of trip t + 1 are gathered before trip t computes (entries read two trips ahead)
#endif
  PendingRows R{0ull, 0u, 0u, 0u, 0u, 0.f};
#if EXACT_PREFETCH
  const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
  uint4 e = zero, e_next = zero;
  if (base + lane < n_pairs) e = __ldcs(G.pairq2 + base + lane);
  if (base + stride + lane < n_pairs) e_next = __ldcs(G.pairq2 + base + stride + lane);
  ExactOperands X;
  exact_gather(G, e, base + lane < n_pairs, X);
  for (; base < n_pairs; base += stride) {
    const bool have = base + lane < n_pairs, have_n = base + stride + lane < n_pairs;
    const uint4 e_cur = e;
    const ExactOperands X_cur = X;
    e = e_next;
    e_next = zero;
    if (base + 2ull * stride + lane < n_pairs) e_next = __ldcs(G.pairq2 + base + 2ull * stride + lane);
    exact_gather(G, e, have_n, X);                              // the next trip's records fly ...
    if (__any_sync(0xffffffffu, R.m != 0u)) flush_rows(G, R);  // ... under the previous trip's rows
    exact_batch(G, P, e_cur, have, X_cur, lane, K, R);          // ... and this trip's ladder
  }
#else
  uint4 e_next = make_uint4(0u, 0u, 0u, 0u);  // read one trip ahead
  if (base + lane < n_pairs) e_next = __ldcs(G.pairq2 + base + lane);
  for (; base < n_pairs; base += stride) {
    const uint4 e = e_next;
    const bool have = base + lane < n_pairs;
    if (base + stride + lane < n_pairs) e_next = __ldcs(G.pairq2 + base + stride + lane);
    ExactOperands X;
    exact_gather(G, e, have, X);
    if (__any_sync(0xffffffffu, R.m != 0u)) flush_rows(G, R);  // the previous trip's rows, while the gathers fly
    exact_batch(G, P, e, have, X, lane, K, R);
  }
#endif
  if (__any_sync(0xffffffffu, R.m != 0u)) flush_rows(G, R);
  unsigned long long n_in = K.n_in, n_cls = K.n_cls;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_in += __shfl_xor_sync(0xffffffffu, n_in, o);
    n_cls += __shfl_xor_sync(0xffffffffu, n_cls, o);
  }
  if (lane == 0) {
    if (n_in) atomicAdd(&G.ctr->in_cutoff, n_in);
    if (n_cls) atomicAdd(&G.ctr->classified, n_cls);
  }
}

// ---- stage A ----------------------------------------------------------------------------------------------------
struct Cand {
  float x[SLOTS], y[SLOTS], z[SLOTS], h[SLOTS];
};

#if SCREEN_F32X2
// sm_100 has packed FP32: fma.rn.f32x2 does two independent FMAs per issue slot.  The kernel is issue-bound, so the
// three FMAs of a test are issued once per PAIR of candidate slots: 6.5 instead of 8 instructions per test.
struct CandP {
  unsigned long long x[SLOTS / 2], y[SLOTS / 2], z[SLOTS / 2], h[SLOTS / 2];  // (slot 2p, slot 2p + 1)
};
__device__ __forceinline__ unsigned long long pack2(float a, float b) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
// inside = tri && s >= k_cert && s < k_near;  push = tri && s >= k_hit && !inside; predicated add / or
// (written in PTX: the compiler otherwise turns both updates into SEL + ALU pairs)
template <bool TRI>
__device__ __forceinline__ void screen_test(float s, float k_hit, float k_cert, float k_near, uint32_t w, uint32_t bit, uint32_t i, uint32_t lim,
                                            uint32_t& cnt, uint32_t& mask) {
  if (TRI) {
    asm("{\n\t.reg .pred pt, pi, pp;\n\t"
        "setp.lt.u32 pt, %8, %9;\n\t"
        "setp.ge.and.f32 pi, %2, %3, pt;\n\t"
        "setp.lt.and.f32 pi, %2, %4, pi;\n\t"
        "setp.ge.and.f32 pp, %2, %5, pt;\n\t"
        "and.pred pp, pp, !pi;\n\t"
        "@pi add.u32 %0, %0, %6;\n\t"
        "@pp or.b32 %1, %1, %7;\n\t}"
        : "+r"(cnt), "+r"(mask)
        : "f"(s), "f"(k_cert), "f"(k_near), "f"(k_hit), "r"(w), "r"(bit), "r"(i), "r"(lim));
  } else {
    asm("{\n\t.reg .pred pi, pp;\n\t"
        "setp.ge.f32 pi, %2, %3;\n\t"
        "setp.lt.and.f32 pi, %2, %4, pi;\n\t"
        "setp.ge.and.f32 pp, %2, %5, !pi;\n\t"
        "@pi add.u32 %0, %0, %6;\n\t"
        "@pp or.b32 %1, %1, %7;\n\t}"
        : "+r"(cnt), "+r"(mask)
        : "f"(s), "f"(k_cert), "f"(k_near), "f"(k_hit), "r"(w), "r"(bit));
  }
}
// one from-atom against the NS candidates of this lane
template <bool ALL, int NS>
__device__ __forceinline__ void screen_one(const FromRec* f, int i, const CandP& C, const uint32_t (&lim)[SLOTS], uint32_t (&mask)[SLOTS],
                                           uint32_t (&cnt)[SLOTS]) {
  const ulonglong2 A = *reinterpret_cast<const ulonglong2*>(f);        // (x, x), (y, y)
  const ulonglong2 B = *(reinterpret_cast<const ulonglong2*>(f) + 1);  // (z, z), (k_hit, k_cert)
  const uint2 D = *reinterpret_cast<const uint2*>(&f->k_near);         // k_near, w
  float k_hit, k_cert;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(k_hit), "=f"(k_cert) : "l"(B.y));
  const float k_near = __uint_as_float(D.x);
  const uint32_t w = D.y, bit = 1u << i;
#pragma unroll
  for (int p = 0; p < (NS + 1) / 2; ++p) {
    const unsigned long long s2 = fma2(A.x, C.x[p], fma2(A.y, C.y[p], fma2(B.x, C.z[p], C.h[p])));
    float s0, s1;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(s0), "=f"(s1) : "l"(s2));
    if (ALL || p == 0)
      screen_test<true>(s0, k_hit, k_cert, k_near, w, bit, (uint32_t)i, lim[2 * p], cnt[2 * p], mask[2 * p]);
    else
      screen_test<false>(s0, k_hit, k_cert, k_near, w, bit, 0u, 0u, cnt[2 * p], mask[2 * p]);
    if (2 * p + 1 < NS) {
      if (ALL)
        screen_test<true>(s1, k_hit, k_cert, k_near, w, bit, (uint32_t)i, lim[2 * p + 1], cnt[2 * p + 1], mask[2 * p + 1]);
      else
        screen_test<false>(s1, k_hit, k_cert, k_near, w, bit, 0u, 0u, cnt[2 * p + 1], mask[2 * p + 1]);
    }
  }
}
template <bool ALL, int NS>
__device__ __forceinline__ void screen_pass(const FromRec* sf, int nf, const CandP& C, const uint32_t (&lim)[SLOTS], uint32_t (&mask)[SLOTS],
                                            uint32_t (&cnt)[SLOTS]) {
#pragma unroll 1
  for (int i = 0; i < nf; i += 2) {  // two from-atoms per trip; sf has FROM_PASS + 1 entries
    screen_one<ALL, NS>(sf + i, i, C, lim, mask, cnt);
    if (i + 1 < nf) screen_one<ALL, NS>(sf + i + 1, i + 1, C, lim, mask, cnt);
  }
}
#else
// one from-atom (f = x, y, z, k_hit; g = k_cert, k_near, w) against the NS candidates of this lane
template <bool ALL, int NS>
__device__ __forceinline__ void screen_one(const float4 f, const float4 g, int i, const Cand& C, const uint32_t (&lim)[SLOTS],
                                           uint32_t (&mask)[SLOTS], uint32_t (&cnt)[SLOTS]) {
  const uint32_t w = __float_as_uint(g.z);
  const uint32_t bit = 1u << i;
#pragma unroll
  for (int u = 0; u < NS; ++u) {
    const float s = fmaf(f.x, C.x[u], fmaf(f.y, C.y[u], fmaf(f.z, C.z[u], C.h[u])));
    // inside = tri && s >= k_cert && s < k_near;  push = tri && s >= k_hit && !inside; predicated add / or
    // (written in PTX: the compiler otherwise turns both updates into SEL + ALU pairs)
    if (ALL || u == 0) {
      asm("{\n\t.reg .pred pt, pi, pp;\n\t"
          "setp.lt.u32 pt, %8, %9;\n\t"
          "setp.ge.and.f32 pi, %2, %3, pt;\n\t"
          "setp.lt.and.f32 pi, %2, %4, pi;\n\t"
          "setp.ge.and.f32 pp, %2, %5, pt;\n\t"
          "and.pred pp, pp, !pi;\n\t"
          "@pi add.u32 %0, %0, %6;\n\t"
          "@pp or.b32 %1, %1, %7;\n\t}"
          : "+r"(cnt[u]), "+r"(mask[u])
          : "f"(s), "f"(g.x), "f"(g.y), "f"(f.w), "r"(w), "r"(bit), "r"((uint32_t)i), "r"(lim[u]));
    } else {
      asm("{\n\t.reg .pred pi, pp;\n\t"
          "setp.ge.f32 pi, %2, %3;\n\t"
          "setp.lt.and.f32 pi, %2, %4, pi;\n\t"
          "setp.ge.and.f32 pp, %2, %5, !pi;\n\t"
          "@pi add.u32 %0, %0, %6;\n\t"
          "@pp or.b32 %1, %1, %7;\n\t}"
          : "+r"(cnt[u]), "+r"(mask[u])
          : "f"(s), "f"(g.x), "f"(g.y), "f"(f.w), "r"(w), "r"(bit));
    }
  }
}

// The screen: every from-atom of the pass against the NS candidates of this lane, two from-atoms per trip.
// lim[u] = number of leading from-atoms candidate u may pair with (own-cell candidates pair only with the atoms that
// precede them, everything else with all of them).  ALL = false: only slot 0 can hold own-cell atoms (cells of at most
// 32 atoms, chunk 0), the other slots skip the comparison.  NS = slots of the chunk that hold candidates (the last
// chunk of an item is rarely full).  All variants together are a few KB of code (the L1.5 I-cache is 32 KB).
template <bool ALL, int NS>
__device__ __forceinline__ void screen_pass(const FromRec* sf, int nf, const Cand& C, const uint32_t (&lim)[SLOTS], uint32_t (&mask)[SLOTS],
                                            uint32_t (&cnt)[SLOTS]) {
#pragma unroll 1
  for (int i = 0; i < nf; i += 2) {
    const float4 f0 = *reinterpret_cast<const float4*>(&sf[i]);
    const float4 g0 = *(reinterpret_cast<const float4*>(&sf[i]) + 1);
    const float4 f1 = *reinterpret_cast<const float4*>(&sf[i + 1]);  // sf has FROM_PASS + 1 entries
    const float4 g1 = *(reinterpret_cast<const float4*>(&sf[i + 1]) + 1);
    screen_one<ALL, NS>(f0, g0, i, C, lim, mask, cnt);
    if (i + 1 < nf) screen_one<ALL, NS>(f1, g1, i + 1, C, lim, mask, cnt);
  }
}

#endif

// Per-warp window into the pair queue: space is reserved CHUNKQ entries at a time (one atomicAdd per ~10 work items
// instead of one per item); the unused tail of a window is filled with sentinel entries that k_filter skips.
struct QWindow {
  unsigned long long pos, end;  // next entry to write / end of the reserved window
};
__device__ __forceinline__ void qwindow_close(const PairArgs& G, QWindow& Q, int lane) {
  for (unsigned long long p = Q.pos + lane; p < Q.end; p += 32)
    if (p < G.pairq_cap) __stcs(G.pairq + p, make_uint2(Q_SENTINEL, Q_SENTINEL));
  Q.pos = Q.end;
}

// masks -> pair queue in HBM.  Entries are (from slot, candidate slot).  Writes beyond the capacity are dropped; the
// counter keeps counting, the host grows the queue and re-runs the job.
__device__ __forceinline__ void expand_masks(const PairArgs& G, uint32_t (&mask)[SLOTS], const uint32_t* tbuf, uint32_t from_slot0,
                                             QWindow& Q, int lane) {
  uint32_t mine = 0;
#pragma unroll
  for (int u = 0; u < SLOTS; ++u) mine += __popc(mask[u]);
  if (!__any_sync(0xffffffffu, mine != 0u)) return;
  const uint32_t inc = warp_inclusive_sum(mine, lane);
  const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
  if (Q.pos + total > Q.end) {  // warp-uniform
    qwindow_close(G, Q, lane);
    const unsigned long long need = (unsigned long long)max(total, CHUNKQ);
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(&G.ctr->q_len, need);
    Q.pos = __shfl_sync(0xffffffffu, base, 0);
    Q.end = Q.pos + need;
  }
  if (Q.end <= G.pairq_cap) {  // warp-uniform fast path: no per-entry bound check
    uint2* wp = G.pairq + (Q.pos + (inc - mine));
#pragma unroll
    for (int u = 0; u < SLOTS; ++u) {
      if (mask[u]) {
        const uint32_t t = tbuf[32 * u + lane];
        do {
          const uint32_t i = __ffs(mask[u]) - 1;
          mask[u] &= mask[u] - 1;
          __stcs(wp++, make_uint2(from_slot0 + i, t));  // streamed: the queue must not push the atom records out of L2
        } while (mask[u]);
      }
    }
  }
  Q.pos += total;
}

// ---- asynchronous staging of the next unit's candidates (cp.async: no registers are held while the copy is in flight) ----
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// candidates [chunk0, chunk0 + CHUNK) of the item whose record is spread over the lanes (word): records -> buf, slots -> tbuf.
// Lane l stages (and later reads) entries l, l + 32, ...: no cross-lane dependency.  Candidate e lives in sorted slot
// e + cb (run 0) or e + delta[r] (run r + 1, from candidate off[r] on).
__device__ __forceinline__ void stage_chunk(const PairArgs& G, uint32_t word, uint32_t chunk0, int lane, uint4* buf, uint32_t* tbuf) {
  const uint32_t cb = __shfl_sync(0xffffffffu, word, 0), total = __shfl_sync(0xffffffffu, word, 2);
  uint32_t off[4], delta[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    off[k] = __shfl_sync(0xffffffffu, word, 4 + k);
    delta[k] = __shfl_sync(0xffffffffu, word, 8 + k);
  }
#pragma unroll 1
  for (uint32_t e0 = chunk0; e0 < min(total, chunk0 + (uint32_t)CHUNK); e0 += 32u) {  // slots beyond the item are never read
    const uint32_t e = e0 + lane;
    uint32_t d = cb;
    d = e >= off[0] ? delta[0] : d;
    d = e >= off[1] ? delta[1] : d;
    d = e >= off[2] ? delta[2] : d;
    d = e >= off[3] ? delta[3] : d;
    const uint32_t t = e + d;
    tbuf[e - chunk0] = t;
    if (e < total)
      cp_async16(buf + (e - chunk0), G.recA + t);  // first half of the record: position + attr
    else
      buf[e - chunk0] = make_uint4(0x5d5e0b6bu, 0x5d5e0b6bu, 0x5d5e0b6bu, 0u);  // "no candidate": a point 1e18 A away
  }
  cp_async_commit();
}

// lane 0 draws the next item index; the value is broadcast only when it is needed (one iteration later).
// `cursor` is this lane's copy of &ctr->work read from shared memory: with an address it can prove warp-uniform, ptxas
// rewrites the atomic into its warp-aggregated form, whose result shuffle right behind the atomic stalled the warp for
// the whole round trip (12 % of the stall samples).
__device__ __forceinline__ uint32_t draw_cursor(unsigned int* cursor, int lane) {
  uint32_t w = 0;
  if (lane == 0) w = atomicAdd(cursor, 1u);
  return w;
}
// word `lane` of item w (lanes 16..31 and out-of-range items read nothing); total = 0 marks the end
__device__ __forceinline__ uint32_t fetch_item(const PairArgs& G, uint32_t w, uint32_t n_items, int lane) {
  uint32_t v = 0;
  if (w < n_items && lane < 16) v = reinterpret_cast<const uint32_t*>(G.items + w)[lane];
  return v;
}

// largest squared distance at which a label can fire for this atom with ANY partner (guard included)
__device__ __forceinline__ float atom_limit2(const DevParams& P, const float* lim1, uint32_t w) {
  float l = lim1[w & ARP_ATTR_ELEM_MASK];
  if (w & (ARP_ATTR_POS_CONF | ARP_ATTR_NEG_CONF)) l = fmaxf(l, 16.0f + P.guard);
  if (w & (ARP_ATTR_DONOR | ARP_ATTR_WEAK_DONOR | ARP_ATTR_ACCEPTOR)) l = fmaxf(l, P.polar_d2);
  if (w & ARP_ATTR_HYDROPHOBIC) l = fmaxf(l, 20.25f + P.guard);
  return l;
}

// ---- k_screen --------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PAIR_THREADS, PAIR_MIN_BLOCKS) k_screen(const PairArgs G, const DevParams P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  ScreenSmem& S = *reinterpret_cast<ScreenSmem*>(smem_raw);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x < 16) {  // per element class: the largest limit over all partner classes
    float m = 0.f;
    for (int eb = 0; eb < 16; ++eb) m = fmaxf(m, class_pair_limit2(P, threadIdx.x, eb));
    S.lim1[threadIdx.x] = m;
  }
  if (threadIdx.x < 32) S.cursor[threadIdx.x] = &G.ctr->work;
  __syncthreads();
  unsigned int* const cursor = S.cursor[lane];
  FromRec* sfrom = S.from[wid];
  unsigned long long n_in = 0, n_tests = 0;
  QWindow Q{0ull, 0ull};
  const uint32_t n_items = G.ctr->n_items;
  const float half_hit = 0.5f * P.scr_hit, half_cert = 0.5f * P.scr_cert;

  // ---- pipeline prologue ------------------------------------------------------------------------------------------
  uint32_t word = fetch_item(G, __shfl_sync(0xffffffffu, draw_cursor(cursor, lane), 0), n_items, lane);    // current item
  uint32_t word_n = fetch_item(G, __shfl_sync(0xffffffffu, draw_cursor(cursor, lane), 0), n_items, lane);  // next item
  uint32_t w_raw = draw_cursor(cursor, lane);                                                              // the one after
  uint32_t chunk0 = 0;
  int pb = 0;
  stage_chunk(G, word, 0, lane, S.buf[wid][0], S.tbuf[wid][0]);

  while (__shfl_sync(0xffffffffu, word, 2) != 0u) {
    const uint32_t cb = __shfl_sync(0xffffffffu, word, 0), n_own = __shfl_sync(0xffffffffu, word, 1),
                   total = __shfl_sync(0xffffffffu, word, 2);
    const float ccx = __uint_as_float(__shfl_sync(0xffffffffu, word, 12)), ccy = __uint_as_float(__shfl_sync(0xffffffffu, word, 13)),
                ccz = __uint_as_float(__shfl_sync(0xffffffffu, word, 14));
    const int ns = (int)min((uint32_t)SLOTS, (total - chunk0 + 31u) >> 5);
    // ---- this chunk's candidates: shared memory -> registers, cell-local frame --------------------------------------
    cp_async_wait_all();
    const uint4* buf = S.buf[wid][pb];
    const uint32_t* tbuf = S.tbuf[wid][pb];
    Cand C1;
    uint32_t eligm = 0;                // bit u: candidate u may start a row / is counted as `from`
    const uint4 own_raw = buf[lane];  // chunk 0: lane < n_own holds own atom `lane`
#pragma unroll
    for (int u = 0; u < SLOTS; ++u) {
      if (u < ns) {
        const uint4 r = buf[32 * u + lane];
        const float x = __uint_as_float(r.x) - ccx, y = __uint_as_float(r.y) - ccy, z = __uint_as_float(r.z) - ccz;
        C1.x[u] = x;
        C1.y[u] = y;
        C1.z[u] = z;
        C1.h[u] = -0.5f * fmaf(z, z, fmaf(y, y, x * x));  // "no candidate": h = -1.5e36 fails every test
        eligm |= (r.w >> 31) << u;
      } else {
        C1.x[u] = C1.y[u] = C1.z[u] = 0.f;
        C1.h[u] = -1.0e30f;
      }
    }
#if SCREEN_F32X2
    CandP C;
#pragma unroll
    for (int p = 0; p < SLOTS / 2; ++p) {
      C.x[p] = pack2(C1.x[2 * p], C1.x[2 * p + 1]);
      C.y[p] = pack2(C1.y[2 * p], C1.y[2 * p + 1]);
      C.z[p] = pack2(C1.z[2 * p], C1.z[2 * p + 1]);
      C.h[p] = pack2(C1.h[2 * p], C1.h[2 * p + 1]);
    }
#else
    const Cand& C = C1;
#endif
    // ---- stage the next unit: the next chunk of this item, or chunk 0 of the next item ------------------------------
    const bool more = chunk0 + CHUNK < total;
    uint32_t word_nn = word_n, w_raw_n = w_raw;
    if (more) {
      stage_chunk(G, word, chunk0 + CHUNK, lane, S.buf[wid][pb ^ 1], S.tbuf[wid][pb ^ 1]);
    } else {
      stage_chunk(G, word_n, 0, lane, S.buf[wid][pb ^ 1], S.tbuf[wid][pb ^ 1]);
      word_nn = fetch_item(G, __shfl_sync(0xffffffffu, w_raw, 0), n_items, lane);
      w_raw_n = draw_cursor(cursor, lane);
    }
    // distance tests this item stands for: own-cell pairs once, every other candidate against every own atom
    if (chunk0 == 0) n_tests += (unsigned long long)n_own * (total - n_own) + (unsigned long long)n_own * (n_own - 1) / 2;

    // ---- from-atoms in passes of 32 (one pass for every realistic cell) ------------------------------------------------
    for (uint32_t fbo = 0; fbo < n_own; fbo += FROM_PASS) {
      const int nf = (int)min((uint32_t)FROM_PASS, n_own - fbo);
      __syncwarp();
      if (lane < nf && (chunk0 == 0 || n_own > (uint32_t)FROM_PASS)) {  // a single pass stays staged for the later chunks
        uint4 fr = own_raw;
        if (chunk0 != 0 || fbo != 0) fr = *reinterpret_cast<const uint4*>(G.recA + cb + fbo + lane);
        FromRec f;
        f.x = __uint_as_float(fr.x) - ccx;
        f.y = __uint_as_float(fr.y) - ccy;
        f.z = __uint_as_float(fr.z) - ccz;
#if SCREEN_F32X2
        f.x2 = f.x, f.y2 = f.y, f.z2 = f.z;
#endif
        const float hf = 0.5f * fmaf(f.z, f.z, fmaf(f.y, f.y, f.x * f.x));
        // no label can fire for this atom beyond its own limit: a per-atom T_near (conservative; B1 applies the pair limit)
        const float t_near = P.count_classified ? P.scr_near : fminf(P.scr_near, atom_limit2(P, S.lim1, fr.w) + P.guard_a);
        f.k_hit = hf - half_hit;
        f.k_cert = hf - half_cert;
        f.k_near = hf - 0.5f * t_near;
        f.w = ((fr.w >> 31) << 16) | 1u;
#if SCREEN_F32X2
        f.pad[0] = f.pad[1] = 0u;
#else
        f.pad = 0u;
#endif
        sfrom[lane] = f;
      }
      __syncwarp();
      uint32_t lim[SLOTS], mask[SLOTS], cnt[SLOTS];
#pragma unroll
      for (int u = 0; u < SLOTS; ++u) {
        const uint32_t e = chunk0 + 32u * u + lane;
        lim[u] = e < n_own ? (e > fbo ? e - fbo : 0u) : 0xffffffffu;
        mask[u] = cnt[u] = 0u;
      }
      // own-cell atoms beyond slot 0 (more than 32 atoms in a cell) are rare: one variant, all slots
      switch ((chunk0 < n_own && chunk0 + 32u < n_own) ? 0 : ns) {
        case 0: screen_pass<true, SLOTS>(sfrom, nf, C, lim, mask, cnt); break;
        case 1: screen_pass<false, 1>(sfrom, nf, C, lim, mask, cnt); break;
        case 2: screen_pass<false, 2>(sfrom, nf, C, lim, mask, cnt); break;
        case 3: screen_pass<false, 3>(sfrom, nf, C, lim, mask, cnt); break;
        case 4: screen_pass<false, 4>(sfrom, nf, C, lim, mask, cnt); break;
        case 5: screen_pass<false, 5>(sfrom, nf, C, lim, mask, cnt); break;
        default: screen_pass<false, 6>(sfrom, nf, C, lim, mask, cnt); break;
      }
      if (P.count_in_cutoff) {
        uint32_t c = 0;
#pragma unroll
        for (int u = 0; u < SLOTS; ++u) c += (cnt[u] >> 16) + (cnt[u] & 0xffffu) * ((eligm >> u) & 1u);
        n_in += c;
      }
      expand_masks(G, mask, tbuf, cb + fbo, Q, lane);
    }

    // ---- advance ------------------------------------------------------------------------------------------------------
    pb ^= 1;
    if (more) {
      chunk0 += CHUNK;
    } else {
      chunk0 = 0;
      word = word_n;
      word_n = word_nn;
      w_raw = w_raw_n;
    }
  }
  cp_async_wait_all();
  qwindow_close(G, Q, lane);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n_in += __shfl_xor_sync(0xffffffffu, n_in, o);
  if (lane == 0) {
    atomicAdd(&G.ctr->screened, n_tests);
    if (n_in) atomicAdd(&G.ctr->in_cutoff, n_in);
  }
}

// ---- k_screen_atoms: the screen at the emit radius (ARP_FLAG_EMIT_RADIUS_SCREEN), one LANE per from-atom -------------------
// With cells at r_emit = 4.5 A instead of dist_cutoff a half shell holds ~40 candidates and a cell ~4 atoms: the
// warp-per-cell screen above then spends nine tenths of its instructions on per-cell bookkeeping (measured: 0.195 ms for
// 47 M tests against 0.155 ms for 119 M at dist_cutoff).  Here a warp takes 32 consecutive sorted slots (atoms of 6-9
// neighbouring cells), each lane walks the five runs of ITS cell's half shell -- run 0 from its own slot + 1, so every
// unordered pair is tested once.  A lane finds its runs through its slot's cell id and boundary bits (SlotCell) and the
// cell table itself; no per-cell work item is built or read in this mode.
// Candidates are read two at a time from the pair-interleaved array pos2 (x0 x1 | y0 y1 | z0 z1 | t0 t1, 32 bytes per slot
// pair, written by k_scatter_atoms; t = residue tag): two 16-byte loads through L1 (lanes of a cell read the same
// addresses: broadcasts) feed packed f32x2 arithmetic -- 3 add, 1 mul, 2 fma for TWO tests, the same operations in the same
// order as k_filter's difference form, so the same value -- one compare each against the from-atom's own largest label
// limit, and the in-chain neighbour rule of complex.rs:112-113 decided right there from the tags (|tag difference| <= 1
// <=> same chain, residue ordinals at most 1 apart: no orientation of should_compare_entities can pass; two hits in
// three end here, with no gather).  A hit appends (candidate slot, d2) to a per-lane list in shared memory (a predicated
// store and add: no vote, no atomic in the loop).  Every SA_ROUND trips the warp looks at the lists; a flush runs stage
// B1 on them -- per-pair label limit, both orientations of the pair filter (one 16-byte gather per hit) -- and sends the
// survivors, with the input indices the rows will carry, to the queue of k_exact 32 at a time (one atomicAdd per 32).
// r2c: the flush reads the 32 lists as ONE list, 32 consecutive hits per row (SA_DENSE_FLUSH): lists are 0-5 entries long,
// so "entry j of every lane" kept under half of a row's lanes busy through the ~60 instructions of stage B1; now every
// lane of every row but the last works (k_screen_atoms 0.138 -> 0.121 ms on config 3, 0.116 with rounds of two trips).
constexpr int SA_THREADS = 256;
constexpr int SA_WARPS = SA_THREADS / 32;
constexpr uint32_t SA_WINDOW = 512;
#ifndef SA_BUF_N
#define SA_BUF_N 12
#endif
#ifndef SA_ROUND_N
#define SA_ROUND_N 2  // r2c: with the dense flush, fewer and fuller flushes win over longer straight-line rounds (4: 0.121 ms, 2: 0.116 ms)
#endif
constexpr int SA_BUF = SA_BUF_N;      // entries of a lane's hit list
constexpr int SA_ROUND = SA_ROUND_N;  // trips between two looks at the lists (a trip appends at most 2)
static_assert(SA_BUF > 2 * SA_ROUND, "a round must fit the list");
#ifndef SA_MIN_BLOCKS
#define SA_MIN_BLOCKS 4
#endif
#ifndef SA_TAG_IN_TRIP
#define SA_TAG_IN_TRIP 1  // 1: the in-chain rule is decided in the trip (tags ride in the position records: 32 bytes per trip);
#endif                    // 0: in the flush, with one 4-byte gather per hit (24 bytes per trip, three times as many hits)
#ifndef SA_LD256
#define SA_LD256 1  // the candidate pair comes in with one 32-byte load
#endif
#ifndef SA_FLUSH_DEPTH
#define SA_FLUSH_DEPTH 2  // hit-list rows (= gathers per lane) in flight in a flush
#endif

__device__ __forceinline__ unsigned long long sa_add2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long sa_mul2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long sa_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ unsigned long long sa_dup(float a) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %1};" : "=l"(r) : "f"(a));
  return r;
}

constexpr int SA_RUNS = 6;  // rows of a thread's run table: up to 5 non-empty runs + the closing row
struct SaSmem {
  float lim1[16];
  float lim2[256];                                                   // per element-class pair: see k_filter
  uint32_t run_b[SA_RUNS][SA_THREADS], run_e[SA_RUNS][SA_THREADS];  // every thread's non-empty runs, then the closing row
  uint2 hits[SA_BUF][SA_THREADS];                                   // per-lane hit lists: (candidate slot, d2 bits)
  uint4 stash2[SA_WARPS][64];                                       // pairs for k_exact waiting for a full 32-entry flush
};
// a warp's carry-over between flushes, tiles and the final drain
struct SaCarry {
  int n2;
  unsigned long long n_q;  // hits that passed the in-chain rule (statistics: arp_stats.n_screened_pairs)
};

// the warp's second stash -> the queue of k_exact, 32 entries per atomicAdd (drain: whatever is left)
__device__ __forceinline__ void sa_emit(const PairArgs& G, SaSmem& S, int wid, int lane, SaCarry& K, bool drain) {
  uint4* s2 = S.stash2[wid];
  while (K.n2 >= 32 || (drain && K.n2 > 0)) {
    const int k = min(K.n2, 32);
    unsigned long long w = 0;
    if (lane == 0) w = atomicAdd(&G.ctr->q2_len, (unsigned long long)k);
    w = __shfl_sync(0xffffffffu, w, 0) + lane;
    K.n2 -= k;
    ARP_CHK(G.ctr, K.n2 >= 0 && K.n2 + k <= 64, 1);  // the stash holds 64 entries
    if (lane < k && w < G.pairq2_cap) __stcs(G.pairq2 + w, s2[K.n2 + lane]);
    __syncwarp();
  }
}

// Stage B1 on the hit lists of a warp, entry j of every lane at a time (SA_FLUSH_DEPTH gathers in flight): per-pair label
// limit (no label of complex.rs:217-296 fires beyond it), both orientations of should_compare_entities
// (complex.rs:76-131); the survivors of a row are ranked by a ballot into the warp's stash.  `mine` = the lane's own RecK.
#ifndef SA_DENSE_FLUSH
#define SA_DENSE_FLUSH 1  // 1: the lanes' lists are read as ONE list, 32 consecutive hits per row (every lane busy);
#endif                    // 0: entry j of every lane per row (lanes with shorter lists idle: ~45 % of a row works)
#if SA_DENSE_FLUSH
// Hit g of the warp (lists concatenated in lane order) belongs to the last lane whose list starts at or before g; a lane
// finds it by a 5-step search over the lanes' start offsets, reads the entry from that lane's column of the hit lists
// and fetches the from-atom's record from that lane's registers.
__device__ __forceinline__ void sa_flush(const PairArgs& G, const DevParams& P, SaSmem& S, const unsigned long long* __restrict__ pos2, int tid,
                                         int lane, uint32_t slot, const RecK mine, uint32_t tf1, uint32_t cnt, SaCarry& K) {
  (void)pos2, (void)tf1;
  static_assert(SA_TAG_IN_TRIP, "the dense flush reads lists that already passed the in-chain rule");
  const uint32_t lt = (1u << lane) - 1u;
  const int wid = tid >> 5, t0 = tid & ~31;
  uint4* s2 = S.stash2[wid];
  uint32_t inc = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, inc, 31), off = inc - cnt;
  K.n_q += cnt;
  int n2 = K.n2;
  for (uint32_t g0 = 0; g0 < total; g0 += 32u) {
    const uint32_t g = g0 + lane;
    const bool have = g < total;
    uint32_t src = 0;
#pragma unroll
    for (int st = 16; st > 0; st >>= 1) {
      const uint32_t probe = src + st;  // <= 31
      const uint32_t v = __shfl_sync(0xffffffffu, off, probe);
      if (v <= g) src = probe;
    }
    // (lanes past the end of the list land on some lane as well: their entry is not read)
    const uint32_t o_src = __shfl_sync(0xffffffffu, off, src);
    const uint32_t a_attr = __shfl_sync(0xffffffffu, mine.attr, src), a_ck = __shfl_sync(0xffffffffu, mine.chain_key, src);
    const uint32_t a_ro = __shfl_sync(0xffffffffu, mine.res_ordinal, src), a_perm = __shfl_sync(0xffffffffu, mine.perm, src);
    const uint32_t a_slot = slot - (uint32_t)lane + src;
    uint2 h = make_uint2(0u, 0x7f800000u);  // no hit: d2 = +inf fails every limit
    uint4 kb = make_uint4(0u, 0u, 0u, 0u);
    if (have) {
      const uint32_t j = g - o_src;
      ARP_CHK(G.ctr, j < (uint32_t)SA_BUF, 2);  // a lane's list never outgrows its SA_BUF entries
      h = S.hits[j][t0 + (int)src];
      ARP_CHK(G.ctr, h.x < G.n_atoms && h.x != a_slot, 3);  // a hit is another atom's sorted slot
      kb = *reinterpret_cast<const uint4*>(G.recK + h.x);
    }
    const float d2 = __uint_as_float(h.y);
    const bool pot = d2 <= label_limit2(P, S.lim2, a_attr, kb.x);
    const int dir = pot ? pair_direction(a_attr, a_ck, a_ro, kb.x, kb.y, kb.z) : 0;
    const bool go = dir != 0;
    const uint32_t m = __ballot_sync(0xffffffffu, go);
    ARP_CHK(G.ctr, n2 + __popc(m) <= 64, 4);
    if (go) s2[n2 + __popc(m & lt)] = make_uint4(a_slot | QF_B, h.x | (dir == 2 ? QF_A : 0u) | QF_B, a_perm, kb.w);
    n2 += __popc(m);
    __syncwarp();
    if (n2 >= 32) {  // the stash holds 64: never more than 31 + 32 before this point
      K.n2 = n2;
      sa_emit(G, S, wid, lane, K, false);
      n2 = K.n2;
    }
  }
  K.n2 = n2;
  __syncwarp();  // the lists are rewritten by their own lanes next: every lane has read what it needed
}
#else
__device__ __forceinline__ void sa_flush(const PairArgs& G, const DevParams& P, SaSmem& S, const unsigned long long* __restrict__ pos2, int tid,
                                         int lane, uint32_t slot, const RecK mine, uint32_t tf1, uint32_t cnt, SaCarry& K) {
  const uint32_t maxc = __reduce_max_sync(0xffffffffu, cnt);
  const uint32_t lt = (1u << lane) - 1u;
  const int wid = tid >> 5;
  uint4* s2 = S.stash2[wid];
#if SA_TAG_IN_TRIP
  K.n_q += cnt;
#endif
  for (uint32_t j0 = 0; j0 < maxc; j0 += SA_FLUSH_DEPTH) {
    uint2 h[SA_FLUSH_DEPTH];
    uint4 kb[SA_FLUSH_DEPTH];
#pragma unroll
    for (int u = 0; u < SA_FLUSH_DEPTH; ++u) {
      h[u] = make_uint2(0u, 0x7f800000u);  // no hit: d2 = +inf fails every limit
      kb[u] = make_uint4(0u, 0u, 0u, 0u);
      if (j0 + u < cnt) {
        ARP_CHK(G.ctr, j0 + u < (uint32_t)SA_BUF, 2);  // a lane's list never outgrows its SA_BUF entries
        h[u] = S.hits[j0 + u][tid];
        ARP_CHK(G.ctr, h[u].x < G.n_atoms && h[u].x != slot, 3);  // a hit is another atom's sorted slot
        kb[u] = *reinterpret_cast<const uint4*>(G.recK + h[u].x);
#if !SA_TAG_IN_TRIP
        const uint32_t tc = reinterpret_cast<const uint32_t*>(pos2)[8u * (size_t)(h[u].x >> 1) + 6u + (h[u].x & 1u)];
        if (tc - tf1 <= 2u) h[u].y = 0x7f800000u;  // same chain, residue ordinals at most 1 apart: no orientation can pass
        else K.n_q += 1u;
#endif
      }
    }
    int n2 = K.n2;
#pragma unroll
    for (int u = 0; u < SA_FLUSH_DEPTH; ++u) {
      if (j0 + u >= maxc) break;  // warp-uniform
      const float d2 = __uint_as_float(h[u].y);
      const bool pot = d2 <= label_limit2(P, S.lim2, mine.attr, kb[u].x);
      const int dir = pot ? pair_direction(mine.attr, mine.chain_key, mine.res_ordinal, kb[u].x, kb[u].y, kb[u].z) : 0;
      const bool go = dir != 0;
      const uint32_t m = __ballot_sync(0xffffffffu, go);
      ARP_CHK(G.ctr, n2 + __popc(m) <= 64, 4);
      if (go) s2[n2 + __popc(m & lt)] = make_uint4(slot | QF_B, h[u].x | (dir == 2 ? QF_A : 0u) | QF_B, mine.perm, kb[u].w);
      n2 += __popc(m);
      __syncwarp();
      if (n2 >= 32) {  // the stash holds 64: never more than 31 + 32 before this point
        K.n2 = n2;
        sa_emit(G, S, wid, lane, K, false);
        n2 = K.n2;
      }
    }
    K.n2 = n2;
  }
}

#endif

// The trips of one tile.  A lane is active while s < re.  The trips are branch-free (a lane that has finished keeps
// computing on slot pair 0 and appends nothing), so a round of SA_ROUND trips is straight-line code; the candidates of
// trip t + 1 are loaded before trip t computes.  Candidates come straight from pos2 through L1 (r2b: staging the union of
// the warp's five candidate spans in shared memory first -- cp.async, 4 KB per warp, trips on ld.shared -- was built and
// measured: 0.185 ms against 0.138 ms; the kernel is bound by issue slots, not by L1, and the staging adds instructions).
__device__ __forceinline__ void sa_trips(const PairArgs& G, const DevParams& P, SaSmem& S, const unsigned long long* __restrict__ pos2,
                                         int tid, int lane, uint32_t slot, const RecK mine, unsigned long long nax, unsigned long long nay,
                                         unsigned long long naz, float tlim, uint32_t tf1, uint32_t s, uint32_t rb, uint32_t re, uint32_t rp,
                                         SaCarry& K) {
  constexpr uint32_t ROW = 4u * SA_THREADS, B_TO_E = 4u * SA_THREADS * SA_RUNS;  // byte strides of the run tables
  constexpr uint32_t HROW = 8u * SA_THREADS;                                      // byte stride of the hit lists
  const uint32_t hp0 = (uint32_t)__cvta_generic_to_shared(&S.hits[0][tid]);
  auto load = [&](uint32_t ld, ulonglong2& XY, ulonglong2& ZT) {
    // slot pair (ld, ld + 1): 32 bytes, (x0 x1 | y0 y1) and (z0 z1 | t0 t1)
#if SA_TAG_IN_TRIP && SA_LD256
    // one 256-bit load (LDG.E.256, sm_100): half the load instructions and L1 requests of two 16-byte loads
    asm("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(XY.x), "=l"(XY.y), "=l"(ZT.x), "=l"(ZT.y) : "l"(pos2 + 2u * ld));
    return;
#endif
    XY = *reinterpret_cast<const ulonglong2*>(pos2 + 2u * ld);
#if SA_TAG_IN_TRIP
    ZT = *reinterpret_cast<const ulonglong2*>(pos2 + 2u * ld + 2u);
#else
    ZT.x = pos2[2u * ld + 2u];
    ZT.y = 0ull;
#endif
  };
  uint32_t ld = s < re ? s : 0u;
  ulonglong2 XY, ZT;
  load(ld, XY, ZT);
  uint32_t hp = hp0;  // shared-memory address of the next free entry of this lane's hit list
  while (__any_sync(0xffffffffu, s < re)) {
#pragma unroll
    for (int trip = 0; trip < SA_ROUND; ++trip) {
      // where the next trip reads: two slots on, or the start of the next run
      uint32_t sn = s + 2u, rbn = rb, ren = re;
      asm volatile("{\n\t.reg .pred p, q;\n\t"
                   "setp.lt.u32 q, %6, %2;\n\t"
                   "setp.ge.and.u32 p, %0, %2, q;\n\t"
                   "@p ld.shared.u32 %1, [%3];\n\t"
                   "@p ld.shared.u32 %2, [%3+%4];\n\t"
                   "@p add.u32 %3, %3, %5;\n\t"
                   "@p and.b32 %0, %1, 0xfffffffe;\n\t}"
                   : "+r"(sn), "+r"(rbn), "+r"(ren), "+r"(rp) : "n"(B_TO_E), "n"(ROW), "r"(s) : "memory");
      ld = sn < ren ? sn : 0u;
      ulonglong2 XYn, ZTn;
      load(ld, XYn, ZTn);
      const unsigned long long dx = sa_add2(XY.x, nax), dy = sa_add2(XY.y, nay), dz = sa_add2(ZT.x, naz);
      const unsigned long long d2 = sa_fma2(dz, dz, sa_fma2(dy, dy, sa_mul2(dx, dx)));
      uint32_t d0, d1, t0, t1;
      asm("mov.b64 {%0, %1}, %2;" : "=r"(d0), "=r"(d1) : "l"(d2));
      asm("mov.b64 {%0, %1}, %2;" : "=r"(t0), "=r"(t1) : "l"(ZT.y));
      const uint32_t s1 = s + 1u;
      // predicated store + add (written in PTX: the compiler otherwise keeps a count and rebuilds the address).
      // hit = within the from-atom's label limit, inside the run, and NOT (same chain, residue ordinals <= 1 apart)
#if SA_TAG_IN_TRIP
      asm volatile("{\n\t.reg .pred p;\n\t.reg .u32 t;\n\t"
                   "setp.le.f32 p, %1, %2;\n\t"
                   "setp.ge.and.u32 p, %3, %4, p;\n\t"
                   "sub.u32 t, %6, %7;\n\t"
                   "setp.gt.and.u32 p, t, 2, p;\n\t"
                   "@p st.shared.v2.u32 [%0], {%3, %8};\n\t"
                   "@p add.u32 %0, %0, %5;\n\t}"
                   : "+r"(hp) : "f"(__uint_as_float(d0)), "f"(tlim), "r"(s), "r"(rb), "n"(HROW), "r"(t0), "r"(tf1), "r"(d0) : "memory");
      asm volatile("{\n\t.reg .pred p;\n\t.reg .u32 t;\n\t"
                   "setp.le.f32 p, %1, %2;\n\t"
                   "setp.lt.and.u32 p, %3, %4, p;\n\t"
                   "sub.u32 t, %6, %7;\n\t"
                   "setp.gt.and.u32 p, t, 2, p;\n\t"
                   "@p st.shared.v2.u32 [%0], {%3, %8};\n\t"
                   "@p add.u32 %0, %0, %5;\n\t}"
                   : "+r"(hp) : "f"(__uint_as_float(d1)), "f"(tlim), "r"(s1), "r"(re), "n"(HROW), "r"(t1), "r"(tf1), "r"(d1) : "memory");
#else
      asm volatile("{\n\t.reg .pred p;\n\t"
                   "setp.le.f32 p, %1, %2;\n\t"
                   "setp.ge.and.u32 p, %3, %4, p;\n\t"
                   "@p st.shared.v2.u32 [%0], {%3, %6};\n\t"
                   "@p add.u32 %0, %0, %5;\n\t}"
                   : "+r"(hp) : "f"(__uint_as_float(d0)), "f"(tlim), "r"(s), "r"(rb), "n"(HROW), "r"(d0) : "memory");
      asm volatile("{\n\t.reg .pred p;\n\t"
                   "setp.le.f32 p, %1, %2;\n\t"
                   "setp.lt.and.u32 p, %3, %4, p;\n\t"
                   "@p st.shared.v2.u32 [%0], {%3, %6};\n\t"
                   "@p add.u32 %0, %0, %5;\n\t}"
                   : "+r"(hp) : "f"(__uint_as_float(d1)), "f"(tlim), "r"(s1), "r"(re), "n"(HROW), "r"(d1) : "memory");
      (void)t0, (void)t1;
#endif
      s = sn, rb = rbn, re = ren, XY = XYn, ZT = ZTn;
    }
    ARP_CHK(G.ctr, hp <= hp0 + (uint32_t)SA_BUF * HROW, 9);  // a round appends at most 2 * SA_ROUND entries to a list that had room for them
    if (__any_sync(0xffffffffu, hp > hp0 + (uint32_t)(SA_BUF - 2 * SA_ROUND) * HROW)) {
      sa_flush(G, P, S, pos2, tid, lane, slot, mine, tf1, (hp - hp0) / HROW, K);
      hp = hp0;
    }
  }
  if (__any_sync(0xffffffffu, hp != hp0)) sa_flush(G, P, S, pos2, tid, lane, slot, mine, tf1, (hp - hp0) / HROW, K);
}

__global__ void __launch_bounds__(SA_THREADS, SA_MIN_BLOCKS) k_screen_atoms(const PairArgs G, const DevParams P,
                                                                            const unsigned long long* __restrict__ pos2, uint32_t n_atoms) {
  // dynamic shared memory (sizeof(SaSmem), above the 48 KB a static array may take: api_entry.inc raises the kernel's limit)
  extern __shared__ __align__(16) unsigned char sa_smem_raw[];
  SaSmem& S = *reinterpret_cast<SaSmem*>(sa_smem_raw);
  const int lane = threadIdx.x & 31, tid = threadIdx.x, wid = threadIdx.x >> 5;
  if (tid < 16) {
    float m = 0.f;
    for (int eb = 0; eb < 16; ++eb) m = fmaxf(m, class_pair_limit2(P, tid, eb));
    S.lim1[tid] = m;
  }
  for (int k = tid; k < 256; k += SA_THREADS) S.lim2[k] = class_pair_limit2(P, k >> 4, k & 15);
  __syncthreads();
  const uint32_t n_tiles = (n_atoms + 31u) >> 5;
  const uint32_t n_warps = gridDim.x * SA_WARPS;
  // single-structure jobs: the grid's row pitches are the same for every lane
  uint32_t nx0 = 0, nxny0 = 0;
  if (!G.atom_struct) {
    const StructDesc& d = G.descs[0];
    nx0 = (uint32_t)d.nx, nxny0 = (uint32_t)d.nx * (uint32_t)d.ny;
  }
  SaCarry K{0, 0ull};
  unsigned long long n_tests = 0;
  const uint32_t rp0 = (uint32_t)__cvta_generic_to_shared(&S.run_b[0][tid]);
  constexpr uint32_t ROW = 4u * SA_THREADS;
  const uint32_t* __restrict__ cs = G.cell_start;
  for (uint32_t tile = blockIdx.x * SA_WARPS + wid; tile < n_tiles; tile += n_warps) {
    const uint32_t slot = (tile << 5) + lane;
    uint32_t tf1 = 0;
    RecK mine{0u, 0u, 0u, 0u};
    unsigned long long nax = 0, nay = 0, naz = 0;  // (-x, -x), (-y, -y), (-z, -z) of the from-atom
    float tlim = -1.f;
    // the five runs of this lane's half shell as sorted slots [b, e): rows (dy, dz) = (+1,0) (-1,+1) (0,+1) (+1,+1) with x
    // from ix-1 to ix+1, then the rest of the own cell and its +x neighbour from the lane's own slot + 1
    uint32_t b5[5], e5[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) b5[k] = 0xffffffffu, e5[k] = 0u;
    if (slot < n_atoms) {
      mine = G.recK[slot];
      const SlotCell sc = G.slot_cell[slot];
      const float* q = reinterpret_cast<const float*>(pos2) + 8u * (size_t)(slot >> 1) + (slot & 1u);
      nax = sa_dup(-q[0]), nay = sa_dup(-q[2]), naz = sa_dup(-q[4]);
      tf1 = __float_as_uint(q[6]) - 1u;  // |tag - own tag| <= 1  <=>  tag - (own tag - 1) <= 2 (unsigned)
      tlim = fminf(P.thr_near, atom_limit2(P, S.lim1, mine.attr));  // no label can fire for this atom beyond its own limit
      uint32_t nx = nx0, nxny = nxny0;
      if (G.atom_struct) {  // batch jobs: every structure has its own grid
        const StructDesc& d = G.descs[G.atom_struct[mine.perm]];
        nx = (uint32_t)d.nx, nxny = (uint32_t)d.nx * (uint32_t)d.ny;
      }
      const uint32_t c = sc.cell, xl = sc.bits & 1u, xh = (sc.bits >> 1) & 1u;
      const bool yl = sc.bits & 4u, yh = sc.bits & 8u, zh = sc.bits & 16u;
      const uint32_t rc4[4] = {c + nx, c - nx + nxny, c + nxny, c + nx + nxny};
      const bool ok4[4] = {yh, yl && zh, zh, yh && zh};
      ARP_CHK(G.ctr, c < G.n_cells && c + 1u + xh <= G.n_cells, 5);  // the slot's cell lies in the cell table
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ok4[k]) {
          ARP_CHK(G.ctr, rc4[k] >= xl && rc4[k] + xh + 1u <= G.n_cells, 6);  // and so do the rows of its half shell
          const uint32_t b = cs[rc4[k] - xl], e = cs[rc4[k] + xh + 1u];
          ARP_CHK(G.ctr, b <= e && e <= n_atoms, 7);  // runs are slot ranges
          if (b < e) b5[k] = b, e5[k] = e;
        }
      {
        const uint32_t e = cs[c + 1u + xh];
        if (slot + 1u < e) b5[4] = slot + 1u, e5[4] = e;
      }
    }
    // ---- this lane's run table: rows 0..3 first (the lanes of a cell walk them in step), the own row last -----------------
    uint32_t n = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      if (b5[k] < e5[k]) {
        S.run_b[n][tid] = b5[k], S.run_e[n][tid] = e5[k];
        ++n;
        n_tests += e5[k] - b5[k];
      }
    }
#ifdef ARP_CHECK_SELFTEST  // negative control of the checked build: a bound that real inputs violate must fail the job
    ARP_CHK(G.ctr, n < 3u, 8);
#else
    ARP_CHK(G.ctr, n < (uint32_t)SA_RUNS, 8);  // five runs at most, then the closing row
#endif
    S.run_b[n][tid] = 0xffffffffu, S.run_e[n][tid] = 0u;  // the closing row: no s reaches rb or stays below re (s keeps stepping)
    uint32_t rb = 0xffffffffu, re = 0u, s = 0u;
    if (n) rb = S.run_b[0][tid], re = S.run_e[0][tid], s = rb & ~1u;
    sa_trips(G, P, S, pos2, tid, lane, slot, mine, nax, nay, naz, tlim, tf1, s, rb, re, rp0 + ROW, K);
  }
  sa_emit(G, S, wid, lane, K, true);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_tests += __shfl_xor_sync(0xffffffffu, n_tests, o);
    K.n_q += __shfl_xor_sync(0xffffffffu, K.n_q, o);
  }
  if (lane == 0 && K.n_q) atomicAdd(&G.ctr->q_len, K.n_q);
  if (lane == 0 && n_tests) atomicAdd(&G.ctr->screened, n_tests);
}

}  // namespace arpd
