// The 20-column frame on the device (sm_100a): row keys, LSD radix sort, column decode + side-chain statistics.
//   k_frame_keys     <->  the sort keys of get_contacts                         src/contacts/mod.rs:123-134
//   k_rsort_*        <->  DataFrame::sort (10 keys, ascending)                   src/contacts/mod.rs:133
//   k_frame_columns  <->  results_to_df                                         src/contacts/mod.rs:140-181
//                         collect_sc_stats / sc_results_to_df / left join       src/contacts/complex.rs:137-174, mod.rs:111-122, 184-214
// All ten sort keys are functions of the two row ends and the label, and every string key orders like a dense integer
// rank the caller prepared (arp_frame_tables), so a row's position is decided by ONE packed key of
//   [chain_rank(from) | chain_rank(to) | ent_rank(from) | ent_rank(to) | label_rank | tie(from) | tie(to)]
// (most significant first; model is folded into chain_rank, both ends of a row share the model).  The widths come from
// the table's bounds, typically 50-60 bits in total; the sort runs ceil(bits / 8) counting passes and skips the upper
// key word when everything fits 64 bits.
#pragma once
#include "dev_common.cuh"

namespace arpd {

// A tile (32 lanes x `rounds` elements) is walked in order by ONE warp: that is what makes a pass stable without any
// block-level synchronisation.  `rounds` grows with the input (host: rs_rounds) so that the digit-major histogram and
// the scattered runs stay coarse when the rows no longer fit L2.
constexpr int RS_WARPS = 8, RS_THREADS = 32 * RS_WARPS;

struct KeyLayout {  // bit position of the least significant bit of every field, widths; total <= 128
  int sh_tie_t, sh_tie_f, sh_label, sh_rank_t, sh_rank_f, sh_chain_t, sh_chain_f, bits;
};

struct FrameArgs {
  const arp_row* rows;
  uint32_t n_rows, n_atoms, n_ents, n_sides;
  const arp_frame_side* sides;
  const double* side_sc;
  const uint32_t* ent_side;
  const int32_t* ent_atomn;
  const int32_t* ent_atomi;
  const uint32_t* ent_rank;
  const uint32_t* ent_tie;
  uint8_t label_rank[32];
  KeyLayout L;
  Counters* ctr;
};

struct FrameOut {  // device pointers into one block, same layout as the pinned host block
  uint32_t* model;  // NULL (with col[]) in the compact frame: the fifteen per-entity columns stay on the host's side
  uint8_t* interaction;
  float* distance;
  int32_t* col[14];
  float *scd, *sch, *sca;
  uint8_t* sc_valid;
  uint32_t *from, *to;
};

__device__ __forceinline__ uint32_t entity_of(uint32_t v, uint32_t n_atoms) { return (v & ARP_RING_FLAG) ? n_atoms + (v & ~ARP_RING_FLAG) : v; }

__device__ __forceinline__ void key_put(unsigned long long& lo, unsigned long long& hi, unsigned long long v, int sh) {
  if (sh < 64) {
    lo |= v << sh;
    if (sh > 0) hi |= v >> (64 - sh);  // the part of the field that crosses into the upper word (v < 2^32)
  } else {
    hi |= v << (sh - 64);
  }
}

// error bits raised here: 8 = a row references an entity that cannot be narrowed (ARP_ERR_NARROWING), 16 = a row end
// outside the tables
__global__ void __launch_bounds__(256) k_frame_keys(const FrameArgs A, unsigned long long* __restrict__ klo, unsigned long long* __restrict__ khi,
                                                    uint32_t* __restrict__ idx) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n_rows) return;
  const arp_row r = A.rows[i];
  const uint32_t ef = entity_of(r.from, A.n_atoms), et = entity_of(r.to, A.n_atoms);
  unsigned long long lo = 0ull, hi = 0ull;
  if (ef >= A.n_ents || et >= A.n_ents || r.label >= (uint32_t)ARP_N_LABELS) {
    atomicOr(&A.ctr->error, 16u);
  } else {
    const uint32_t sf = A.ent_side[ef], st = A.ent_side[et];
    const uint32_t tf = A.ent_tie[ef], tt = A.ent_tie[et];
    if (sf >= A.n_sides || st >= A.n_sides) {
      atomicOr(&A.ctr->error, 16u);
    } else {
      if ((tf | tt) & ARP_ENT_NARROW) atomicOr(&A.ctr->error, 8u);
      key_put(lo, hi, tt & ~ARP_ENT_NARROW, A.L.sh_tie_t);
      key_put(lo, hi, tf & ~ARP_ENT_NARROW, A.L.sh_tie_f);
      key_put(lo, hi, A.label_rank[r.label], A.L.sh_label);
      key_put(lo, hi, A.ent_rank[et], A.L.sh_rank_t);
      key_put(lo, hi, A.ent_rank[ef], A.L.sh_rank_f);
      key_put(lo, hi, A.sides[st].chain_rank, A.L.sh_chain_t);
      key_put(lo, hi, A.sides[sf].chain_rank, A.L.sh_chain_f);
    }
  }
  klo[i] = lo;
  if (khi) khi[i] = hi;
  idx[i] = i;
}

__device__ __forceinline__ uint32_t rs_digit(unsigned long long lo, unsigned long long hi, int shift) {
  return (uint32_t)((shift < 64 ? lo >> shift : hi >> (shift - 64)) & 255ull);  // shift is a multiple of 8: never straddles
}

// One counting pass = k_rsort_hist -> exclusive scan of hist (k_scan_lookback) -> k_rsort_scatter.
// hist is digit-major: hist[d * n_tiles + tile], so its flat exclusive scan is the first output slot of (digit, tile).
__global__ void __launch_bounds__(RS_THREADS) k_rsort_hist(const unsigned long long* __restrict__ klo, const unsigned long long* __restrict__ khi,
                                                           uint32_t n, int shift, uint32_t n_tiles, int rounds, uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[RS_WARPS][256];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int d = lane; d < 256; d += 32) sh[wid][d] = 0u;
  __syncwarp();
  const uint32_t tile = blockIdx.x * RS_WARPS + wid;
  if (tile >= n_tiles) return;
  const unsigned long long* src = shift < 64 ? klo : khi;
  const int s = shift & 63;
  const uint32_t base = tile * (uint32_t)(32 * rounds);
#pragma unroll 8
  for (int r = 0; r < rounds; ++r) {
    const uint32_t e = base + 32u * r + lane;
    if (e < n) atomicAdd(&sh[wid][(uint32_t)(src[e] >> s) & 255u], 1u);
  }
  __syncwarp();
  for (int d = lane; d < 256; d += 32) hist[(size_t)d * n_tiles + tile] = sh[wid][d];
}

__global__ void __launch_bounds__(RS_THREADS) k_rsort_scatter(const unsigned long long* __restrict__ klo_in, const unsigned long long* __restrict__ khi_in,
                                                              const uint32_t* __restrict__ idx_in, unsigned long long* __restrict__ klo_out,
                                                              unsigned long long* __restrict__ khi_out, uint32_t* __restrict__ idx_out, uint32_t n,
                                                              int shift, uint32_t n_tiles, int rounds, const uint32_t* __restrict__ hist) {
  __shared__ uint32_t offs[RS_WARPS][256];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t tile = blockIdx.x * RS_WARPS + wid;
  if (tile >= n_tiles) return;
  for (int d = lane; d < 256; d += 32) offs[wid][d] = hist[(size_t)d * n_tiles + tile];
  __syncwarp();
  const uint32_t base = tile * (uint32_t)(32 * rounds);
  const uint32_t lt = (1u << lane) - 1u;
  // software pipeline: the next round's loads are in flight while this round is ranked
  uint32_t e = base + lane;
  unsigned long long lo = 0ull, hi = 0ull;
  uint32_t id = 0u;
  if (e < n) {
    lo = klo_in[e];
    if (khi_in) hi = khi_in[e];
    id = idx_in[e];
  }
  for (int r = 0; r < rounds; ++r) {
    const uint32_t e_n = e + 32u;
    unsigned long long lo_n = 0ull, hi_n = 0ull;
    uint32_t id_n = 0u;
    if (r + 1 < rounds && e_n < n) {
      lo_n = klo_in[e_n];
      if (khi_in) hi_n = khi_in[e_n];
      id_n = idx_in[e_n];
    }
    const bool valid = e < n;
    const uint32_t d = valid ? rs_digit(lo, hi, shift) : 256u;  // lanes past the end form their own group and write nothing
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const uint32_t rank = __popc(peers & lt);
    uint32_t pos = 0u;
    if (valid) pos = offs[wid][d] + rank;
    __syncwarp();
    if (valid && rank == 0u) offs[wid][d] += __popc(peers);
    __syncwarp();
    if (valid) {
      klo_out[pos] = lo;
      if (khi_out) khi_out[pos] = hi;
      idx_out[pos] = id;
    }
    e = e_n, lo = lo_n, hi = hi_n, id = id_n;
  }
}

// ---- column decode + the side-chain plane join ----------------------------------------------------------------------
// Plane maths exactly as the host states it (residues.rs:31-75; host/geometry.cpp): left-to-right dot products without
// contraction, IEEE sqrt and division, acos -> fold to <= 90 degrees -> degrees.  Only acos is not correctly rounded on
// the device, so a value whose f32 rounding could change under an error of 4 ulp of pi in the angle is reported as
// sc_valid = 2 (a handful of rows per million); the distance column needs no such care.
__device__ __forceinline__ double fold_deg_dev(double c) {
  const double kPi = 3.14159265358979323846, kHalfPi = 1.57079632679489661923;
  double rad = acos(c);
  if (rad > kHalfPi) rad = dsub(kPi, rad);
  return dmul(rad, 180.0 / kPi);
}
__device__ __forceinline__ float angle_to_f32(double a, bool& doubtful) {
  const float f = __double2float_rn(a);
  if (!(a == a) || __double2float_rn(a - 2.0e-13) != f || __double2float_rn(a + 2.0e-13) != f) doubtful = true;
  return f;
}

__global__ void __launch_bounds__(256) k_frame_columns(const FrameArgs A, const uint32_t* __restrict__ idx, const FrameOut O) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= A.n_rows) return;
  const arp_row r = A.rows[idx[k]];
  const uint32_t e2[2] = {entity_of(r.from, A.n_atoms), entity_of(r.to, A.n_atoms)};
  O.from[k] = r.from;
  O.to[k] = r.to;
  O.interaction[k] = (uint8_t)r.label;
  O.distance[k] = r.distance;
  if (e2[0] >= A.n_ents || e2[1] >= A.n_ents) return;  // k_frame_keys raised the error
  uint32_t sid[2];
  arp_frame_side S[2];
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    sid[s] = A.ent_side[e2[s]];
    if (sid[s] >= A.n_sides) return;
    const uint4* p = reinterpret_cast<const uint4*>(A.sides + sid[s]);
    const uint4 a = p[0], b = p[1];
    S[s].chain = (int32_t)a.x, S[s].resn = (int32_t)a.y, S[s].resi = (int32_t)a.z, S[s].icode = (int32_t)a.w;
    S[s].altloc = (int32_t)b.x, S[s].model = b.y, S[s].chain_rank = b.z, S[s].has_sc = b.w;
    if (O.model) {
      O.col[7 * s + 0][k] = S[s].chain;
      O.col[7 * s + 1][k] = S[s].resn;
      O.col[7 * s + 2][k] = S[s].resi;
      O.col[7 * s + 3][k] = S[s].icode;
      O.col[7 * s + 4][k] = S[s].altloc;
      O.col[7 * s + 5][k] = A.ent_atomn[e2[s]];
      O.col[7 * s + 6][k] = A.ent_atomi[e2[s]];
    }
  }
  if (O.model) O.model[k] = S[0].model;
  // complex.rs:152-161: statistics exist iff both residues have a side-chain plane
  const float qnan = __int_as_float(0x7fc00000);
  float scd = qnan, sch = qnan, sca = qnan;
  uint8_t valid = 0;
  if (S[0].has_sc && S[1].has_sc) {
    const double* pf = A.side_sc + 6ull * sid[0];
    const double* pt = A.side_sc + 6ull * sid[1];
    const double c1[3] = {pf[0], pf[1], pf[2]}, n1[3] = {pf[3], pf[4], pf[5]};
    const double c2[3] = {pt[0], pt[1], pt[2]}, n2[3] = {pt[3], pt[4], pt[5]};
    const double vx = dsub(c2[0], c1[0]), vy = dsub(c2[1], c1[1]), vz = dsub(c2[2], c1[2]);
    const double lv = __dsqrt_rn(dot3(vx, vy, vz, vx, vy, vz));
    const double ln1 = norm3(n1[0], n1[1], n1[2]), ln2 = norm3(n2[0], n2[1], n2[2]);
    bool doubtful = false;
    scd = __double2float_rn(lv);                                                                                         // point_vec_dist
    sch = angle_to_f32(fold_deg_dev(__ddiv_rn(dot3(n1[0], n1[1], n1[2], n2[0], n2[1], n2[2]), dmul(ln1, ln2))), doubtful);  // dihedral
    sca = angle_to_f32(fold_deg_dev(__ddiv_rn(dot3(n1[0], n1[1], n1[2], vx, vy, vz), dmul(ln1, lv))), doubtful);           // point_vec_angle
    valid = doubtful ? 2 : 1;
  }
  O.scd[k] = scd;
  O.sch[k] = sch;
  O.sca[k] = sca;
  O.sc_valid[k] = valid;
}

}  // namespace arpd
