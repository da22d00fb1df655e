// Device-side data layout shared by the kernels and the C-ABI glue (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "arpeggia_cuda.h"

namespace arpd {

// Per structure of a job: atom grid (edge = screening radius) and ring-centroid grid (edge = 6.0 A).
struct StructDesc {
  double ox, oy, oz;       // grid origin = bounding-box minimum of the structure's atoms
  int nx, ny, nz;          // atom grid
  int rnx, rny, rnz;       // ring-centroid grid
  uint32_t cell_base;      // first cell id of this structure in the job-wide cell arrays
  uint32_t rcell_base;
  uint32_t atom_begin, atom_end;
  uint32_t ring_begin, ring_end;
};

// Sorted (by cell) atom records, 32 bytes = one sector each.
// RecA: everything the f32 stages need about an atom.  The screen reads the first 16 bytes (origin-relative f32 position +
//       attribute word: one 16-byte cp.async per candidate), the pair filter reads the whole sector in one gather.
// RecB: everything the exact stage reads about an atom, ONE sector: f64 position, attribute word, residue id.
// RecK: what the pair filter of the emit-radius mode and the ring kernel read: attribute word, chain key, residue ordinal,
//       input index.  (The f32 positions of that mode live in the pair-interleaved array pos2, see k_screen_atoms.)
struct __align__(32) RecA {
  float x, y, z;
  uint32_t attr;         // ARP_ATTR_* | bit 31: may start a row (k_scatter_atoms)
  uint32_t chain_key;
  uint32_t res_ordinal;
  uint32_t res_id;       // job-global residue id (H lists, CB / SG slots)
  uint32_t perm;         // job-global atom index in input order (output rows)
};
struct __align__(32) RecB {
  double x, y, z;
  uint32_t attr;         // as RecA.attr
  uint32_t res_id;       // job-global residue id (H lists, CB / SG slots)
};
struct __align__(16) RecK {
  uint32_t attr;         // as RecA.attr
  uint32_t chain_key;
  uint32_t res_ordinal;
  uint32_t perm;         // job-global atom index in input order (output rows)
};
// emit-radius mode: the cell of a sorted slot and which of its neighbours exist
//   bits: 1 = ix > 0, 2 = ix + 1 < nx, 4 = iy > 0, 8 = iy + 1 < ny, 16 = iz + 1 < nz
struct __align__(8) SlotCell {
  uint32_t cell;         // job-wide cell id
  uint32_t bits;
};

// Work item of the pair kernel: one non-empty cell and its half shell (the cell plus its 13 forward neighbours),
// which in the x-fastest cell order are 5 contiguous runs of cell-sorted atoms.  The candidates of the item are the
// concatenation of the runs; run 0 starts at the own cell, so the first n_own candidates are the cell's own atoms.
// Candidate e lives in sorted slot  e + cb  (e < off[0])  or  e + delta[r]  (off[r] <= e < off[r+1]).
// 64 bytes: a warp reads one item with a single coalesced load (word k in lane k).
struct __align__(16) Item {
  uint32_t cb, n_own, total, sid;   // own cell: first slot, atoms; candidates in the half shell; structure
  uint32_t off[4];                  // first candidate index of runs 1..4
  uint32_t delta[4];                // run r+1: first slot - off[r]
  float cx, cy, cz;                 // cell centre relative to the structure origin (f32 local frame of the screen)
  uint32_t pad;
};

// Cosine-domain images of the angle comparisons (host-computed, see host/geometry.cpp).
struct DevAngles {
  double a_ge_90, a_ge_130, a_ge_60, a_le_120;
  double f_le_30_pos, f_le_30_neg, f_le_60_pos, f_le_60_neg;
  double f_lt_30_pos, f_lt_30_neg, f_lt_60_pos, f_lt_60_neg;
};

// Everything a kernel needs that is not a pointer; passed by value (constant bank).
struct DevParams {
  double rc2;            // dist_cutoff * dist_cutoff                      (complex.rs:191)
  double comp;           // vdw_comp
  double cov[16];        // covalent-single radius per element class       (vdw.rs:24-27)
  double vdw[16];        // van der Waals radius per element class         (vdw.rs:28)
  double emit_d2_max;    // no label can fire beyond this squared distance (conservative)
  double inv_edge;       // 1 / atom-grid cell edge
  double inv_redge;      // 1 / ring-grid cell edge
  float thr_hit;         // f32 re-test: d2 <= thr_hit  -> may be within the screening radius  (r_screen^2 + g)
  float thr_cert;        //              d2 <= thr_cert -> certainly within dist_cutoff         (rc^2 - g)
  float thr_near;        //              d2 <= thr_near -> may be within the emit radius        (r_emit^2 + g)
  float scr_hit;         // the same three thresholds for the dot-product screen (wider guard band gA >= g)
  float scr_cert;
  float scr_near;
  float guard;           // g (added to the per-pair label limits)
  float guard_a;         // gA (added to the per-atom label limits of the screen)
  float polar_d2;        // largest squared distance of an H-bond / polar label: 4.0^2 with hydrogens in the job, else 3.5^2
  float edge;            // atom-grid cell edge
  int count_classified;  // 1: every in-cutoff pair goes through the exact filter (pairs_classified is counted)
  uint32_t flags;        // ARP_FLAG_*
  int count_in_cutoff;   // 1: screening radius == dist_cutoff, pairs_in_cutoff is counted
  int n_struct;
  DevAngles ang;
};

// job-wide counters, one cache line each to keep the atomics apart
struct Counters {
  unsigned long long rows;        uint64_t pad0[15];
  unsigned long long screened;    uint64_t pad1[15];
  unsigned long long in_cutoff;   uint64_t pad2[15];
  unsigned long long classified;  uint64_t pad3[15];
  unsigned int work;              uint32_t padw[31];   // work-item cursor of the persistent pair kernel
  unsigned int n_items;           uint32_t padi[31];   // non-empty cells
  unsigned int n_cations;         uint32_t padc[31];   // atoms that can take part in a cation-pi contact (aromatic.rs:18)
  unsigned int error;             uint32_t pade[31];   // 1 = CYS residue without CB / SG (vdw.rs:58 unwrap)
  unsigned long long q_len;       uint64_t padq[15];   // pairs pushed by k_screen (may exceed the queue's capacity)
  unsigned long long q2_len;      uint64_t padr[15];   // pairs pushed by k_filter (likewise)
};

// Checked build (-DARP_CHECKED=1, scratch/checked.sh): the index arithmetic the kernels rely on is verified on the device
// -- slots below the atom count, cells inside the cell table, per-lane list and stash positions inside their arrays,
// queue entries decoded to valid slots -- and a violated check sets bit (16 + id) of Counters.error, which fails the job
// on the host with the check's number.  compute-sanitizer is closed on this pool; this is the stand-in (profiles/).
#ifndef ARP_CHECKED
#define ARP_CHECKED 0
#endif
#if ARP_CHECKED
#define ARP_CHK(ctr, cond, id) do { if (!(cond)) atomicOr(&(ctr)->error, 0x10000u << (id)); } while (0)
#else
#define ARP_CHK(ctr, cond, id) do { } while (0)
#endif

struct PairArgs {
  const RecA* recA;            // counting mode
  const RecB* recB;
  const RecK* recK;
  const SlotCell* slot_cell;   // emit-radius mode
  const uint32_t* atom_struct; // input-order atom -> structure; nullptr when the job has one structure
  const uint32_t* cell_start;  // n_cells + 1
  uint32_t n_atoms, n_cells;   // sizes of the sorted arrays / the cell table (checked build)
  const Item* items;           // non-empty cells
  const StructDesc* descs;
  // original-order arrays (gathers for disulfide / hydrogen lists)
  const double* x;
  const double* y;
  const double* z;
  const int32_t* res_cb;
  const int32_t* res_sg;
  const uint32_t* res_h_begin;
  const double* hx;
  const double* hy;
  const double* hz;
  uint2* pairq;                // k_screen -> k_filter: (from slot, candidate slot)
  unsigned long long pairq_cap;
  uint4* pairq2;               // k_filter / k_screen_atoms -> k_exact: (slot a | flags, slot b | flags, input index a, input index b)
  unsigned long long pairq2_cap;
  arp_row* rows;
  unsigned long long row_cap;
  Counters* ctr;
};

struct RingArgs {
  const double* center;        // 3 * n_rings
  const double* normal;
  const uint32_t* attr;
  const uint32_t* chain_key;
  const uint32_t* res_ordinal;
  const uint32_t* ring_struct;  // nullptr when the job has one structure
  uint32_t n_rings;
  // ring-centroid cell list
  const uint32_t* rcell_start;
  const uint32_t* rperm;        // sorted slot -> ring index
  // cation list: sorted slots of the atoms with ARP_ATTR_POS_RES (k_scatter_atoms)
  const uint32_t* cations;
  const uint32_t* atom_struct;  // input-order atom -> structure; nullptr when the job has one structure
  const RecK* recK;
  const RecB* recB;
  const StructDesc* descs;
  arp_row* rows;
  unsigned long long row_cap;
  Counters* ctr;
};

// ---- exact FP64 helpers: every operation individually rounded (no FMA contraction), the order the reference uses ----
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dot3(double ax, double ay, double az, double bx, double by, double bz) {
  return dadd(dadd(dmul(ax, bx), dmul(ay, by)), dmul(az, bz));
}
__device__ __forceinline__ double norm3(double ax, double ay, double az) { return __dsqrt_rn(dot3(ax, ay, az, ax, ay, az)); }

// folded angle <= 30 / <= 60 / < 30 / < 60 in the cosine domain; |c| > 1 and NaN give false like acos -> NaN
__device__ __forceinline__ bool fold_pred(double c, double pos, double neg) {
  return (c >= pos && c <= 1.0) || (c <= neg && c >= -1.0);
}

}  // namespace arpd
