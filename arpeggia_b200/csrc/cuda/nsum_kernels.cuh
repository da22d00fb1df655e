// Weighted neighbour sum over the atoms of the resident job: the radius query of the SAP score,
// src/sap.rs:181-215 -- for every selected atom x: sum of weight[y] over the selected atoms y (x itself included) with
// distance_2(x, y) <= radius^2, the same create_hierarchy_rtree().locate_within_distance primitive as the contacts path
// (inclusive, rstar), with the squared radius formed in f32 and widened (sap.rs:184).  Same cell-list build as the
// contacts job (k_cell_assign + k_scan_lookback), its own sorted record and a full 27-cell sweep: the sum is not
// symmetric in (x, y), every atom needs all of its neighbours.
#pragma once
#include "grid_kernels.cuh"

namespace arpd {

struct __align__(32) NsRec {  // one sector per selected atom
  double x, y, z;
  float w;
  uint32_t perm;  // input-order index
};

struct __align__(16) NsRecF {  // what the f32 screen reads: position relative to the grid origin, weight
  float x, y, z, w;
};

constexpr uint32_t NS_UNSELECTED = 0xffffffffu;

// The cell list holds the SELECTED atoms only (SAP selects the side-chain atoms, about half of a protein): an unselected
// atom neither contributes nor receives (sap.rs:181-215 iterates over the selected set), so it gets its 0 here and is
// never sorted, screened or given a lane in the sum.
__global__ void __launch_bounds__(256)
k_ns_cell_assign(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z,
                 const unsigned char* __restrict__ select, uint32_t n, const uint32_t* __restrict__ atom_struct,
                 const StructDesc* __restrict__ descs, double inv_edge, uint32_t* __restrict__ cell_of, uint32_t* __restrict__ rank_of,
                 uint32_t* __restrict__ cell_count, float* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (!select[i]) {
    cell_of[i] = NS_UNSELECTED;
    out[i] = 0.f;
    return;
  }
  const StructDesc& d = descs[atom_struct ? atom_struct[i] : 0u];  // batch jobs: every structure has its own grid
  const int ix = cell_coord(x[i], d.ox, inv_edge, d.nx), iy = cell_coord(y[i], d.oy, inv_edge, d.ny), iz = cell_coord(z[i], d.oz, inv_edge, d.nz);
  const uint32_t c = d.cell_base + (uint32_t)((iz * d.ny + iy) * d.nx + ix);
  cell_of[i] = c;
  rank_of[i] = atomicAdd(&cell_count[c], 1u);
}

__global__ void __launch_bounds__(256)
k_ns_scatter(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z, const float* __restrict__ weight,
             const uint32_t* __restrict__ cell_of, const uint32_t* __restrict__ rank_of, const uint32_t* __restrict__ cell_start, uint32_t n,
             const uint32_t* __restrict__ atom_struct, const StructDesc* __restrict__ descs, NsRec* __restrict__ rec,
             NsRecF* __restrict__ recf) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n) return;
  const uint32_t c = cell_of[a];
  if (c == NS_UNSELECTED) return;
  NsRec r;
  r.x = x[a], r.y = y[a], r.z = z[a];
  r.w = weight[a];
  r.perm = a;
  const uint32_t slot = cell_start[c] + rank_of[a];
  rec[slot] = r;
  const StructDesc& d = descs[atom_struct ? atom_struct[a] : 0u];
  recf[slot] = NsRecF{(float)(r.x - d.ox), (float)(r.y - d.oy), (float)(r.z - d.oz), r.w};
}

// One lane per sorted slot (= per selected atom; cell_start[n_cells] is their number).  Every candidate of the 27-cell
// shell is screened in f32 on the origin-relative positions
// (difference form; |d2_f32 - d2| < g, the guard band of the contacts path, api_run.inc): d2_f32 <= r2 - g is certainly
// inside, d2_f32 > r2 + g certainly outside, and only the shell in between takes the exact FP64 test in the reference's
// operation order ((dx*dx + dy*dy) + dz*dz, every operation rounded once, inclusive comparison) -- so the neighbour SET
// is the reference's.  The sum is accumulated in f64 and rounded to f32 once (the reference adds f32 terms in the
// R*-tree's iteration order, which is unspecified: any order of the same terms lies within n * 2^-24 * sum|w| of this).
__global__ void __launch_bounds__(256)
k_neighbor_sum(const NsRec* __restrict__ rec, const NsRecF* __restrict__ recf, const uint32_t* __restrict__ cell_start,
               const uint32_t* __restrict__ atom_struct, const StructDesc* __restrict__ descs, uint32_t n_cells, double inv_edge, double r2,
               float thr_cert, float thr_hit, float* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cell_start[n_cells]) return;
  const NsRec me = rec[i];
  const NsRecF mf = recf[i];
  const StructDesc d = descs[atom_struct ? atom_struct[me.perm] : 0u];  // the shell never leaves the structure's own cells
  const int ix = cell_coord(me.x, d.ox, inv_edge, d.nx), iy = cell_coord(me.y, d.oy, inv_edge, d.ny), iz = cell_coord(me.z, d.oz, inv_edge, d.nz);
  const int x0 = max(ix - 1, 0), x1 = min(ix + 1, d.nx - 1);
  double acc = 0.0;
  for (int dz = -1; dz <= 1; ++dz) {
    const int zc = iz + dz;
    if (zc < 0 || zc >= d.nz) continue;
    for (int dy = -1; dy <= 1; ++dy) {
      const int yc = iy + dy;
      if (yc < 0 || yc >= d.ny) continue;
      const uint32_t row = d.cell_base + (uint32_t)((zc * d.ny + yc) * d.nx);
      const uint32_t b = cell_start[row + (uint32_t)x0], e = cell_start[row + (uint32_t)x1 + 1u];
      double exact = 0.0;
      for (uint32_t j = b; j < e; ++j) {
        const NsRecF q = recf[j];
        const float fx = q.x - mf.x, fy = q.y - mf.y, fz = q.z - mf.z;
        const float d2f = fmaf(fz, fz, fmaf(fy, fy, fx * fx));
        if (d2f <= thr_cert) {
          exact += (double)q.w;
        } else if (d2f <= thr_hit) {  // too close to the radius to call in f32
          const NsRec p = rec[j];
          const double dx = dsub(p.x, me.x), dyy = dsub(p.y, me.y), dzz = dsub(p.z, me.z);
          const double d2 = dadd(dadd(dmul(dx, dx), dmul(dyy, dyy)), dmul(dzz, dzz));
          if (d2 <= r2) exact += (double)q.w;
        }
      }
      acc += exact;
    }
  }
  out[me.perm] = (float)acc;
}

}  // namespace arpd
