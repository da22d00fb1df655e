// Cell-list build for sm_100a: bounding boxes -> cell id + in-cell rank (one atomic per atom) -> exclusive scan
// of the cell counts -> scatter into cell-sorted records.  This is a single-digit radix (counting) sort with
// radix = number of cells: two passes over the atoms and one scan over the cells, no key/value ping-pong.
// Replaces pdbtbx create_hierarchy_rtree (rstar bulk load), src/contacts/complex.rs:190 and :302.
#pragma once
#include "dev_common.cuh"

namespace arpd {

// ---- order-preserving u64 encoding of doubles, for atomicMin / atomicMax --------------------------------
__device__ __forceinline__ unsigned long long enc_f64(double d) {
  unsigned long long b = (unsigned long long)__double_as_longlong(d);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
static inline double dec_f64_host(unsigned long long e) {
  unsigned long long b = (e & 0x8000000000000000ull) ? (e & 0x7fffffffffffffffull) : ~e;
  double d;
  memcpy(&d, &b, 8);
  return d;
}

// bbox[s*6 + 0..2] = min xyz (encoded), [3..5] = max xyz (encoded)
__global__ void k_bbox_init(unsigned long long* bbox, int n_struct) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_struct * 6) bbox[i] = (i % 6) < 3 ? ~0ull : 0ull;
}

// one CTA per structure of the job writes the structure id of its atoms / rings (batch jobs only)
__global__ void k_fill_struct(const StructDesc* __restrict__ descs, uint32_t* __restrict__ atom_struct,
                              uint32_t* __restrict__ ring_struct) {
  const StructDesc d = descs[blockIdx.x];
  for (uint32_t a = d.atom_begin + threadIdx.x; a < d.atom_end; a += blockDim.x) atom_struct[a] = blockIdx.x;
  for (uint32_t r = d.ring_begin + threadIdx.x; r < d.ring_end; r += blockDim.x) ring_struct[r] = blockIdx.x;
}

// Bounding boxes.  Each CTA reads BBOX_PER_CTA consecutive atoms (coalesced), reduces the atoms of the first
// structure it sees in registers / shuffles and sends the few others straight to the atomics.
constexpr int BBOX_THREADS = 256;
constexpr int BBOX_PER_THREAD = 8;
__global__ void __launch_bounds__(BBOX_THREADS)
k_bbox(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z, uint32_t n,
       const uint32_t* __restrict__ atom_struct, unsigned long long* __restrict__ bbox) {
  const uint32_t base = blockIdx.x * (BBOX_THREADS * BBOX_PER_THREAD);
  if (base >= n) return;
  const uint32_t s0 = atom_struct ? atom_struct[base] : 0u;
  double lo[3] = {1e308, 1e308, 1e308}, hi[3] = {-1e308, -1e308, -1e308};
#pragma unroll
  for (int k = 0; k < BBOX_PER_THREAD; ++k) {
    const uint32_t a = base + k * BBOX_THREADS + threadIdx.x;
    if (a >= n) break;
    const double p[3] = {x[a], y[a], z[a]};
    const uint32_t s = atom_struct ? atom_struct[a] : 0u;
    if (s == s0) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        lo[c] = fmin(lo[c], p[c]);
        hi[c] = fmax(hi[c], p[c]);
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        atomicMin(&bbox[s * 6 + c], enc_f64(p[c]));
        atomicMax(&bbox[s * 6 + 3 + c], enc_f64(p[c]));
      }
    }
  }
  __shared__ double red[6][BBOX_THREADS / 32];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    double a = lo[c], b = hi[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a = fmin(a, __shfl_xor_sync(0xffffffffu, a, o));
      b = fmax(b, __shfl_xor_sync(0xffffffffu, b, o));
    }
    if ((threadIdx.x & 31) == 0) {
      red[c][threadIdx.x >> 5] = a;
      red[3 + c][threadIdx.x >> 5] = b;
    }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    double v = red[threadIdx.x][0];
    for (int w = 1; w < BBOX_THREADS / 32; ++w) v = threadIdx.x < 3 ? fmin(v, red[threadIdx.x][w]) : fmax(v, red[threadIdx.x][w]);
    if (threadIdx.x < 3) {
      if (v < 1e307) atomicMin(&bbox[s0 * 6 + threadIdx.x], enc_f64(v));
    } else {
      if (v > -1e307) atomicMax(&bbox[s0 * 6 + threadIdx.x], enc_f64(v));
    }
  }
}

// cell coordinate along one axis, clamped into the grid
__device__ __forceinline__ int cell_coord(double p, double origin, double inv_edge, int dim) {
  int c = (int)floor((p - origin) * inv_edge);
  return min(max(c, 0), dim - 1);
}

// Pass 1 of the counting sort: cell id of every point and its arrival rank inside the cell.
// `ring_grid` selects which of the two grids of StructDesc is addressed; points are read with a stride so that
// both the SoA atom arrays (stride 1) and the interleaved ring centres (stride 3) fit.
#ifndef CA_PER_THREAD
#define CA_PER_THREAD 1  // points per thread.  The kernel waits for the atomics' return values (79 % of its stall samples), but
#endif                   // two or four points per thread -- loads, then atomics, in flight together -- measured SLOWER on config 3
                         // (grid stage 0.081 -> 0.085 / 0.086 ms, r2c): with one point per thread the SMs already hold 64 warps
__global__ void __launch_bounds__(256)
k_cell_assign(const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz, int stride, uint32_t n,
              const uint32_t* __restrict__ point_struct, const StructDesc* __restrict__ descs, double inv_edge, int ring_grid,
              uint32_t* __restrict__ cell_of, uint32_t* __restrict__ rank_of, uint32_t* __restrict__ cell_count,
              unsigned int* __restrict__ box_error) {
  const uint32_t i0 = blockIdx.x * (blockDim.x * CA_PER_THREAD) + threadIdx.x;
  uint32_t c[CA_PER_THREAD];
#pragma unroll
  for (int u = 0; u < CA_PER_THREAD; ++u) {
    const uint32_t i = i0 + u * blockDim.x;
    if (i >= n) continue;
    const StructDesc& d = descs[point_struct ? point_struct[i] : 0u];
    const int nx = ring_grid ? d.rnx : d.nx, ny = ring_grid ? d.rny : d.ny, nz = ring_grid ? d.rnz : d.nz;
    const double x = px[(size_t)i * stride], y = py[(size_t)i * stride], z = pz[(size_t)i * stride];
    if (box_error) {
      // the grid was laid out from the caller's bounding-box hint: an atom outside it (or a NaN) voids the layout
      const double tx = (x - d.ox) * inv_edge, ty = (y - d.oy) * inv_edge, tz = (z - d.oz) * inv_edge;
      const double eps = 1e-6;
      if (!(tx > -eps && tx < nx + eps && ty > -eps && ty < ny + eps && tz > -eps && tz < nz + eps)) atomicOr(box_error, 2u);
    }
    const int ix = cell_coord(x, d.ox, inv_edge, nx);
    const int iy = cell_coord(y, d.oy, inv_edge, ny);
    const int iz = cell_coord(z, d.oz, inv_edge, nz);
    c[u] = (ring_grid ? d.rcell_base : d.cell_base) + (uint32_t)((iz * ny + iy) * nx + ix);
  }
  uint32_t r[CA_PER_THREAD];
#pragma unroll
  for (int u = 0; u < CA_PER_THREAD; ++u)
    if (i0 + u * blockDim.x < n) r[u] = atomicAdd(&cell_count[c[u]], 1u);
#pragma unroll
  for (int u = 0; u < CA_PER_THREAD; ++u) {
    const uint32_t i = i0 + u * blockDim.x;
    if (i < n) cell_of[i] = c[u], rank_of[i] = r[u];
  }
}

// ---- exclusive scan of the cell counts (in place) ------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
#ifndef SCAN_PT
#define SCAN_PT 16  // with the 4-window look-back: 8 -> 23 us, 16 -> 16 us, 32 -> 14 us for the 2.6 M cells of config 3 (r1s)
#endif
constexpr int SCAN_PER_THREAD = SCAN_PT;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_PER_THREAD;
#ifndef SCAN_LOOK
#define SCAN_LOOK 4  // windows of 32 predecessor tiles fetched per look-back round trip
#endif

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
  __shared__ uint32_t wsum[SCAN_THREADS / 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) wsum[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t s = lane < SCAN_THREADS / 32 ? wsum[lane] : 0u;
#pragma unroll
    for (int o = 1; o < SCAN_THREADS / 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, s, o);
      if (lane >= o) s += t;
    }
    if (lane < SCAN_THREADS / 32) wsum[lane] = s;
  }
  __syncthreads();
  const uint32_t before = w ? wsum[w - 1] : 0u;
  *total = wsum[SCAN_THREADS / 32 - 1];
  __syncthreads();
  return before + inc - v;
}

// Exclusive scan in place, ONE pass over the data (decoupled look-back, Merrill & Garland): tiles take a ticket in
// scheduling order, publish their aggregate, and sum the aggregates / inclusive prefixes of their predecessors 32 at a
// time.  status[t] = flag << 62 | value (flag 1: aggregate of tile t, 2: inclusive prefix up to tile t), written with
// a single 64-bit store; status[n_tiles] is the ticket counter.  A tile only ever waits for tiles with smaller tickets,
// which are running or done; the wait is bounded all the same and reports through *error instead of hanging.
#define SCAN_FLAG_A (1ull << 62)
#define SCAN_FLAG_P (2ull << 62)
#define SCAN_VALUE ((1ull << 62) - 1ull)
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_lookback(uint32_t* __restrict__ data, uint32_t n, unsigned long long* status,
                                                                uint32_t n_tiles, unsigned int* __restrict__ error) {
  __shared__ uint32_t s_tile, s_prefix;
  if (threadIdx.x == 0) s_tile = atomicAdd(reinterpret_cast<unsigned int*>(status + n_tiles), 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint32_t base = tile * SCAN_TILE + threadIdx.x * SCAN_PER_THREAD;
  uint32_t v[SCAN_PER_THREAD], sum = 0;
  // whole tiles move as 16-byte words (cudaMalloc'ed base, SCAN_PER_THREAD a multiple of 4); the last tile goes word by word
  const bool wide = (SCAN_PER_THREAD % 4 == 0) && (unsigned long long)(tile + 1u) * SCAN_TILE <= n && (reinterpret_cast<uintptr_t>(data) & 15u) == 0u;
  if (wide) {
#pragma unroll
    for (int k = 0; k < SCAN_PER_THREAD / 4; ++k) {
      const uint4 q = reinterpret_cast<const uint4*>(data + base)[k];
      v[4 * k] = q.x, v[4 * k + 1] = q.y, v[4 * k + 2] = q.z, v[4 * k + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_PER_THREAD; ++k) v[k] = base + k < n ? data[base + k] : 0u;
  }
#pragma unroll
  for (int k = 0; k < SCAN_PER_THREAD; ++k) sum += v[k];
  uint32_t total;
  uint32_t run = block_exclusive_scan(sum, &total);
  volatile unsigned long long* vs = status;
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    if (tile == 0) {
      if (lane == 0) {
        vs[0] = SCAN_FLAG_P | total;
        s_prefix = 0u;
      }
    } else {
      if (lane == 0) vs[tile] = SCAN_FLAG_A | total;
      // The look-back is the serial part of the scan (all tiles of a cell table are resident at once, so the inclusive
      // prefixes spread at one window per memory round trip): a round trip fetches SCAN_LOOK windows of 32 predecessors,
      // lane l of window k looking at tile first - 32 k - l, and consumes them in order up to the first window that still
      // has an unpublished tile.
      unsigned long long part = 0;            // per lane; summed over the warp once at the end
      long long first = (long long)tile - 1;
      uint32_t spins = 0;
      bool found = false;
      while (!found) {
        unsigned long long w[SCAN_LOOK];
#pragma unroll
        for (int k = 0; k < SCAN_LOOK; ++k) {
          const long long j = first - 32 * k - lane;
          w[k] = j >= 0 ? vs[j] : SCAN_FLAG_P;  // before tile 0: an inclusive prefix of 0
        }
        int used = 0;
#pragma unroll
        for (int k = 0; k < SCAN_LOOK; ++k) {
          if (!found && used == k) {
            const uint32_t flag = (uint32_t)(w[k] >> 62);
            if (!__any_sync(0xffffffffu, flag == 0u)) {
              const uint32_t pm = __ballot_sync(0xffffffffu, flag == 2u);
              const int stop = pm ? (__ffs(pm) - 1) : 31;  // nearest predecessor that already has its inclusive prefix
              if (lane <= stop) part += w[k] & SCAN_VALUE;
              used = k + 1;
              found = pm != 0u;
            }
          }
        }
        first -= 32 * used;
        if (used == 0 && ++spins > (1u << 22)) {  // cannot happen (see above); never hang a GPU over it
          if (lane == 0) atomicOr(error, 4u);
          break;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      if (lane == 0) {
        vs[tile] = SCAN_FLAG_P | ((part + total) & SCAN_VALUE);
        s_prefix = (uint32_t)part;
      }
    }
  }
  __syncthreads();
  run += s_prefix;
  if (wide) {
#pragma unroll
    for (int k = 0; k < SCAN_PER_THREAD / 4; ++k) {
      uint4 q;
      q.x = run, run += v[4 * k];
      q.y = run, run += v[4 * k + 1];
      q.z = run, run += v[4 * k + 2];
      q.w = run, run += v[4 * k + 3];
      reinterpret_cast<uint4*>(data + base)[k] = q;
    }
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_PER_THREAD; ++k) {
      if (base + k < n) data[base + k] = run;
      run += v[k];
    }
  }
}

// res_id from the residues' atom ranges (arp_structure.res_atom_begin): a warp takes 32 consecutive residues and writes
// their atoms -- one contiguous range -- 32 at a time; an atom's residue is found among the warp's 33 range starts by a
// 5-step shuffle search (r2c; one thread per atom with a 17-step binary search in the whole CSR took 21 us for config 3).
__global__ void __launch_bounds__(256) k_expand_res_id(const uint32_t* __restrict__ res_atom_begin, uint32_t n_res, uint32_t n_atoms,
                                                       uint32_t* __restrict__ res_id) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t r0 = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) << 5;  // the warp's first residue
  if (r0 >= n_res) return;
  const uint32_t r = min(r0 + lane, n_res);                  // lanes past the last residue hold the closing entry
  const uint32_t b = min(res_atom_begin[r], n_atoms);        // (an entry beyond n_atoms would be the caller's error: clamp)
  const uint32_t first = __shfl_sync(0xffffffffu, b, 0);
  const uint32_t last = min(res_atom_begin[min(r0 + 32u, n_res)], n_atoms);
  for (uint32_t a = first + lane; a - lane < last; a += 32u) {  // whole warp in step (shuffles inside)
    uint32_t k = 0;  // largest lane index with b_k <= a
#pragma unroll
    for (uint32_t st = 16; st > 0; st >>= 1) {
      const uint32_t v = __shfl_sync(0xffffffffu, b, k + st);
      if (v <= a) k += st;
    }
    if (a < last) res_id[a] = r0 + k;
  }
}

// ---- upload: the arrays of one merged host->device copy go to their own buffers ---------------------------------------
constexpr int SEG_MAX = 24;
struct Segments {
  const unsigned char* src[SEG_MAX];
  unsigned char* dst[SEG_MAX];
  size_t bytes[SEG_MAX];  // multiples of 4; src and dst 4-byte aligned
  int n;
};
__global__ void __launch_bounds__(256) k_scatter_segments(const Segments S) {
  const int s = blockIdx.y;
  const unsigned char* __restrict__ src = S.src[s];
  unsigned char* __restrict__ dst = S.dst[s];
  const size_t bytes = S.bytes[s], tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, step = (size_t)gridDim.x * blockDim.x;
  size_t done = 0;
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15u) == 0) {
    const size_t n16 = bytes / 16;
    for (size_t i = tid; i < n16; i += step) reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
    done = n16 * 16;
  }
  for (size_t i = done / 4 + tid; i < bytes / 4; i += step) reinterpret_cast<uint32_t*>(dst)[i] = reinterpret_cast<const uint32_t*>(src)[i];
}

// f64 coordinates from the on-lattice form (arp_structure.x_milli): one IEEE division per coordinate, the correctly
// rounded double of the decimal v/1000 -- what the packer verified against the parsed value bit for bit
__global__ void __launch_bounds__(256) k_expand_milli(const int32_t* __restrict__ xm, const int32_t* __restrict__ ym,
                                                      const int32_t* __restrict__ zm, uint32_t n_atoms, double* __restrict__ x,
                                                      double* __restrict__ y, double* __restrict__ z) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n_atoms) return;
  x[a] = __ddiv_rn((double)xm[a], 1000.0);
  y[a] = __ddiv_rn((double)ym[a], 1000.0);
  z[a] = __ddiv_rn((double)zm[a], 1000.0);
}

// ... and from the 21-bit form (arp_structure.xyz_milli21: dx | dy << 21 | dz << 42 relative to an integer origin), with
// the 16-bit attribute words widened in the same pass (attr16 may be NULL: attr then came as 32-bit words)
__global__ void __launch_bounds__(256) k_expand_milli21(const unsigned long long* __restrict__ w, int ox, int oy, int oz,
                                                        const unsigned short* __restrict__ attr16, uint32_t n_atoms,
                                                        double* __restrict__ x, double* __restrict__ y, double* __restrict__ z,
                                                        uint32_t* __restrict__ attr) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n_atoms) return;
  if (w) {
    const unsigned long long v = w[a];
    x[a] = __ddiv_rn((double)(ox + (int)(v & 0x1fffffull)), 1000.0);
    y[a] = __ddiv_rn((double)(oy + (int)((v >> 21) & 0x1fffffull)), 1000.0);
    z[a] = __ddiv_rn((double)(oz + (int)((v >> 42) & 0x1fffffull)), 1000.0);
  }
  if (attr16) attr[a] = attr16[a];
}

#ifndef BUILD_BY_SLOT
#define BUILD_BY_SLOT 0  // 1: emit-radius mode builds its records in two kernels (k_scatter_atoms: slot -> atom; k_build_records)
#endif

// Pass 2 of the counting sort: write the cell-sorted records.  The atom that arrived first in a cell also
// publishes the cell as a work item of the pair kernel.
__global__ void __launch_bounds__(256)
k_scatter_atoms(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z,
                const uint32_t* __restrict__ attr, const uint32_t* __restrict__ chain_key, const uint32_t* __restrict__ res_ordinal,
                const uint32_t* __restrict__ res_id, const uint32_t* __restrict__ res_chain_key, const uint32_t* __restrict__ res_res_ordinal,
                const uint32_t* __restrict__ atom_struct, const StructDesc* __restrict__ descs, const uint32_t* __restrict__ cell_of, const uint32_t* __restrict__ rank_of, const uint32_t* __restrict__ cell_start,
                uint32_t n, double edge, double inv_edge, RecA* __restrict__ recA, RecB* __restrict__ recB, Item* __restrict__ items,
                uint32_t* __restrict__ cations, Counters* __restrict__ ctr, RecK* __restrict__ recK, float* __restrict__ pos2,
                SlotCell* __restrict__ slot_cell, const uint32_t* __restrict__ res_tag, uint32_t* __restrict__ slot_atom) {
  // slots in `items` / `cations` are reserved per BLOCK (shared-memory counters, one global atomicAdd each): the 110 000
  // same-address atomics of a one-per-cell reservation were 40 % of this kernel's stall samples
  __shared__ uint32_t s_items, s_cat, s_items_base, s_cat_base;
  __shared__ uint32_t s_fc[256], s_fs[256];  // cell, structure and position of the block's first arrivals, compacted
  __shared__ double s_fp[3][256];
  if (threadIdx.x == 0) s_items = s_cat = 0u;
  __syncthreads();
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = a < n;
  const int lane = threadIdx.x & 31;
  const uint32_t lt = (1u << lane) - 1u;
  uint32_t s = 0u, c = 0u, r = 1u, slot = 0u, w = 0u;
  double px = 0.0;
  double py = 0.0, pz = 0.0;
  if (valid) {
  s = atom_struct ? atom_struct[a] : 0u;
  const StructDesc& d = descs[s];
  c = cell_of[a], r = rank_of[a];
  slot = cell_start[c] + r;
  ARP_CHK(ctr, slot < n && slot < cell_start[c + 1], 10);  // the atom's place lies inside its cell's slot range
  if (slot_atom) {  // two-kernel build (BUILD_BY_SLOT): this pass only says which atom sits in which slot
    slot_atom[slot] = a;
    w = attr[a] & 0x7fffffffu;
  } else {
  px = x[a], py = y[a], pz = z[a];
  RecA ra;
  ra.x = (float)(px - d.ox);
  ra.y = (float)(py - d.oy);
  ra.z = (float)(pz - d.oz);
  // bit 31 of the sorted attribute word (internal): the atom may start a row / counts as `from` (complex.rs:200-202)
  w = attr[a] & 0x7fffffffu;
  ra.attr = w | (((w & ARP_ATTR_LIGAND) && (w & ARP_ATTR_ELEM_MASK) && !(w & ARP_ATTR_GHOST)) ? 0x80000000u : 0u);
  ra.res_id = res_id[a];
  // per-atom arrays, or the compact per-residue form (arp_structure.res_chain_key): the residue tables are L2-resident
  ra.chain_key = res_chain_key ? res_chain_key[ra.res_id] : chain_key[a];
  ra.res_ordinal = res_chain_key ? res_res_ordinal[ra.res_id] : res_ordinal[a];
  ra.perm = a;
  RecB rb;
  rb.x = px;
  rb.y = py;
  rb.z = pz;
  rb.attr = ra.attr;
  rb.res_id = ra.res_id;
  recB[slot] = rb;
  recK[slot] = RecK{ra.attr, ra.chain_key, ra.res_ordinal, a};
  if (pos2) {
    // emit-radius mode, what k_screen_atoms reads: positions and residue tags interleaved per slot PAIR
    // (x0 x1 | y0 y1 | z0 z1 | t0 t1: 32 bytes), and the slot's cell with its boundary bits
    float* q = pos2 + 8u * (size_t)(slot >> 1) + (slot & 1u);
    q[0] = ra.x, q[2] = ra.y, q[4] = ra.z;
    q[6] = __uint_as_float(res_tag[ra.res_id]);
    const int ix = cell_coord(px, d.ox, inv_edge, d.nx), iy = cell_coord(py, d.oy, inv_edge, d.ny), iz = cell_coord(pz, d.oz, inv_edge, d.nz);
    slot_cell[slot] = SlotCell{c, (ix > 0 ? 1u : 0u) | (ix + 1 < d.nx ? 2u : 0u) | (iy > 0 ? 4u : 0u) | (iy + 1 < d.ny ? 8u : 0u) | (iz + 1 < d.nz ? 16u : 0u)};
  } else {
    recA[slot] = ra;
  }
  }
  }
  // the few atoms that can make a cation-pi contact (aromatic.rs:18) are listed for the ring kernel; the first arrival
  // of a cell publishes the cell as a work item
  const bool is_cat = valid && (w & ARP_ATTR_POS_RES), first = valid && r == 0u;
  uint32_t my_cat = 0u, my_item = 0u;
  {
    const uint32_t bc = __ballot_sync(0xffffffffu, is_cat), bi = __ballot_sync(0xffffffffu, first);
    uint32_t wc = 0u, wi = 0u;
    if (lane == 0) {
      if (bc) wc = atomicAdd(&s_cat, (uint32_t)__popc(bc));
      if (bi) wi = atomicAdd(&s_items, (uint32_t)__popc(bi));
    }
    my_cat = __shfl_sync(0xffffffffu, wc, 0) + __popc(bc & lt);
    my_item = __shfl_sync(0xffffffffu, wi, 0) + __popc(bi & lt);
  }
  if (first && !pos2) s_fc[my_item] = c, s_fs[my_item] = s, s_fp[0][my_item] = px, s_fp[1][my_item] = py, s_fp[2][my_item] = pz;
  __syncthreads();
  if (threadIdx.x == 0) {
    s_cat_base = s_cat ? atomicAdd(&ctr->n_cations, s_cat) : 0u;
    s_items_base = s_items ? atomicAdd(&ctr->n_items, s_items) : 0u;
  }
  __syncthreads();
  if (is_cat) cations[s_cat_base + my_cat] = slot;
  // the cells' half shells as 5 runs of sorted slots, one THREAD per first arrival of the block (one atom in four starts a
  // cell at the emit-radius edge: built in place, three quarters of every warp idled through these 150 instructions)
  const uint32_t n_build = pos2 ? 0u : s_items;  // the emit-radius screen finds its runs through slot_cell + cell_start: no work items
  for (uint32_t i = threadIdx.x; i < n_build; i += blockDim.x) {
    const uint32_t fc = s_fc[i], fs = s_fs[i];
    const StructDesc& d = descs[fs];
    const int nx = d.nx, ny = d.ny, nz = d.nz;
    // the cell coordinates from the position, exactly as k_cell_assign computed them (no integer divisions)
    const int ix = cell_coord(s_fp[0][i], d.ox, inv_edge, nx), iy = cell_coord(s_fp[1][i], d.oy, inv_edge, ny),
              iz = cell_coord(s_fp[2][i], d.oz, inv_edge, nz);
    Item it;
    it.cb = cell_start[fc];
    it.n_own = cell_start[fc + 1] - it.cb;
    it.sid = fs;
    // run 0: own cell and its +x neighbour; runs 1..4: rows (dy, dz) = (+1,0) (-1,+1) (0,+1) (+1,+1), x from ix-1 to ix+1
    uint32_t acc = cell_start[fc + (ix + 1 < nx ? 2 : 1)] - it.cb;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int dy = (k == 0) ? 1 : (k - 2), dz = (k == 0) ? 0 : 1;
      const int y = iy + dy, z = iz + dz;
      uint32_t b = 0, e = 0;
      if (y >= 0 && y < ny && z < nz) {
        const uint32_t row = d.cell_base + (uint32_t)((z * ny + y) * nx);
        b = cell_start[row + (uint32_t)max(ix - 1, 0)];
        e = cell_start[row + (uint32_t)min(ix + 1, nx - 1) + 1u];
      }
      it.off[k] = acc;
      it.delta[k] = b - acc;
      acc += e - b;
    }
    it.total = acc;
    it.cx = (float)(((double)ix + 0.5) * edge);
    it.cy = (float)(((double)iy + 0.5) * edge);
    it.cz = (float)(((double)iz + 0.5) * edge);
    it.pad = 0;
    // a compact list: k_screen hands the items out to warps
    items[s_items_base + i] = it;
  }
}

// Second pass of the two-kernel build of the emit-radius mode (BUILD_BY_SLOT): one thread per sorted SLOT gathers its atom's
// inputs and writes the records -- RecB, RecK, pos2, SlotCell -- so that a warp's stores are contiguous (the one-kernel
// build writes the same 72 bytes per atom from the atom's side: eight scattered stores per lane).
__global__ void __launch_bounds__(256)
k_build_records(const uint32_t* __restrict__ slot_atom, const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ z,
                const uint32_t* __restrict__ attr, const uint32_t* __restrict__ chain_key, const uint32_t* __restrict__ res_ordinal,
                const uint32_t* __restrict__ res_id, const uint32_t* __restrict__ res_chain_key, const uint32_t* __restrict__ res_res_ordinal,
                const uint32_t* __restrict__ atom_struct, const StructDesc* __restrict__ descs, const uint32_t* __restrict__ cell_of,
                uint32_t n, double inv_edge, RecB* __restrict__ recB, RecK* __restrict__ recK, float* __restrict__ pos2,
                SlotCell* __restrict__ slot_cell, const uint32_t* __restrict__ res_tag) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= n) return;
  const uint32_t a = slot_atom[slot];
  const StructDesc& d = descs[atom_struct ? atom_struct[a] : 0u];
  const double px = x[a], py = y[a], pz = z[a];
  const uint32_t w = attr[a] & 0x7fffffffu;
  const uint32_t wa = w | (((w & ARP_ATTR_LIGAND) && (w & ARP_ATTR_ELEM_MASK) && !(w & ARP_ATTR_GHOST)) ? 0x80000000u : 0u);
  const uint32_t rid = res_id[a];
  const uint32_t ck = res_chain_key ? res_chain_key[rid] : chain_key[a];
  const uint32_t ro = res_chain_key ? res_res_ordinal[rid] : res_ordinal[a];
  RecB rb;
  rb.x = px, rb.y = py, rb.z = pz;
  rb.attr = wa;
  rb.res_id = rid;
  recB[slot] = rb;
  recK[slot] = RecK{wa, ck, ro, a};
  float* q = pos2 + 8u * (size_t)(slot >> 1) + (slot & 1u);
  q[0] = (float)(px - d.ox), q[2] = (float)(py - d.oy), q[4] = (float)(pz - d.oz);
  q[6] = __uint_as_float(res_tag[rid]);
  const int ix = cell_coord(px, d.ox, inv_edge, d.nx), iy = cell_coord(py, d.oy, inv_edge, d.ny), iz = cell_coord(pz, d.oz, inv_edge, d.nz);
  slot_cell[slot] = SlotCell{cell_of[a], (ix > 0 ? 1u : 0u) | (ix + 1 < d.nx ? 2u : 0u) | (iy > 0 ? 4u : 0u) | (iy + 1 < d.ny ? 8u : 0u) | (iz + 1 < d.nz ? 16u : 0u)};
}

__global__ void __launch_bounds__(256)
k_scatter_rings(const uint32_t* __restrict__ cell_of, const uint32_t* __restrict__ rank_of, const uint32_t* __restrict__ rcell_start,
                uint32_t n, uint32_t* __restrict__ rperm) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) rperm[rcell_start[cell_of[r]] + rank_of[r]] = r;
}

}  // namespace arpd
