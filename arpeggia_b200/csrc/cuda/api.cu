// C ABI of the device path (include/arpeggia_cuda.h): context, staged upload / run / download, one-shot calls.
// sm_100a only; every kernel of the job is launched on the context's own stream.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <type_traits>
#include <vector>

#include "../host/common.hpp"
#include "grid_kernels.cuh"
#include "pair_kernel.cuh"
#include "ring_kernels.cuh"
#include "frame_kernels.cuh"
#include "nsum_kernels.cuh"
#include "slab_kernels.cuh"

using namespace arpd;

namespace {

struct CudaFail {
  int code;
  std::string msg;
};

#define CK(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess)                                                                               \
      throw CudaFail{e_ == cudaErrorMemoryAllocation ? ARP_ERR_OOM : ARP_ERR_CUDA,                       \
                     std::string(#call) + ": " + cudaGetErrorString(e_)};                                \
  } while (0)

// grow-only device / pinned-host buffers
template <class T>
struct DBuf {
  T* p = nullptr;
  size_t cap = 0;
  void need(size_t n) {
    if (n <= cap) return;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    const size_t want = n + n / 8 + 64;
    CK(cudaMalloc((void**)&p, want * sizeof(T)));
    cap = want;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};
template <class T>
struct HBuf {
  T* p = nullptr;
  size_t cap = 0;
  void need(size_t n) {
    if (n <= cap) return;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    const size_t want = n + n / 8 + 64;
    CK(cudaMallocHost((void**)&p, want * sizeof(T)));
    cap = want;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
};

// one array of an upload that goes straight from the caller's memory to its device buffer
struct UpItem {
  const unsigned char* h;
  unsigned char* d;
  size_t bytes;
  int after;  // 0 = nothing; 1 / 2: a kernel of the upload waits for this array (see flush_uploads)
};

}  // namespace

struct arp_ctx {
  int device = 0;
  int sm_count = 148;
  int occ_screen = 1, occ_filter = 1, occ_exact = 1, occ_screen_atoms = 1;  // resident CTAs per SM (grids are sized to one full wave)
  cudaStream_t stream = nullptr;
  cudaStream_t side = nullptr;  // k_rings runs here, concurrently with the pair kernels (forked / joined with events)
  std::string err;

  // job inputs in HBM (original order)
  int64_t n_struct = 0, n_atoms = 0, n_res = 0, n_rings = 0, n_h = 0;
  DBuf<double> x, y, z, hx, hy, hz, ring_c, ring_n;
  DBuf<uint32_t> attr, chain_key, res_ordinal, res_id, res_h_begin, ring_attr, ring_chain_key, ring_res_ordinal;
  DBuf<uint32_t> res_chain_key, res_res_ordinal;  // compact per-residue form of chain_key / res_ordinal
  DBuf<uint32_t> res_atom_begin;                  // CSR of the residues' atoms (res_id is expanded from it on the device)
  bool per_residue_keys = false;                  // every structure of the job came with it
  DBuf<int32_t> res_cb, res_sg;
  DBuf<int32_t> xm, ym, zm;                       // on-lattice coordinates as uploaded (arp_structure.x_milli)
  DBuf<uint64_t> xyz21;                           // ... or as 21-bit words (arp_structure.xyz_milli21)
  DBuf<uint16_t> attr16;                    // 16-bit attribute words as uploaded (arp_structure.attr16)
  int xyz21_origin[3] = {0, 0, 0};
  DBuf<uint32_t> atom_struct, ring_struct;
  DBuf<StructDesc> descs;
  std::vector<StructDesc> h_descs;
  std::vector<int64_t> atom_begin, ring_begin;
  std::vector<double> hint_lo, hint_hi;  // callers' bounding boxes (arp_structure.bbox_*), 3 per structure
  bool have_hint = false;                // every structure of the job came with one
  bool ignore_hint = false;              // a hint turned out wrong: the job is re-run with the computed boxes

  // global ids of the job's atoms / rings and per-residue tags (arp_structure.atom_gid / ring_gid / res_tag)
  DBuf<uint32_t> atom_gid, ring_gid, res_tag, struct_res_begin;
  DBuf<unsigned int> up_err;       // error word of the upload-time kernels
  HBuf<unsigned int> h_up_err;
  bool have_gid = false;           // every structure of the job came with ids: rows carry them
  bool have_tag = false;           // every structure of the job came with tags (else they are derived on the device)
  bool tags_ok = false;            // res_tag is resident and valid: the emit-radius screen can run
  bool have_res_csr = false;       // res_atom_begin is resident (every structure came with it)
  bool rows_global = false;        // the rows of the last run have been rewritten to global ids (arp_stage_download)

  // slab jobs (arp_slab_exchange): communicator, the owned part of the resident job, exchange scratch
  struct Slab {
    void* comm = nullptr;          // ncclComm_t
    int rank = 0, world = 1;
    std::vector<arp_ctx*> clique;  // arp_ctx_create_multi: every context of the set, in rank order
    bool copy_mode = false;        // a device appears twice in the set: messages move as peer copies
    int64_t own_atoms = -1, own_res = 0, own_rings = 0, own_h = 0;  // -1: no exchange since the upload
    bool have_extent = false;
    std::vector<double> all_box;   // 7 per rank: min xyz, max xyz of the owned atoms, largest residue x-extent
    std::vector<int> peers;        // ranks within reach, in the order of the per-peer arrays
    uint32_t n_blocks = 0;
    DBuf<unsigned long long> extent, gather, row;
    HBuf<unsigned long long> h_gather, h_small;
    DBuf<uint32_t> res_mask, ring_mask;
    DBuf<Tri> block_tri;
    DBuf<HaloCounts> counts;
    DBuf<unsigned int> cursor;
    DBuf<int> peer_rank;
    DBuf<unsigned char> send[HALO_MAX_PEERS], recv[HALO_MAX_PEERS];
    cudaEvent_t ev[5] = {};
    int64_t ghost[4] = {0, 0, 0, 0}, sent[4] = {0, 0, 0, 0}, bytes_sent = 0, bytes_recv = 0;
  } slab;

  // cell lists and sorted records
  DBuf<unsigned long long> bbox;
  DBuf<uint32_t> cell_of, rank_of, cell_start;
  DBuf<unsigned long long> scan_status, rscan_status;
  DBuf<uint32_t> rcell_of, rrank_of, rcell_start, rperm, cations;
  DBuf<float> pos2;          // sorted positions + residue tags interleaved per slot pair (k_screen_atoms)
  DBuf<SlotCell> slot_cell;  // cell and boundary bits per sorted slot (k_screen_atoms)
  DBuf<uint32_t> slot_atom;  // BUILD_BY_SLOT: the atom of every sorted slot (k_scatter_atoms -> k_build_records)
  DBuf<RecA> recA;           // counting mode
  DBuf<RecB> recB;
  DBuf<RecK> recK;
  DBuf<Item> items;
  DBuf<Counters> ctr;
  DBuf<arp_row> rows;
  uint64_t row_cap = 0;
  DBuf<uint2> pairq;          // pairs between k_screen -> k_filter
  DBuf<uint4> pairq2;         // pairs for k_exact (with the input indices of both atoms)
  uint64_t pairq_cap = 0, pairq2_cap = 0;
  int64_t n_cells = 0, n_rcells = 0;

  // the weighted neighbour sum (arp_stage_neighbor_sum): its own cell list over the job's atoms
  DBuf<StructDesc> ns_desc;
  DBuf<uint32_t> ns_cell_of, ns_rank_of, ns_cell_start;
  DBuf<unsigned long long> ns_status;
  DBuf<NsRec> ns_rec;
  DBuf<NsRecF> ns_recf;
  DBuf<float> ns_weight, ns_out;
  DBuf<unsigned char> ns_select;
  cudaEvent_t nev[4] = {};
  float ns_ms_grid = 0.f, ns_ms_sum = 0.f;
  int64_t ns_cells = 0;

  // the frame (arp_stage_frame): the caller's tables, sort ping-pong buffers, output columns
  DBuf<arp_frame_side> f_sides;
  DBuf<double> f_side_sc;
  DBuf<uint32_t> f_ent_side, f_ent_rank, f_ent_tie;
  DBuf<int32_t> f_ent_atomn, f_ent_atomi;
  DBuf<unsigned long long> f_klo[2], f_khi[2], f_status;
  DBuf<uint32_t> f_idx[2], f_hist;
  DBuf<unsigned char> f_out;
  cudaEvent_t fev[4] = {};

  // host staging
  HBuf<unsigned long long> h_bbox;
  HBuf<Counters> h_ctr;
  HBuf<unsigned char> stage;  // concatenation buffer for batch uploads
  std::vector<HBuf<arp_row>> row_pool;  // page-locked row buffers of freed results (every arp_result owns its rows)

  // A whole run as a CUDA graph: captured when the same job is run again with the same parameters and buffer sizes
  // (parameter studies on a resident structure, the timed loop of bench.py), replayed from then on -- one graph launch
  // instead of ~25 stream operations; what it removes is launch gaps, the kernels are the same.
  struct RunKey {
    uint64_t job_gen, row_cap, pairq_cap, pairq2_cap;
    double comp, rc;
    uint32_t flags, hinted;
  };
  RunKey last_key{}, graph_key{};
  bool have_last_key = false, graph_disabled = false;
  uint64_t job_gen = 0;
  cudaGraphExec_t graph_exec = nullptr;
  int64_t graph_launches = 0;
  HBuf<StructDesc> h_descs_pin;  // page-locked copy of h_descs: the source of the descriptor upload (capturable)

  // bookkeeping of the last run
  bool have_job = false, have_run = false;
  bool settled = false;  // the last run has been waited for and its counters checked (settle_run)
  arp_params last_params{};
  DevParams dp{};
  cudaEvent_t ev[13] = {};  // timing marks (recorded as external event nodes inside a capture: still readable after a replay)
  cudaEvent_t udep[3] = {};  // upload: fork / join edges to the side stream (the expand kernels run beside the copies)
  std::vector<UpItem> up_items;  // an upload's direct copies, issued by flush_uploads (neighbours in host memory travel as one copy)
  DBuf<unsigned char> up_arena;  // landing area of such merged copies; k_scatter_segments moves the arrays to their buffers
  int up_copies = 0;             // H2D copies the last upload issued for its arrays (after merging)
  std::function<void(int)> up_after;  // the running upload's kernels that wait for a marked array
  bool h2d_pending = false;      // a one-shot call left the upload's end un-waited: ms_h2d is read once the run has settled
  HBuf<uint32_t> h_words;   // page-locked single words of an upload (CSR closing entries)
  HBuf<StructDesc> h_descs_up;  // page-locked source of the upload's descriptor copy
  cudaEvent_t dep[3] = {};  // fork / join edges between the main and the side stream (plain records: captured as dependencies)
  bool capturing = false;
  int64_t launches = 0;
  int retries = 0;
  float ms_h2d = 0.f, ms_d2h = 0.f;
  float ms_wall[4] = {0.f, 0.f, 0.f, 0.f};  // one-shot call: upload, run, download, whole call (host wall clock)
};

namespace {

void slab_release(arp_ctx* c);  // api_slab.inc

int fail(arp_ctx* c, int code, const std::string& msg) {
  if (c) c->err = msg;
  arp::last_error() = msg;
  return code;
}

template <class F>
int guarded(arp_ctx* c, F&& f) {
  try {
    f();
    return ARP_OK;
  } catch (const CudaFail& e) {
    return fail(c, e.code, e.msg);
  } catch (const arp::Error& e) {
    return fail(c, e.code, e.msg);
  } catch (const std::exception& e) {
    return fail(c, ARP_ERR_OOM, e.what());
  }
}

// Issue the queued direct copies of an upload.  Arrays that follow one another in the caller's memory (less than a page
// apart: the arrays of one arena) travel as ONE copy into c->up_arena and are moved to their buffers by one kernel --
// on this pool's boxes a 20 MB copy runs at 54 GB/s, the same bytes as five copies at 35-43 GB/s
// (profiles/r2c_pcie_copy_rates.json).  The bytes between two such arrays are read too: they lie in pages that hold
// the arrays' own first / last bytes, so they are mapped.  Copies go out in host-address order; `after(k)` is called
// once the stream holds EVERY array marked k (the upload's expand kernels wait for exactly those).
// Which allocation a host pointer belongs to, as far as a merged copy must know: pageable memory (the driver stages it
// anyway), or page-locked memory with the extent of its allocation (one copy must not cross from one page-locked
// allocation into the next: cudaMemcpyAsync refuses that).  Anything else -- device or managed memory, a failed query --
// is `other` and never merged.
struct HostSpan {
  enum Kind { pageable, pinned, other } kind = other;
  uintptr_t base = 0;
  size_t size = 0;
};
HostSpan host_span_of(const void* p) {
  HostSpan r;
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return r;
  }
  if (a.type == cudaMemoryTypeUnregistered) {
    r.kind = HostSpan::pageable;
    return r;
  }
  if (a.type != cudaMemoryTypeHost) return r;
  using Query = int (*)(void*, int, unsigned long long);  // cuPointerGetAttribute
  static Query query = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult st{};
    if (cudaGetDriverEntryPoint("cuPointerGetAttribute", &f, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      f = nullptr;
    }
    return reinterpret_cast<Query>(f);
  }();
  if (!query) return r;
  unsigned long long base = 0;
  size_t size = 0;
  const unsigned long long q = reinterpret_cast<uintptr_t>(p);
  if (query(&base, 11 /* CU_POINTER_ATTRIBUTE_RANGE_START_ADDR */, q) != 0 || query(&size, 12 /* CU_POINTER_ATTRIBUTE_RANGE_SIZE */, q) != 0 || !size) return r;
  r.kind = HostSpan::pinned;
  r.base = (uintptr_t)base;
  r.size = size;
  return r;
}

template <class After>
void flush_uploads_impl(arp_ctx* c, After&& after) {
  std::vector<UpItem>& it = c->up_items;
  if (it.empty()) return;
  constexpr size_t kGap = 4096;
  std::stable_sort(it.begin(), it.end(), [](const UpItem& a, const UpItem& b) { return a.h < b.h; });
  int waiting[3] = {0, 0, 0};  // arrays a marked kernel still waits for
  for (const UpItem& u : it) waiting[u.after] += 1;
  // runs of neighbours, by host address
  struct Run {
    size_t first, last, span;
  };
  std::vector<Run> runs;
  size_t arena = 0;
  // k_scatter_segments moves 4-byte words: a 16-bit array of odd length is moved with its last word completed by whatever
  // follows it in the landing area (the device buffers have room for it)
  const auto words = [](const UpItem& u) { return (reinterpret_cast<uintptr_t>(u.h) & 3u) == 0; };
  for (size_t k = 0; k < it.size();) {
    size_t j = k;
    HostSpan run_span;
    bool have_span = false;
    while (j + 1 < it.size() && j + 1 - k < (size_t)SEG_MAX && words(it[j]) && words(it[j + 1]) && it[j + 1].h >= it[j].h + it[j].bytes &&
           (size_t)(it[j + 1].h - (it[j].h + it[j].bytes)) < kGap) {
      if (!have_span) {
        run_span = host_span_of(it[k].h);
        have_span = true;
      }
      if (run_span.kind == HostSpan::other) break;
      const uintptr_t next_end = reinterpret_cast<uintptr_t>(it[j + 1].h) + it[j + 1].bytes;
      if (run_span.kind == HostSpan::pinned) {  // the next array must lie in the same page-locked allocation
        if (next_end > run_span.base + run_span.size) break;
      } else {  // pageable so far: the next one must be pageable too
        if (host_span_of(it[j + 1].h).kind != HostSpan::pageable || host_span_of(reinterpret_cast<const void*>(next_end - 1)).kind != HostSpan::pageable) break;
      }
      ++j;
    }
    const size_t span = (size_t)(it[j].h + it[j].bytes - it[k].h);
    runs.push_back(Run{k, j, span});
    if (j > k) arena += (span + 255) / 256 * 256;
    k = j + 1;
  }
  c->up_arena.need(std::max<size_t>(arena, 1));  // (grows between uploads only: a free here would wait for the device)
  size_t at = 0;
  for (const Run& r : runs) {
    if (r.first == r.last) {
      const UpItem& u = it[r.first];
      CK(cudaMemcpyAsync(u.d, u.h, u.bytes, cudaMemcpyDefault, c->stream));
    } else {
      unsigned char* land = c->up_arena.p + at;
      at += (r.span + 255) / 256 * 256;
      CK(cudaMemcpyAsync(land, it[r.first].h, r.span, cudaMemcpyDefault, c->stream));
      Segments S{};
      size_t most = 0;
      for (size_t k = r.first; k <= r.last; ++k) {
        const UpItem& u = it[k];
        S.src[S.n] = land + (u.h - it[r.first].h);
        S.dst[S.n] = u.d;
        S.bytes[S.n] = (u.bytes + 3u) & ~(size_t)3u;
        most = std::max(most, u.bytes);
        S.n += 1;
      }
      const unsigned bx = (unsigned)std::min<size_t>((most / 16 + 255) / 256 + 1, 4 * (size_t)c->sm_count);
      k_scatter_segments<<<dim3(bx, (unsigned)S.n), 256, 0, c->stream>>>(S);
      CK(cudaGetLastError());
    }
    c->up_copies += 1;
    for (size_t k = r.first; k <= r.last; ++k)
      if (it[k].after && --waiting[it[k].after] == 0) after(it[k].after);  // the last array that kernel waits for is on the stream
  }
  it.clear();
}
void flush_uploads(arp_ctx* c) {
  flush_uploads_impl(c, [c](int k) {
    if (c->up_after) c->up_after(k);
  });
}

// Concatenate one array of every structure into device memory.  `adjust(s)` is added to every element of
// structure s (index arrays that point into the concatenated job).  n == 1: straight from the caller's memory.
template <class T, class GetPtr, class GetLen>
void upload_concat(arp_ctx* c, DBuf<T>& dst, const arp_structure* const* s, int64_t n, GetPtr ptr, GetLen len, size_t extra_tail,
                   const std::vector<int64_t>* adjust, bool skip_negative) {
  size_t total = extra_tail;
  for (int64_t k = 0; k < n; ++k) total += (size_t)len(s[k]);
  dst.need(std::max<size_t>(total, 1));
  if (total == 0) return;
  if (n == 1 && !adjust) {  // queued: flush_uploads issues it, merged with its neighbours in host memory where it has any
    if (len(s[0]))
      c->up_items.push_back(UpItem{reinterpret_cast<const unsigned char*>(ptr(s[0])), reinterpret_cast<unsigned char*>(dst.p), (size_t)len(s[0]) * sizeof(T), 0});
    return;
  }
  flush_uploads(c);  // the staged path below copies on the stream itself: keep the order
  // the staging buffer is reused by the next array: wait until the previous copy has left it
  CK(cudaStreamSynchronize(c->stream));
  c->stage.need(total * sizeof(T));
  T* h = reinterpret_cast<T*>(c->stage.p);
  size_t off = 0;
  for (int64_t k = 0; k < n; ++k) {
    const size_t m = (size_t)len(s[k]);
    const T* src = ptr(s[k]);
    if (adjust && (*adjust)[k] != 0) {
      const T add = (T)(*adjust)[k];
      for (size_t i = 0; i < m; ++i) h[off + i] = (skip_negative && (int64_t)src[i] < 0 && std::is_signed<T>::value) ? src[i] : (T)(src[i] + add);
    } else if (m) {
      std::memcpy(h + off, src, m * sizeof(T));
    }
    off += m;
  }
  CK(cudaMemcpyAsync(dst.p, h, off * sizeof(T), cudaMemcpyDefault, c->stream));
}

void do_upload(arp_ctx* c, const arp_structure* const* s, int64_t n, bool one_shot = false) {
  arp::NvtxRange nvtx_("arp:upload");
  CK(cudaSetDevice(c->device));
  if (n < 1) throw arp::Error{ARP_ERR_ARG, "a job needs at least one structure"};
  c->have_job = c->have_run = false;
  c->job_gen += 1;  // a new job: a captured run no longer applies
  c->slab.own_atoms = -1;
  c->slab.have_extent = false;
  c->atom_begin.assign(n + 1, 0);
  c->ring_begin.assign(n + 1, 0);
  std::vector<int64_t> res_begin(n + 1, 0), h_begin(n + 1, 0);
  for (int64_t k = 0; k < n; ++k) {
    if (!s[k]) throw arp::Error{ARP_ERR_ARG, "null structure"};
    c->atom_begin[k + 1] = c->atom_begin[k] + s[k]->n_atoms;
    c->ring_begin[k + 1] = c->ring_begin[k] + s[k]->n_rings;
    res_begin[k + 1] = res_begin[k] + s[k]->n_residues;
    h_begin[k + 1] = h_begin[k] + s[k]->n_h;
  }
  c->hint_lo.assign(3 * n, 0.0);
  c->hint_hi.assign(3 * n, 0.0);
  c->have_hint = true;
  c->ignore_hint = false;
  for (int64_t k = 0; k < n; ++k) {
    if (s[k]->n_atoms > 0 && !s[k]->has_bbox) c->have_hint = false;
    for (int a = 0; a < 3; ++a) {
      c->hint_lo[3 * k + a] = s[k]->has_bbox ? s[k]->bbox_min[a] : 0.0;
      c->hint_hi[3 * k + a] = s[k]->has_bbox ? s[k]->bbox_max[a] : 0.0;
      if (s[k]->has_bbox && !(std::isfinite(s[k]->bbox_min[a]) && std::isfinite(s[k]->bbox_max[a]) && s[k]->bbox_min[a] <= s[k]->bbox_max[a]))
        c->have_hint = false;
    }
  }
  c->n_struct = n;
  c->n_atoms = c->atom_begin[n];
  c->n_rings = c->ring_begin[n];
  c->n_res = res_begin[n];
  c->n_h = h_begin[n];
  if (c->n_atoms >= ((int64_t)1 << 30) || c->n_rings >= ((int64_t)1 << 31) || c->n_h >= ((int64_t)1 << 32))
    throw arp::Error{ARP_ERR_LIMIT, "job exceeds 2^30 atoms / 2^31 rings"};

  CK(cudaEventRecord(c->ev[0], c->stream));
  c->h_words.need(32);
  bool side_used = false, use_21 = false, use_attr16 = false, milli = false;
  c->up_items.clear();
  c->up_copies = 0;
  c->h2d_pending = false;
  // the two expand kernels run on the side stream as soon as their input has arrived, beside the remaining copies
  c->up_after = [&](int which) {
    if (which == 1) {
      CK(cudaEventRecord(c->udep[0], c->stream));
      CK(cudaStreamWaitEvent(c->side, c->udep[0], 0));
      if (use_21 || use_attr16)
        k_expand_milli21<<<(unsigned)((c->n_atoms + 255) / 256), 256, 0, c->side>>>(use_21 ? reinterpret_cast<const unsigned long long*>(c->xyz21.p) : nullptr, c->xyz21_origin[0], c->xyz21_origin[1],
                                                                                  c->xyz21_origin[2], use_attr16 ? c->attr16.p : nullptr,
                                                                                  (uint32_t)c->n_atoms, c->x.p, c->y.p, c->z.p, c->attr.p);
      if (!use_21 && milli)
        k_expand_milli<<<(unsigned)((c->n_atoms + 255) / 256), 256, 0, c->side>>>(c->xm.p, c->ym.p, c->zm.p, (uint32_t)c->n_atoms, c->x.p, c->y.p, c->z.p);
    } else {
      CK(cudaEventRecord(c->udep[1], c->stream));
      CK(cudaStreamWaitEvent(c->side, c->udep[1], 0));
      k_expand_res_id<<<(unsigned)((c->n_res + 255) / 256), 256, 0, c->side>>>(c->res_atom_begin.p, (uint32_t)c->n_res, (uint32_t)c->n_atoms, c->res_id.p);  // a warp per 32 residues
    }
    CK(cudaGetLastError());
    side_used = true;
  };
  // a kernel that needs the array queued last: marked for flush_uploads, or launched now if the copy is already on the stream
  const auto then_kernel = [&](size_t queued_before, int which) {
    if (c->up_items.size() > queued_before)
      for (size_t k = queued_before; k < c->up_items.size(); ++k) c->up_items[k].after = which;
    else
      c->up_after(which);
  };
  auto nat = [](const arp_structure* t) { return t->n_atoms; };
  auto nrs = [](const arp_structure* t) { return t->n_residues; };
  auto nrg = [](const arp_structure* t) { return t->n_rings; };
  auto nrg3 = [](const arp_structure* t) { return 3 * t->n_rings; };
  auto nh = [](const arp_structure* t) { return t->n_h; };
#define UP(buf, field, lenf) upload_concat(c, c->buf, s, n, [](const arp_structure* t) { return t->field; }, lenf, 0, nullptr, false)
  // the most compact forms (single-structure jobs): 21-bit lattice words and 16-bit attribute words
  use_21 = n == 1 && c->n_atoms > 0 && s[0]->xyz_milli21 != nullptr;
  use_attr16 = n == 1 && c->n_atoms > 0 && s[0]->attr16 != nullptr;
  milli = c->n_atoms > 0;
  for (int64_t k = 0; k < n; ++k)
    if (!s[k]->x_milli || !s[k]->y_milli || !s[k]->z_milli) milli = false;
  if (use_21 || use_attr16 || milli) {  // on-lattice coordinates: a half or a third of the bytes over PCIe, f64 rebuilt here
    c->x.need(c->n_atoms);
    c->y.need(c->n_atoms);
    c->z.need(c->n_atoms);
    c->attr.need(std::max<int64_t>(c->n_atoms, 1));
    const size_t q = c->up_items.size();
    if (use_21) {
      for (int a3 = 0; a3 < 3; ++a3) c->xyz21_origin[a3] = s[0]->xyz_milli21_origin[a3];
      UP(xyz21, xyz_milli21, nat);
    } else if (milli) {
      UP(xm, x_milli, nat);
      UP(ym, y_milli, nat);
      UP(zm, z_milli, nat);
    }
    if (use_attr16) UP(attr16, attr16, nat);
    then_kernel(q, 1);  // on the side stream: the next copies do not wait for the kernel
  }
  if (!use_21 && !milli) {
    for (int64_t k = 0; k < n; ++k)
      if (s[k]->n_atoms > 0 && (!s[k]->x || !s[k]->y || !s[k]->z)) throw arp::Error{ARP_ERR_ARG, "structure has neither x / y / z nor x_milli / y_milli / z_milli"};
    UP(x, x, nat);
    UP(y, y, nat);
    UP(z, z, nat);
  }
  if (!use_attr16) {
    for (int64_t k = 0; k < n; ++k)
      if (s[k]->n_atoms > 0 && !s[k]->attr) throw arp::Error{ARP_ERR_ARG, "structure has neither attr nor attr16"};
    UP(attr, attr, nat);
  }
  c->per_residue_keys = true;
  for (int64_t k = 0; k < n; ++k)
    if (!s[k]->res_chain_key || !s[k]->res_res_ordinal) c->per_residue_keys = false;
  if (c->per_residue_keys) {
    UP(res_chain_key, res_chain_key, nrs);
    UP(res_res_ordinal, res_res_ordinal, nrs);
  } else {
    for (int64_t k = 0; k < n; ++k)
      if (s[k]->n_atoms > 0 && (!s[k]->chain_key || !s[k]->res_ordinal))
        throw arp::Error{ARP_ERR_ARG, "structure has neither per-atom nor per-residue chain_key / res_ordinal"};
    UP(chain_key, chain_key, nat);
    UP(res_ordinal, res_ordinal, nat);
  }
  UP(hx, h_x, nh);
  UP(hy, h_y, nh);
  UP(hz, h_z, nh);
  UP(ring_c, ring_center, nrg3);
  UP(ring_n, ring_normal, nrg3);
  UP(ring_attr, ring_attr, nrg);
  UP(ring_chain_key, ring_chain_key, nrg);
  UP(ring_res_ordinal, ring_res_ordinal, nrg);
#undef UP
  std::vector<int64_t> res_adj(res_begin.begin(), res_begin.end() - 1), atom_adj(c->atom_begin.begin(), c->atom_begin.end() - 1),
      h_adj(h_begin.begin(), h_begin.end() - 1);
  const bool multi = n > 1;
  bool expand_res_id = c->n_res > 0;
  for (int64_t k = 0; k < n; ++k)
    if (!s[k]->res_atom_begin) expand_res_id = false;
  c->have_res_csr = expand_res_id;
  if (expand_res_id) {
    // like res_h_begin below: every structure contributes its first n_residues offsets, the tail closes the CSR
    const size_t q = c->up_items.size();
    upload_concat(c, c->res_atom_begin, s, n, [](const arp_structure* t) { return t->res_atom_begin; }, nrs, 1, multi ? &atom_adj : nullptr, false);
    // the closing word comes from a page-locked slot of its own (nothing else writes it during this upload: no wait)
    c->h_words.p[0] = (uint32_t)c->n_atoms;
    CK(cudaMemcpyAsync(c->res_atom_begin.p + c->n_res, c->h_words.p, 4, cudaMemcpyDefault, c->stream));
    c->res_id.need(std::max<int64_t>(c->n_atoms, 1));
    if (c->n_atoms > 0) then_kernel(q, 2);  // on the side stream: the copy engine keeps going underneath
  } else {
    for (int64_t k = 0; k < n; ++k)
      if (s[k]->n_atoms > 0 && !s[k]->res_id) throw arp::Error{ARP_ERR_ARG, "structure has neither res_id nor res_atom_begin"};
    upload_concat(c, c->res_id, s, n, [](const arp_structure* t) { return t->res_id; }, nat, 0, multi ? &res_adj : nullptr, false);
  }
  upload_concat(c, c->res_cb, s, n, [](const arp_structure* t) { return t->res_cb; }, nrs, 0, multi ? &atom_adj : nullptr, true);
  upload_concat(c, c->res_sg, s, n, [](const arp_structure* t) { return t->res_sg; }, nrs, 0, multi ? &atom_adj : nullptr, true);
  // res_h_begin: n_res + 1 entries; each structure contributes its first n_residues offsets, the tail closes the CSR
  upload_concat(c, c->res_h_begin, s, n, [](const arp_structure* t) { return t->res_h_begin; }, nrs, 1, multi ? &h_adj : nullptr, false);
  {
    c->h_words.p[16] = (uint32_t)c->n_h;  // its own page-locked slot, see above
    CK(cudaMemcpyAsync(c->res_h_begin.p + c->n_res, c->h_words.p + 16, 4, cudaMemcpyDefault, c->stream));
  }
  // optional global ids: rows then carry them (arp_structure.atom_gid / ring_gid)
  c->have_gid = true;
  for (int64_t k = 0; k < n; ++k)
    if ((s[k]->n_atoms > 0 && !s[k]->atom_gid) || (s[k]->n_rings > 0 && !s[k]->ring_gid)) c->have_gid = false;
  if (c->have_gid) {
    upload_concat(c, c->atom_gid, s, n, [](const arp_structure* t) { return t->atom_gid; }, nat, 0, nullptr, false);
    upload_concat(c, c->ring_gid, s, n, [](const arp_structure* t) { return t->ring_gid; }, nrg, 0, nullptr, false);
  }
  // residue tags: the caller's, or derived from (chain key, ordinal) of neighbouring residues (hierarchy order assumed)
  c->have_tag = c->n_res > 0;
  for (int64_t k = 0; k < n; ++k)
    if (s[k]->n_residues > 0 && !s[k]->res_tag) c->have_tag = false;
  bool derive_tags = false;
  if (c->have_tag) upload_concat(c, c->res_tag, s, n, [](const arp_structure* t) { return t->res_tag; }, nrs, 0, nullptr, false);
  flush_uploads(c);  // every array of the job is on the stream from here on
  if (!c->have_tag && c->n_res > 0) {
    derive_tags = true;
    if (!c->per_residue_keys) {  // per-atom keys only: the residues' keys first (every atom of a residue carries the same pair)
      c->res_chain_key.need((size_t)c->n_res);
      c->res_res_ordinal.need((size_t)c->n_res);
      if (c->n_atoms > 0) {
        k_res_keys_from_atoms<<<(unsigned)((c->n_atoms + 255) / 256), 256, 0, c->stream>>>(c->chain_key.p, c->res_ordinal.p, c->res_id.p, (uint32_t)c->n_atoms,
                                                                                         c->res_chain_key.p, c->res_res_ordinal.p);
        CK(cudaGetLastError());
      }
    }
    c->res_tag.need((size_t)c->n_res + 1);
    c->struct_res_begin.need((size_t)n + 1);
    c->up_err.need(1);
    c->h_up_err.need(1);
    CK(cudaMemsetAsync(c->up_err.p, 0, sizeof(unsigned int), c->stream));
    {
      std::vector<uint32_t> rb(res_begin.begin(), res_begin.end());
      CK(cudaStreamSynchronize(c->stream));  // the staging buffer is reused
      c->stage.need(rb.size() * sizeof(uint32_t));
      std::memcpy(c->stage.p, rb.data(), rb.size() * sizeof(uint32_t));
      CK(cudaMemcpyAsync(c->struct_res_begin.p, c->stage.p, rb.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    }
    k_res_tag_steps<<<(unsigned)((c->n_res + 255) / 256), 256, 0, c->stream>>>(c->res_chain_key.p, c->res_res_ordinal.p, c->struct_res_begin.p,
                                                                              (uint32_t)n, (uint32_t)c->n_res, c->res_tag.p, c->up_err.p);
    CK(cudaGetLastError());
    {
      const uint32_t tiles = ((uint32_t)c->n_res + SCAN_TILE - 1) / SCAN_TILE;
      c->scan_status.need(tiles + 1);
      CK(cudaMemsetAsync(c->scan_status.p, 0, (size_t)(tiles + 1) * sizeof(unsigned long long), c->stream));
      k_scan_lookback<<<tiles, SCAN_THREADS, 0, c->stream>>>(c->res_tag.p, (uint32_t)c->n_res, c->scan_status.p, tiles, c->up_err.p);
      CK(cudaGetLastError());
    }
    CK(cudaMemcpyAsync(c->h_up_err.p, c->up_err.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, c->stream));
  }
  // structure descriptors (grids are filled in by the run, they depend on the cutoff)
  c->h_descs.assign(n, StructDesc{});
  for (int64_t k = 0; k < n; ++k) {
    c->h_descs[k].atom_begin = (uint32_t)c->atom_begin[k];
    c->h_descs[k].atom_end = (uint32_t)c->atom_begin[k + 1];
    c->h_descs[k].ring_begin = (uint32_t)c->ring_begin[k];
    c->h_descs[k].ring_end = (uint32_t)c->ring_begin[k + 1];
  }
  c->descs.need(n);
  c->h_descs_up.need((size_t)n);  // page-locked, this upload's own (h_descs_pin belongs to the run: it compares against it)
  std::memcpy(c->h_descs_up.p, c->h_descs.data(), n * sizeof(StructDesc));
  CK(cudaMemcpyAsync(c->descs.p, c->h_descs_up.p, n * sizeof(StructDesc), cudaMemcpyDefault, c->stream));
  if (side_used) {  // join
    CK(cudaEventRecord(c->udep[2], c->side));
    CK(cudaStreamWaitEvent(c->stream, c->udep[2], 0));
  }
  if (multi) {
    c->atom_struct.need(std::max<int64_t>(c->n_atoms, 1));
    c->ring_struct.need(std::max<int64_t>(c->n_rings, 1));
    k_fill_struct<<<(unsigned)n, 256, 0, c->stream>>>(c->descs.p, c->atom_struct.p, c->ring_struct.p);
    CK(cudaGetLastError());
  }
  CK(cudaEventRecord(c->ev[1], c->stream));
  c->up_after = nullptr;
  if (one_shot && !derive_tags) {
    // arp_contacts*: the caller's arrays stay valid until the call returns, and it returns after the download -- the
    // run is queued behind the copies without a host round trip in between
    c->h2d_pending = true;
  } else {
    CK(cudaStreamSynchronize(c->stream));  // arp_stage_upload: the caller may reuse its arrays once this returns
    CK(cudaEventElapsedTime(&c->ms_h2d, c->ev[0], c->ev[1]));
  }
  c->tags_ok = c->n_res > 0;
  if (derive_tags && *c->h_up_err.p) {
    if (*c->h_up_err.p & 4u) throw arp::Error{ARP_ERR_CUDA, "residue-tag scan: look-back wait exceeded its bound"};
    // residues are not in atoms_with_hierarchy() order (a chain's ordinals do not increase) and the caller gave no
    // res_tag: the emit-radius screen, which needs the tags, is not used for this job (rows are the same either way)
    c->tags_ok = false;
  }
  c->have_job = true;
}

}  // namespace

#include "api_run.inc"
#include "api_nsum.inc"
#include "api_slab.inc"
#include "api_static.inc"
