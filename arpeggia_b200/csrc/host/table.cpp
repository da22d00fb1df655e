// arph_build_table: device rows -> the reference's 20-column frame.
//   results_to_df                      src/contacts/mod.rs:140-181   (u32 / i32 narrowing, f64 -> f32 distance)
//   collect_sc_stats + sc_results_to_df src/contacts/complex.rs:137-174, mod.rs:184-214
//   concat + left join + sort           src/contacts/mod.rs:111-136
// The left join is evaluated per row (both residues' side-chain planes looked up directly); the sort runs on
// integer ranks of the dictionary strings, so no string is compared or materialised per row.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <numeric>
#include <vector>

#include "packed.hpp"

namespace arp {

// Process-wide pool of column blocks for tables gathered on the host (table_from_frame on a compact frame): a fresh
// 50 MB allocation pays its page faults on every call, a block that went round once does not.
struct ColumnBlock {
  unsigned char* p = nullptr;
  size_t cap = 0;
};
namespace {
std::mutex g_block_mu;
std::vector<ColumnBlock> g_blocks;
ColumnBlock block_acquire(size_t bytes) {
  {
    std::lock_guard<std::mutex> g(g_block_mu);
    int best = -1;
    for (int k = 0; k < (int)g_blocks.size(); ++k)
      if (g_blocks[k].cap >= bytes && (best < 0 || g_blocks[k].cap < g_blocks[best].cap)) best = k;
    if (best >= 0) {
      ColumnBlock b = g_blocks[best];
      g_blocks.erase(g_blocks.begin() + best);
      return b;
    }
  }
  ColumnBlock b;
  b.cap = bytes + bytes / 8 + 4096;
  b.p = static_cast<unsigned char*>(std::aligned_alloc(256, (b.cap + 255) / 256 * 256));
  if (!b.p) throw Error{ARP_ERR_OOM, "out of host memory for the table's columns"};
  return b;
}
void block_release(ColumnBlock b) {
  if (!b.p) return;
  {
    std::lock_guard<std::mutex> g(g_block_mu);
    if (g_blocks.size() < 2) {
      g_blocks.push_back(b);
      return;
    }
  }
  std::free(b.p);
}
}  // namespace

struct TableStore {
  arph_table t{};  // must stay the first member (arph_table_free casts back)
  arp_frame* frame = nullptr;  // device-built tables borrow the frame's page-locked columns
  ColumnBlock block;           // ... and hold the columns gathered on the host (compact frames) + the validity bitmap
  std::vector<uint8_t> validity;
  ~TableStore() { block_release(block); }
  std::vector<uint32_t> model;
  std::vector<uint8_t> interaction, sc_valid;
  std::vector<float> distance, scd, sch, sca;
  std::vector<int32_t> col[14];
  std::vector<std::string> strings;
  std::vector<const char*> cstrs;
};

namespace {
int32_t narrow(int64_t v, const char* what) {  // mod.rs:150-178
  if (v > INT32_MAX || v < INT32_MIN) throw Error{ARP_ERR_NARROWING, std::string(what) + " " + std::to_string(v) + " exceeds i32 limits"};
  return (int32_t)v;
}
bool fits_i32(int64_t v) { return v <= INT32_MAX && v >= INT32_MIN; }

// rank of every dictionary string in byte order (polars sorts strings bytewise)
std::vector<int32_t> string_ranks(const Packed& P) {
  const size_t ns = P.strings.size();
  std::vector<int32_t> by_rank(ns), srank(ns);
  std::iota(by_rank.begin(), by_rank.end(), 0);
  std::sort(by_rank.begin(), by_rank.end(), [&](int32_t a, int32_t b) { return P.strings[a] < P.strings[b]; });
  for (size_t r = 0; r < ns; ++r) srank[by_rank[r]] = (int32_t)r;
  return srank;
}
void label_ranks(int32_t* lrank) {  // the interaction column is a string: sort key 10 orders by the label's name
  int order[ARP_N_LABELS];
  std::iota(order, order + ARP_N_LABELS, 0);
  std::sort(order, order + ARP_N_LABELS, [](int a, int b) { return std::strcmp(arp_label_name(a), arp_label_name(b)) < 0; });
  for (int r = 0; r < ARP_N_LABELS; ++r) lrank[order[r]] = r;
}
}  // namespace

// The integer side tables arp_stage_frame sorts and decodes with (include/arpeggia_cuda.h).  Entities = atoms, then
// rings; ent_rank orders (resi, altloc, atomi) = sort keys 4-6 / 7-9 of mod.rs:123-134, ent_tie breaks the ties the
// reference leaves open by entity order -- build_table below uses the same order, so both paths give the same frame.
void build_frame_tables(const Packed& P, FrameTables& T) {
  NvtxRange nvtx_("arph:frame_tables");
  PhaseClock clk;
  const std::vector<int32_t> srank = string_ranks(P);
  const size_t n_sides = P.resconfs.size(), n_atoms = P.x.size(), n_rings = P.ring_rc.size(), E = n_atoms + n_rings;
  T.sides.assign(n_sides, arp_frame_side{});
  T.side_sc.assign(6 * n_sides, 0.0);
  std::vector<uint8_t> side_bad(n_sides, 0);
  {  // chain_rank: dense rank of (model serial, chain id bytes)
    std::vector<std::pair<int64_t, int32_t>> keys(n_sides);
#pragma omp parallel for schedule(static) if (n_sides > 16384)
    for (size_t k = 0; k < n_sides; ++k) keys[k] = {P.resconfs[k].model_serial, srank[P.resconfs[k].chain]};
    // the distinct (model, chain) pairs: residues of a chain follow one another, so the list of pairs that differ from
    // their predecessor is short; sorted and made unique it is the rank table
    std::vector<std::pair<int64_t, int32_t>> uniq;
    for (size_t k = 0; k < n_sides; ++k)
      if (k == 0 || keys[k] != keys[k - 1]) uniq.push_back(keys[k]);
    std::sort(uniq.begin(), uniq.end());
    uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
    T.t.n_chain_ranks = (uint32_t)std::max<size_t>(uniq.size(), 1);
#pragma omp parallel for schedule(static) if (n_sides > 16384)
    for (size_t k = 0; k < n_sides; ++k) {
      const Packed::ResConf& r = P.resconfs[k];
      arp_frame_side& s = T.sides[k];
      s.chain = r.chain, s.resn = r.resn, s.icode = r.icode, s.altloc = r.altloc;
      side_bad[k] = !fits_i32(r.resi) || r.model_serial < 0 || r.model_serial > (int64_t)UINT32_MAX;
      s.resi = side_bad[k] ? 0 : (int32_t)r.resi;
      s.model = side_bad[k] ? 0u : (uint32_t)r.model_serial;
      s.chain_rank = (uint32_t)(std::lower_bound(uniq.begin(), uniq.end(), keys[k]) - uniq.begin());
      s.has_sc = r.has_sc ? 1u : 0u;
      if (r.has_sc)
        for (int a = 0; a < 3; ++a) T.side_sc[6 * k + a] = r.sc.c[a], T.side_sc[6 * k + 3 + a] = r.sc.n[a];
    }
  }
  clk.mark("t1 sides");
  T.ent_side.resize(E);
  T.ent_atomn.resize(E);
  T.ent_atomi.resize(E);
  T.ent_rank.resize(E);
  T.ent_tie.resize(E);
  // ent_rank / ent_tie are only ever compared between entities of the same chain (the chain ranks precede them in the
  // key), so both are counted per chain: small numbers, few key bits, few sort passes
  struct Key {
    int64_t resi, atomi;
    uint32_t chain;
    int32_t alt;
    uint32_t id;
  };
  std::unique_ptr<Key[]> keys_store(new Key[std::max<size_t>(E, 1)]);  // (no zero fill: every element is written below)
  Key* keys = keys_store.get();
#pragma omp parallel for schedule(static) if (E > 65536)
  for (size_t e = 0; e < E; ++e) {
    const bool ring = e >= n_atoms;
    const uint32_t rc = ring ? P.ring_rc[e - n_atoms] : P.atom_rc[e];
    const int64_t serial = ring ? 0 : P.atom_serial[e];  // InteractingEntity::new(.., "Ring", 0), complex.rs:334-342
    T.ent_side[e] = rc;
    T.ent_atomn[e] = ring ? P.ring_name : P.atom_name[e];
    const bool bad = !fits_i32(serial) || side_bad[rc];
    T.ent_atomi[e] = fits_i32(serial) ? (int32_t)serial : 0;
    T.ent_tie[e] = bad ? ARP_ENT_NARROW : 0u;
    keys[e] = Key{P.resconfs[rc].resi, serial, T.sides[rc].chain_rank, srank[P.resconfs[rc].altloc], (uint32_t)e};
  }
  clk.mark("t2 entity keys");
  const auto key_less = [](const Key& a, const Key& b) {  // a total order (id breaks every tie): one result on any number of threads
    if (a.chain != b.chain) return a.chain < b.chain;
    if (a.resi != b.resi) return a.resi < b.resi;
    if (a.alt != b.alt) return a.alt < b.alt;
    if (a.atomi != b.atomi) return a.atomi < b.atomi;
    return a.id < b.id;
  };
  const auto sort_range = [&](Key* b, Key* e) {
    if (!std::is_sorted(b, e, key_less)) std::sort(b, e, key_less);
  };
  const size_t n_chains = T.t.n_chain_ranks;
  uint32_t max_tie = 0, max_rank = 0;
  // ranks of one chain from its atoms [ab, ae) and its rings [rb, re), each in key order: a merge walk, nothing is moved
  const auto rank_chain = [&](const Key* ab, const Key* ae, const Key* rb, const Key* re, uint32_t& top_rank, uint32_t& top_tie) {
    uint32_t rank = 0, tie = 0;
    const Key* prev = nullptr;
    while (ab != ae || rb != re) {
      const Key* k = (rb == re || (ab != ae && !key_less(*rb, *ab))) ? ab++ : rb++;
      if (prev) {
        if (k->resi == prev->resi && k->alt == prev->alt && k->atomi == prev->atomi) {
          tie += 1;
        } else {
          rank += 1;
          tie = 0;
        }
      }
      T.ent_rank[k->id] = rank;
      T.ent_tie[k->id] |= tie;
      top_rank = std::max(top_rank, rank);
      top_tie = std::max(top_tie, tie);
      prev = k;
    }
  };
  // The packer lays the atoms out chain by chain (atoms_with_hierarchy() order) and the rings likewise: a chain is one run
  // of atoms and at most one run of rings.  Each chain is then ranked where it lies -- its runs put in key order if they
  // are not yet (within a chain the atoms come residue by residue, serials rising: usually sorted already) -- chains in
  // parallel.  Anything else (a chain in several runs) goes through a counting sort by chain first.
  struct Run {
    size_t begin, end;
  };
  std::vector<Run> atom_run(n_chains, Run{0, 0}), ring_run(n_chains, Run{0, 0});
  bool one_run_each = true;
  {
    const auto find_runs = [&](size_t lo, size_t hi, std::vector<Run>& out) {
      for (size_t e = lo; e < hi && one_run_each;) {
        size_t f = e + 1;
        const uint32_t c = keys[e].chain;
        while (f < hi && keys[f].chain == c) ++f;
        if (out[c].end != out[c].begin) one_run_each = false;  // the chain has appeared before
        out[c] = Run{e, f};
        e = f;
      }
    };
    find_runs(0, n_atoms, atom_run);
    find_runs(n_atoms, E, ring_run);
    if (std::getenv("ARPH_TABLES_BUCKET")) one_run_each = false;  // tests: the general path on any input
  }
  clk.mark("t3 chain runs");
  if (one_run_each) {
#pragma omp parallel for schedule(dynamic, 1) reduction(max : max_rank, max_tie) if (E > 65536 && n_chains > 1)
    for (size_t c = 0; c < n_chains; ++c) {
      sort_range(keys + atom_run[c].begin, keys + atom_run[c].end);
      sort_range(keys + ring_run[c].begin, keys + ring_run[c].end);
      uint32_t top_rank = 0, top_tie = 0;
      rank_chain(keys + atom_run[c].begin, keys + atom_run[c].end, keys + ring_run[c].begin, keys + ring_run[c].end, top_rank, top_tie);
      max_rank = std::max(max_rank, top_rank);
      max_tie = std::max(max_tie, top_tie);
    }
  } else {
    std::vector<size_t> begin(n_chains + 1, 0);
    for (size_t e = 0; e < E; ++e) begin[keys[e].chain + 1] += 1;
    for (size_t c = 0; c < n_chains; ++c) begin[c + 1] += begin[c];
    std::vector<Key> by_chain(E);
    {
      std::vector<size_t> at(begin.begin(), begin.end() - 1);
      for (size_t e = 0; e < E; ++e) by_chain[at[keys[e].chain]++] = keys[e];
    }
#pragma omp parallel for schedule(dynamic, 1) reduction(max : max_rank, max_tie) if (E > 65536 && n_chains > 1)
    for (size_t c = 0; c < n_chains; ++c) {
      Key *b = by_chain.data() + begin[c], *e = by_chain.data() + begin[c + 1];
      Key* mid = std::find_if(b, e, [&](const Key& k) { return k.id >= n_atoms; });  // first ring of the chain (the sort is stable)
      sort_range(b, mid);
      sort_range(mid, e);
      uint32_t top_rank = 0, top_tie = 0;
      rank_chain(b, mid, mid, e, top_rank, top_tie);
      max_rank = std::max(max_rank, top_rank);
      max_tie = std::max(max_tie, top_tie);
    }
  }
  clk.mark("t5 ranks");
  int32_t lrank[ARP_N_LABELS];
  label_ranks(lrank);
  std::memset(T.t.label_rank, 0, sizeof T.t.label_rank);
  for (int l = 0; l < ARP_N_LABELS; ++l) T.t.label_rank[l] = (uint8_t)lrank[l];
  T.t.n_atoms = (int64_t)n_atoms;
  T.t.n_rings = (int64_t)n_rings;
  T.t.n_sides = (int64_t)n_sides;
  T.t.sides = T.sides.data();
  T.t.side_sc = T.side_sc.data();
  T.t.ent_side = T.ent_side.data();
  T.t.ent_atomn = T.ent_atomn.data();
  T.t.ent_atomi = T.ent_atomi.data();
  T.t.ent_rank = T.ent_rank.data();
  T.t.ent_tie = T.ent_tie.data();
  T.t.n_ent_ranks = max_rank + 1;
  T.t.max_tie = max_tie;
}

arph_table* build_table(const Packed& P, const arp_row* rows, int64_t n) {
  auto* T = new TableStore();
  try {
    T->strings = P.strings;
    const std::vector<int32_t> srank = string_ranks(P);
    int32_t lrank[ARP_N_LABELS];
    label_ranks(lrank);
    const int64_t n_atoms = (int64_t)P.x.size(), n_rings = (int64_t)P.ring_rc.size();

    struct Side {
      uint32_t rc;
      int32_t atomn, atomi;
    };
    auto side = [&](uint32_t v) -> Side {
      if (v & ARP_RING_FLAG) {  // InteractingEntity::new(.., "Ring", 0), complex.rs:334-342
        const uint32_t r = v & ~ARP_RING_FLAG;
        if ((int64_t)r >= n_rings) throw Error{ARP_ERR_ARG, "row references a ring outside the structure"};
        return Side{P.ring_rc[r], P.ring_name, 0};
      }
      if ((int64_t)v >= n_atoms) throw Error{ARP_ERR_ARG, "row references an atom outside the structure"};
      return Side{P.atom_rc[v], P.atom_name[v], narrow(P.atom_serial[v], "Atom index")};
    };
    auto entity_pos = [&](uint32_t v) -> uint64_t { return (v & ARP_RING_FLAG) ? (uint64_t)n_atoms + (v & ~ARP_RING_FLAG) : v; };
    std::vector<Side> F((size_t)n), G((size_t)n);
    for (int64_t k = 0; k < n; ++k) {
      F[k] = side(rows[k].from);
      G[k] = side(rows[k].to);
      if (rows[k].label >= (uint32_t)ARP_N_LABELS) throw Error{ARP_ERR_ARG, "row carries an unknown label"};
    }
    // mod.rs:123-134 keys, then the remaining columns so that the order is total and reproducible
    std::vector<int64_t> order((size_t)n);
    std::iota(order.begin(), order.end(), (int64_t)0);
    auto less = [&](int64_t ia, int64_t ib) {
      const Packed::ResConf &fa = P.resconfs[F[ia].rc], &fb = P.resconfs[F[ib].rc];
      const Packed::ResConf &ta = P.resconfs[G[ia].rc], &tb = P.resconfs[G[ib].rc];
      if (fa.model_serial != fb.model_serial) return fa.model_serial < fb.model_serial;
      if (fa.chain != fb.chain) return srank[fa.chain] < srank[fb.chain];
      if (ta.chain != tb.chain) return srank[ta.chain] < srank[tb.chain];
      if (fa.resi != fb.resi) return fa.resi < fb.resi;
      if (fa.altloc != fb.altloc) return srank[fa.altloc] < srank[fb.altloc];
      if (F[ia].atomi != F[ib].atomi) return F[ia].atomi < F[ib].atomi;
      if (ta.resi != tb.resi) return ta.resi < tb.resi;
      if (ta.altloc != tb.altloc) return srank[ta.altloc] < srank[tb.altloc];
      if (G[ia].atomi != G[ib].atomi) return G[ia].atomi < G[ib].atomi;
      if (rows[ia].label != rows[ib].label) return lrank[rows[ia].label] < lrank[rows[ib].label];
      // ties on all ten keys (unspecified in the reference): entity order, atoms before rings -- the order
      // arp_frame_tables.ent_tie encodes for the device sort
      const uint64_t pfa = entity_pos(rows[ia].from), pfb = entity_pos(rows[ib].from);
      if (pfa != pfb) return pfa < pfb;
      return entity_pos(rows[ia].to) < entity_pos(rows[ib].to);
    };
    std::sort(order.begin(), order.end(), less);

    T->model.resize(n);
    T->interaction.resize(n);
    T->sc_valid.resize(n);
    T->distance.resize(n);
    T->scd.resize(n);
    T->sch.resize(n);
    T->sca.resize(n);
    for (auto& c : T->col) c.resize(n);
    for (int64_t k = 0; k < n; ++k) {
      const int64_t src = order[k];
      const Side sides[2] = {F[src], G[src]};
      const Packed::ResConf& f = P.resconfs[sides[0].rc];
      const Packed::ResConf& g = P.resconfs[sides[1].rc];
      if (f.model_serial < 0 || f.model_serial > (int64_t)UINT32_MAX)
        throw Error{ARP_ERR_NARROWING, "Model number " + std::to_string(f.model_serial) + " exceeds u32 limits"};
      T->model[k] = (uint32_t)f.model_serial;
      T->interaction[k] = (uint8_t)rows[src].label;
      T->distance[k] = rows[src].distance;
      for (int s = 0; s < 2; ++s) {
        const Packed::ResConf& r = s ? g : f;
        T->col[s * 7 + 0][k] = r.chain;
        T->col[s * 7 + 1][k] = r.resn;
        T->col[s * 7 + 2][k] = narrow(r.resi, "Residue index");
        T->col[s * 7 + 3][k] = r.icode;
        T->col[s * 7 + 4][k] = r.altloc;
        T->col[s * 7 + 5][k] = sides[s].atomn;
        T->col[s * 7 + 6][k] = sides[s].atomi;
      }
      const bool ok = f.has_sc && g.has_sc;  // complex.rs:152-161
      T->sc_valid[k] = ok ? 1 : 0;
      if (ok) {
        T->scd[k] = (float)plane_point_dist(f.sc, g.sc.c);
        T->sch[k] = (float)plane_dihedral(f.sc.n, g.sc.n);
        T->sca[k] = (float)plane_point_angle(f.sc, g.sc.c);
      } else {
        T->scd[k] = T->sch[k] = T->sca[k] = std::nanf("");
      }
    }
  } catch (...) {
    delete T;
    throw;
  }
  arph_table& t = T->t;
  t.n_rows = n;
  t.model = T->model.data();
  t.interaction = T->interaction.data();
  t.distance = T->distance.data();
  int32_t** dst[14] = {&t.from_chain, &t.from_resn, &t.from_resi, &t.from_insertion, &t.from_altloc, &t.from_atomn, &t.from_atomi,
                       &t.to_chain,   &t.to_resn,   &t.to_resi,   &t.to_insertion,   &t.to_altloc,   &t.to_atomn,   &t.to_atomi};
  for (int k = 0; k < 14; ++k) *dst[k] = T->col[k].data();
  t.sc_centroid_dist = T->scd.data();
  t.sc_dihedral = T->sch.data();
  t.sc_centroid_angle = T->sca.data();
  t.sc_valid = T->sc_valid.data();
  T->validity.assign((size_t)(n + 7) / 8, 0);
  for (int64_t k = 0; k < n; ++k) {
    if (T->sc_valid[k]) T->validity[k >> 3] |= (uint8_t)(1u << (k & 7));
    else t.sc_null_count += 1;
  }
  t.sc_validity = T->validity.data();
  for (auto& s : T->strings) T->cstrs.push_back(s.c_str());
  t.n_strings = (int32_t)T->cstrs.size();
  t.strings = T->cstrs.data();
  t.n_atoms = (int64_t)P.x.size();
  return &T->t;
}

// A frame built on the device becomes the host table without copying a column: the table borrows the frame's
// page-locked arrays.  Rows the device marked sc_valid = 2 (an angle too close to an f32 rounding boundary for the
// device's acos, or NaN) are recomputed with the host libm, so the sc_* columns carry the same bits as build_table's.
// A COMPACT frame (arp_stage_frame_compact: no model / from_* / to_* columns) gets those fifteen columns here: pure
// gathers entity -> side -> field through the frame tables, rows split over the host's threads (mod.rs:140-181 builds
// the same columns from its row structs).  One pass does all of it, plus the Arrow validity bitmap of the sc_* columns.
arph_table* table_from_frame(const Packed& P, arp_frame* f, const arp_frame_tables* ft) {
  auto* T = new TableStore();
  try {
    T->strings = P.strings;
    const int64_t n = f->n_rows, n_atoms = (int64_t)P.x.size(), n_rings = (int64_t)P.ring_rc.size();
    const bool compact = !f->from_chain;
    FrameTables local;
    if (compact && n > 0 && !ft) {
      build_frame_tables(P, local);
      ft = &local.t;
    }
    if (compact && n > 0 && (ft->n_atoms != n_atoms || ft->n_rings != n_rings || !ft->ent_atomn || !ft->ent_atomi))
      throw Error{ARP_ERR_ARG, "frame tables do not describe this structure"};
    // one block: [15 gathered columns] + validity bitmap, every section 256-byte aligned
    const size_t col_bytes = ((size_t)n * 4 + 255) / 256 * 256, map_bytes = ((size_t)(n + 7) / 8 + 255) / 256 * 256;
    T->block = block_acquire((compact ? 15 * col_bytes : 0) + map_bytes + 256);
    unsigned char* base = T->block.p;
    uint32_t* model = f->model;
    int32_t* col[14] = {f->from_chain, f->from_resn, f->from_resi, f->from_insertion, f->from_altloc, f->from_atomn, f->from_atomi,
                        f->to_chain,   f->to_resn,   f->to_resi,   f->to_insertion,   f->to_altloc,   f->to_atomn,   f->to_atomi};
    if (compact) {
      model = reinterpret_cast<uint32_t*>(base);
      for (int k = 0; k < 14; ++k) col[k] = reinterpret_cast<int32_t*>(base + (size_t)(k + 1) * col_bytes);
      base += 15 * col_bytes;
    }
    uint8_t* bitmap = base;
    constexpr int64_t kChunk = 8192;  // a multiple of 8: every chunk owns whole bitmap bytes
    const int64_t n_chunks = (n + kChunk - 1) / kChunk;
    int64_t nulls = 0;
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(+ : nulls) reduction(| : bad) if (n > 4 * kChunk)
    for (int64_t c = 0; c < n_chunks; ++c) {
      const int64_t k0 = c * kChunk, k1 = std::min(n, k0 + kChunk);
      for (int64_t k = k0; k < k1; ++k) {
        const uint32_t v2[2] = {f->from[k], f->to[k]};
        uint32_t rc[2];
        bool ok = true;
        for (int s = 0; s < 2; ++s) {
          const uint32_t v = v2[s], i = v & ~ARP_RING_FLAG;
          const bool ring = (v & ARP_RING_FLAG) != 0;
          if (ring ? (int64_t)i >= n_rings : (int64_t)v >= n_atoms) {
            ok = false;
            break;
          }
          rc[s] = ring ? P.ring_rc[i] : P.atom_rc[v];
          if (compact) {
            const int64_t e = ring ? n_atoms + (int64_t)i : (int64_t)v;
            const arp_frame_side& sd = ft->sides[rc[s]];
            col[7 * s + 0][k] = sd.chain;
            col[7 * s + 1][k] = sd.resn;
            col[7 * s + 2][k] = sd.resi;
            col[7 * s + 3][k] = sd.icode;
            col[7 * s + 4][k] = sd.altloc;
            col[7 * s + 5][k] = ft->ent_atomn[e];
            col[7 * s + 6][k] = ft->ent_atomi[e];
            if (s == 0) model[k] = sd.model;
          }
        }
        if (!ok) {
          bad |= 1;
          continue;
        }
        if (f->sc_valid[k] == 2) {
          const Packed::ResConf& a = P.resconfs[rc[0]];
          const Packed::ResConf& b = P.resconfs[rc[1]];
          f->sc_dihedral[k] = (float)plane_dihedral(a.sc.n, b.sc.n);
          f->sc_centroid_angle[k] = (float)plane_point_angle(a.sc, b.sc.c);
          f->sc_valid[k] = 1;
        }
        // null statistics carry the same quiet-NaN bits as build_table's (the validity column says they are null)
        if (!f->sc_valid[k]) {
          f->sc_centroid_dist[k] = f->sc_dihedral[k] = f->sc_centroid_angle[k] = std::nanf("");
          nulls += 1;
        }
      }
      for (int64_t k = k0; k < k1; k += 8) {
        uint8_t byte = 0;
        for (int64_t j = k; j < std::min(k1, k + 8); ++j) byte |= (uint8_t)((f->sc_valid[j] ? 1u : 0u) << (j - k));
        bitmap[k >> 3] = byte;
      }
    }
    if (bad) throw Error{ARP_ERR_ARG, "frame row references an entity outside the structure"};
    arph_table& t = T->t;
    t.n_rows = n;
    t.model = model;
    t.from_chain = col[0], t.from_resn = col[1], t.from_resi = col[2], t.from_insertion = col[3];
    t.from_altloc = col[4], t.from_atomn = col[5], t.from_atomi = col[6];
    t.to_chain = col[7], t.to_resn = col[8], t.to_resi = col[9], t.to_insertion = col[10];
    t.to_altloc = col[11], t.to_atomn = col[12], t.to_atomi = col[13];
    t.sc_validity = bitmap;
    t.sc_null_count = nulls;
  } catch (...) {
    delete T;
    throw;
  }
  T->frame = f;
  arph_table& t = T->t;
  t.interaction = f->interaction;
  t.distance = f->distance;
  t.sc_centroid_dist = f->sc_centroid_dist;
  t.sc_dihedral = f->sc_dihedral;
  t.sc_centroid_angle = f->sc_centroid_angle;
  t.sc_valid = f->sc_valid;
  for (auto& s : T->strings) T->cstrs.push_back(s.c_str());
  t.n_strings = (int32_t)T->cstrs.size();
  t.strings = T->cstrs.data();
  t.n_atoms = (int64_t)P.x.size();
  t.pairs_in_cutoff = f->pairs_in_cutoff;
  t.pairs_classified = f->pairs_classified;
  return &T->t;
}

}  // namespace arp

extern "C" {
int arph_build_table(const arph_packed* p, const arp_row* rows, int64_t n_rows, arph_table** out) {
  arp::NvtxRange nvtx_("arph:build_table");
  *out = nullptr;
  try {
    *out = arp::build_table(arph_packed_ref(p), rows, n_rows);
    return ARP_OK;
  } catch (const arp::Error& e) {
    arp::last_error() = e.msg;
    return e.code;
  } catch (const std::exception& e) {
    arp::last_error() = e.what();
    return ARP_ERR_OOM;
  }
}
int arph_table_from_frame(const arph_packed* p, arp_frame* f, arph_table** out) {
  arp::NvtxRange nvtx_("arph:table_from_frame");
  *out = nullptr;
  if (!p || !f) {
    if (f) arp_frame_free(nullptr, f);
    arp::last_error() = "null argument";
    return ARP_ERR_ARG;
  }
  try {
    *out = arp::table_from_frame(arph_packed_ref(p), f, arph_packed_cached_frame_tables(p));
    return ARP_OK;
  } catch (const arp::Error& e) {
    arp_frame_free(nullptr, f);
    arp::last_error() = e.msg;
    return e.code;
  } catch (const std::exception& e) {
    arp_frame_free(nullptr, f);
    arp::last_error() = e.what();
    return ARP_ERR_OOM;
  }
}
void arph_table_free(arph_table* t) {
  if (!t) return;
  auto* T = reinterpret_cast<arp::TableStore*>(t);
  if (T->frame) arp_frame_free(nullptr, T->frame);
  delete T;
}
}
