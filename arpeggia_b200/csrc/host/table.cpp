// arph_build_table: device rows -> the reference's 20-column frame.
//   results_to_df                      src/contacts/mod.rs:140-181   (u32 / i32 narrowing, f64 -> f32 distance)
//   collect_sc_stats + sc_results_to_df src/contacts/complex.rs:137-174, mod.rs:184-214
//   concat + left join + sort           src/contacts/mod.rs:111-136
// The left join is evaluated per row (both residues' side-chain planes looked up directly); the sort runs on
// integer ranks of the dictionary strings, so no string is compared or materialised per row.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <vector>

#include "packed.hpp"

namespace arp {

struct TableStore {
  arph_table t{};  // must stay the first member (arph_table_free casts back)
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
}  // namespace

arph_table* build_table(const Packed& P, const arp_row* rows, int64_t n) {
  auto* T = new TableStore();
  try {
    T->strings = P.strings;
    // rank of every dictionary string in byte order (polars sorts strings bytewise)
    const size_t ns = P.strings.size();
    std::vector<int32_t> by_rank(ns), srank(ns);
    std::iota(by_rank.begin(), by_rank.end(), 0);
    std::sort(by_rank.begin(), by_rank.end(), [&](int32_t a, int32_t b) { return P.strings[a] < P.strings[b]; });
    for (size_t r = 0; r < ns; ++r) srank[by_rank[r]] = (int32_t)r;
    int32_t lrank[ARP_N_LABELS];
    {
      int order[ARP_N_LABELS];
      std::iota(order, order + ARP_N_LABELS, 0);
      std::sort(order, order + ARP_N_LABELS, [](int a, int b) { return std::strcmp(arp_label_name(a), arp_label_name(b)) < 0; });
      for (int r = 0; r < ARP_N_LABELS; ++r) lrank[order[r]] = r;
    }
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
      if (fa.icode != fb.icode) return srank[fa.icode] < srank[fb.icode];
      if (ta.icode != tb.icode) return srank[ta.icode] < srank[tb.icode];
      if (F[ia].atomn != F[ib].atomn) return srank[F[ia].atomn] < srank[F[ib].atomn];
      if (G[ia].atomn != G[ib].atomn) return srank[G[ia].atomn] < srank[G[ib].atomn];
      if (fa.resn != fb.resn) return srank[fa.resn] < srank[fb.resn];
      if (ta.resn != tb.resn) return srank[ta.resn] < srank[tb.resn];
      uint32_t da, db;
      std::memcpy(&da, &rows[ia].distance, 4);
      std::memcpy(&db, &rows[ib].distance, 4);
      return da < db;
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
  for (auto& s : T->strings) T->cstrs.push_back(s.c_str());
  t.n_strings = (int32_t)T->cstrs.size();
  t.strings = T->cstrs.data();
  t.n_atoms = (int64_t)P.x.size();
  return &T->t;
}

}  // namespace arp

extern "C" {
int arph_build_table(const arph_packed* p, const arp_row* rows, int64_t n_rows, arph_table** out) {
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
void arph_table_free(arph_table* t) {
  if (t) delete reinterpret_cast<arp::TableStore*>(t);
}
}
