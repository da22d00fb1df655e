// arph_pack: atom records -> device SoA (arp_structure) + the dictionaries needed to decode rows.
//
// Host equivalent of everything the reference does between reading the file and starting the search:
//   load_model residue filter                     src/utils.rs:60, src/contacts/residues.rs:131-161
//   optional zero-occupancy filter                src/python.rs:45-47, src/cli/contacts.rs:87-90
//   InteractionComplex::new                       src/contacts/complex.rs:36-68
//     parse_groups                                src/utils.rs:71-115
//     build_residue_index                         src/contacts/complex.rs:411-440
//     build_ring_positions / build_sc_plane_positions   src/contacts/complex.rs:442-514
// and the per-atom evaluation of the name tables (ionic.rs:83-99, hbond.rs:137-207, hydrophobic.rs:27-45,
// residues.rs:163-268, vdw.rs:50-53) into one attribute word per atom.
//
// The pdbtbx hierarchy (chain by id within a model, residue by (resi, icode) within a chain, conformer by
// (name, altloc) within a residue, blank-altloc atoms appended to every other conformer of the residue;
// SURVEY.md 8.2) is rebuilt with interned strings and flat index vectors.
#include <algorithm>
#include <cmath>
#include <array>
#include <map>
#include <set>
#include <unordered_map>
#include <vector>

#include "packed.hpp"

namespace arp {

namespace {

// ---- name tables ------------------------------------------------------------------------------------
enum : uint32_t {
  T_POS = 1u << 0, T_NEG = 1u << 1, T_ACC = 1u << 2, T_DON = 1u << 3, T_HYD = 1u << 4, T_RING = 1u << 5, T_SCP = 1u << 6
};
struct Rule {
  const char* res;
  const char* atoms;  // space separated
  uint32_t bits;
};
const Rule kRules[] = {
    // ionic.rs:83-99
    {"ARG", "NE CZ NH1 NH2", T_POS}, {"HIS", "CG ND1 CE1 NE2 CD2", T_POS}, {"LYS", "NZ", T_POS},
    {"ASP", "OD1 OD2", T_NEG}, {"GLU", "OE1 OE2", T_NEG},
    // hbond.rs:144-158 (side-chain acceptors; backbone O / OXT handled in code)
    {"ASN", "OD1", T_ACC}, {"ASP", "OD1 OD2", T_ACC}, {"GLN", "OE1", T_ACC}, {"GLU", "OE1 OE2", T_ACC},
    {"HIS", "ND1 NE2", T_ACC}, {"SER", "OG", T_ACC}, {"THR", "OG1", T_ACC}, {"TYR", "OH", T_ACC},
    {"MET", "SD", T_ACC}, {"CYS", "SG", T_ACC},
    // hbond.rs:167-180 (side-chain donors; backbone N handled in code)
    {"ARG", "NE NH1 NH2", T_DON}, {"ASN", "ND2", T_DON}, {"GLN", "NE2", T_DON}, {"HIS", "ND1 NE2", T_DON},
    {"LYS", "NZ", T_DON}, {"SER", "OG", T_DON}, {"THR", "OG1", T_DON}, {"TRP", "NE1", T_DON}, {"TYR", "OH", T_DON},
    {"CYS", "SG", T_DON},
    // hydrophobic.rs:32-44 (CB of everything but SER handled in code)
    {"ARG", "CG", T_HYD}, {"GLN", "CG", T_HYD}, {"GLU", "CG", T_HYD}, {"PRO", "CG", T_HYD},
    {"ILE", "CG1 CD1 CG2", T_HYD}, {"LEU", "CG CD1 CD2", T_HYD}, {"LYS", "CG CD", T_HYD}, {"MET", "CG CE SD", T_HYD},
    {"PHE", "CG CD1 CD2 CE1 CE2 CZ", T_HYD}, {"THR", "CG2", T_HYD}, {"TRP", "CG CD2 CE3 CZ3 CH2 CZ2", T_HYD},
    {"TYR", "CG CD1 CD2 CE1 CE2", T_HYD}, {"VAL", "CG1 CG2", T_HYD},
    // residues.rs:163-186 ring atoms
    {"HIS", "CG ND1 CE1 NE2 CD2", T_RING}, {"PHE", "CG CD1 CD2 CE1 CE2 CZ", T_RING}, {"TYR", "CG CD1 CD2 CE1 CE2 CZ", T_RING},
    {"TRP", "CG CD1 CD2 NE1 CE2 CE3 CZ2 CZ3 CH2", T_RING},
    // residues.rs:188-268 side-chain plane atoms
    {"ARG", "NE CZ NH1 NH2", T_SCP}, {"ASN", "CB CG OD1 ND2", T_SCP}, {"ASP", "CB CG OD1 OD2", T_SCP}, {"CYS", "CA CB SG", T_SCP},
    {"GLU", "CG CD OE1 OE2", T_SCP}, {"GLN", "CG CD OE1 NE2", T_SCP}, {"HIS", "CG ND1 CE1 NE2 CD2", T_SCP},
    {"ILE", "CB CG1 CG2 CD1", T_SCP}, {"LEU", "CB CG CD1 CD2", T_SCP}, {"LYS", "CG CD CE NZ", T_SCP}, {"MET", "CG SD CE", T_SCP},
    {"PHE", "CG CD1 CD2 CE1 CE2 CZ", T_SCP}, {"TYR", "CG CD1 CD2 CE1 CE2 CZ", T_SCP}, {"PRO", "N CA CB CG CD", T_SCP},
    {"SER", "CA CB OG", T_SCP}, {"THR", "CA CB OG1 CG2", T_SCP},
    {"TRP", "CG CD1 CD2 NE1 CE2 CE3 CZ2 CZ3 CH2", T_SCP}, {"VAL", "CA CB CG1 CG2", T_SCP},
};

const std::unordered_map<std::string, uint32_t>& rule_map() {
  static const std::unordered_map<std::string, uint32_t> m = [] {
    std::unordered_map<std::string, uint32_t> out;
    for (const Rule& r : kRules) {
      std::string atoms(r.atoms);
      size_t p = 0;
      while (p < atoms.size()) {
        size_t q = atoms.find(' ', p);
        if (q == std::string::npos) q = atoms.size();
        out[std::string(r.res) + "/" + atoms.substr(p, q - p)] |= r.bits;
        p = q + 1;
      }
    }
    return out;
  }();
  return m;
}
uint32_t table_bits(const std::string& res, const std::string& atom) {
  auto it = rule_map().find(res + "/" + atom);
  uint32_t b = it == rule_map().end() ? 0u : it->second;
  if (atom == "N") b |= T_DON;                                  // hbond.rs:164-166
  if ((atom == "O" || atom == "OXT") && res != "HOH") b |= T_ACC;  // hbond.rs:139-142
  if (atom == "CB" && res != "SER") b |= T_HYD;                 // hydrophobic.rs:29-31
  return b;
}

bool known_residue(std::string name) {  // residues.rs:131-161 (upper-cased name)
  static const std::set<std::string> k = {"ALA", "ARG", "ASN", "ASP", "CYS", "GLN", "GLU", "GLY", "HIS", "ILE", "LEU",
                                          "LYS", "MET", "PHE", "PRO", "SER", "THR", "TRP", "TYR", "VAL", "HOH"};
  for (auto& ch : name) ch = (char)std::toupper((unsigned char)ch);
  return k.count(name) != 0;
}

// ---- hierarchy with interned strings ----------------------------------------------------------------
struct ConfT {
  int32_t name, altloc;
  std::vector<int64_t> recs;
};
struct ResT {
  int64_t resi;
  int32_t icode;
  std::vector<ConfT> confs;
};
struct ResKeyHash {
  size_t operator()(const std::pair<int64_t, int32_t>& k) const { return std::hash<int64_t>()(k.first * 1315423911LL + k.second); }
};
struct ChainT {
  int32_t id;
  std::vector<ResT> residues;
  std::unordered_map<std::pair<int64_t, int32_t>, size_t, ResKeyHash> lookup;
};
struct ModelT {
  int64_t serial;
  std::vector<ChainT> chains;
  std::unordered_map<int32_t, size_t> lookup;
};

}  // namespace

int32_t Packed::intern(const std::string& s) {
  auto it = string_ids.find(s);
  if (it != string_ids.end()) return it->second;
  const int32_t id = (int32_t)strings.size();
  strings.push_back(s);
  string_ids.emplace(s, id);
  return id;
}

// utils.rs:71-115
void parse_groups(const std::set<std::string>& all, const std::string& groups, std::set<std::string>& lig, std::set<std::string>& rec) {
  std::vector<std::string> fields;
  for (size_t p = 0;;) {
    const size_t q = groups.find('/', p);
    fields.push_back(groups.substr(p, q == std::string::npos ? std::string::npos : q - p));
    if (q == std::string::npos) break;
    p = q + 1;
  }
  if (fields.size() < 2) throw Error{ARP_ERR_BAD_GROUPS, "Invalid chain groups format! Use '/' for all-to-all comparisons."};
  auto ids = [](const std::string& s) {
    std::set<std::string> out;
    for (size_t p = 0;;) {
      const size_t q = s.find(',', p);
      const std::string tok = s.substr(p, q == std::string::npos ? std::string::npos : q - p);
      if (!tok.empty()) out.insert(tok);
      if (q == std::string::npos) break;
      p = q + 1;
    }
    return out;
  };
  lig = ids(fields[0]);
  rec = ids(fields[1]);
  if (lig.empty() && rec.empty()) {
    lig = rec = all;
    return;
  }
  if (lig.empty())
    std::set_difference(all.begin(), all.end(), rec.begin(), rec.end(), std::inserter(lig, lig.end()));
  else if (rec.empty())
    std::set_difference(all.begin(), all.end(), lig.begin(), lig.end(), std::inserter(rec, rec.end()));
  if (lig.empty() || rec.empty()) throw Error{ARP_ERR_EMPTY_GROUP, "Empty chain groups!"};
}

static bool is_ring_residue_name(const std::string& s) { return s == "HIS" || s == "PHE" || s == "TYR" || s == "TRP"; }

void pack_records(const arp_atom_record* recs, int64_t n, const std::string& groups, bool ignore_zero_occupancy, uint32_t flags,
                  Packed& P) {
  if (n >= (int64_t)1 << 31) throw Error{ARP_ERR_LIMIT, "more than 2^31 atom records"};
  // ---- 1. group records into the hierarchy --------------------------------------------------------
  std::vector<ModelT> models;
  std::vector<int32_t> name_id((size_t)n);
  {
    struct NameSlot {
      uint64_t raw;
      int32_t id;
    };
    NameSlot name_cache[64];
    for (auto& e : name_cache) e = NameSlot{0, -1};
    int64_t mi = -1, ci = -1, ri = -1;
    // one-entry caches so that runs of equal fields cost no hashing
    char last_chain[8] = {1}, last_resn[8] = {1}, last_alt[2] = {1}, last_ic[2] = {1};
    int32_t chain_id = -1, resn_id = -1, alt_id = -1, ic_id = -1;
    for (int64_t k = 0; k < n; ++k) {
      const arp_atom_record& r = recs[k];
      if (std::memcmp(last_chain, r.chain, 8) != 0 || chain_id < 0) {
        std::memcpy(last_chain, r.chain, 8);
        chain_id = P.intern(fixed_field(r.chain, 8));
      }
      if (std::memcmp(last_resn, r.resn, 8) != 0 || resn_id < 0) {
        std::memcpy(last_resn, r.resn, 8);
        resn_id = P.intern(fixed_field(r.resn, 8));
      }
      if (std::memcmp(last_alt, r.altloc, 2) != 0 || alt_id < 0) {
        std::memcpy(last_alt, r.altloc, 2);
        alt_id = P.intern(fixed_field(r.altloc, 2));
      }
      if (std::memcmp(last_ic, r.icode, 2) != 0 || ic_id < 0) {
        std::memcpy(last_ic, r.icode, 2);
        ic_id = P.intern(fixed_field(r.icode, 2));
      }
      if (mi < 0 || models[mi].serial != r.model) {
        mi = -1;
        for (size_t m = 0; m < models.size(); ++m)
          if (models[m].serial == r.model) mi = (int64_t)m;
        if (mi < 0) {
          models.push_back(ModelT{r.model, {}, {}});
          mi = (int64_t)models.size() - 1;
        }
        ci = ri = -1;
      }
      ModelT& M = models[mi];
      if (ci < 0 || M.chains[ci].id != chain_id) {
        auto it = M.lookup.find(chain_id);
        if (it == M.lookup.end()) {
          M.lookup.emplace(chain_id, M.chains.size());
          M.chains.push_back(ChainT{chain_id, {}, {}});
          ci = (int64_t)M.chains.size() - 1;
        } else {
          ci = (int64_t)it->second;
        }
        ri = -1;
      }
      ChainT& Cn = M.chains[ci];
      if (ri < 0 || Cn.residues[ri].resi != r.resi || Cn.residues[ri].icode != ic_id) {
        const auto key = std::make_pair((int64_t)r.resi, ic_id);
        auto it = Cn.lookup.find(key);
        if (it == Cn.lookup.end()) {
          Cn.lookup.emplace(key, Cn.residues.size());
          Cn.residues.push_back(ResT{r.resi, ic_id, {}});
          ri = (int64_t)Cn.residues.size() - 1;
        } else {
          ri = (int64_t)it->second;
        }
      }
      ResT& R = Cn.residues[ri];
      ConfT* cf = nullptr;
      for (auto& c : R.confs)
        if (c.name == resn_id && c.altloc == alt_id) cf = &c;
      if (!cf) {
        R.confs.push_back(ConfT{resn_id, alt_id, {}});
        cf = &R.confs.back();
      }
      cf->recs.push_back(k);
      // atom-name id of the record: a direct-mapped cache on the 8 raw bytes in front of the string table
      uint64_t raw;
      std::memcpy(&raw, r.name, 8);
      NameSlot& ns = name_cache[(raw * 0x9E3779B97F4A7C15ull) >> 58];
      if (ns.id < 0 || ns.raw != raw) {
        ns.raw = raw;
        ns.id = P.intern(fixed_field(r.name, 8));
      }
      name_id[k] = ns.id;
    }
  }
  const int32_t blank = P.intern("");
  // ---- 2. blank-altloc conformer joins every other conformer (pdbtbx reader post-pass) ----------------
  for (auto& M : models)
    for (auto& Cn : M.chains)
      for (auto& R : Cn.residues) {
        if (R.confs.size() < 2) continue;
        int which = -1;
        for (size_t i = 0; i < R.confs.size(); ++i)
          if (R.confs[i].altloc == blank) which = (int)i;
        if (which < 0) continue;
        const std::vector<int64_t> shared = R.confs[which].recs;
        R.confs.erase(R.confs.begin() + which);
        for (auto& c : R.confs) c.recs.insert(c.recs.end(), shared.begin(), shared.end());
      }
  // ---- 3. load_model filter + zero occupancy ---------------------------------------------------------
  auto uniform_name = [&](const ResT& R) -> int32_t {  // pdbtbx Residue::name()
    if (R.confs.empty()) return -1;
    for (auto& c : R.confs)
      if (c.name != R.confs[0].name) return -1;
    return R.confs[0].name;
  };
  for (auto& M : models)
    for (auto& Cn : M.chains) {
      std::vector<ResT> kept;
      kept.reserve(Cn.residues.size());
      for (auto& R : Cn.residues) {
        const int32_t nm = uniform_name(R);
        if (nm < 0) throw Error{ARP_ERR_RESIDUE_NAME, "residue has conformers with different names (Residue::name() is None)"};
        if ((flags & ARPH_FLAG_SKIP_LOAD_FILTER) || known_residue(P.strings[nm])) kept.push_back(std::move(R));
      }
      Cn.residues.swap(kept);
      Cn.lookup.clear();
      if (ignore_zero_occupancy)
        for (auto& R : Cn.residues)
          for (auto& c : R.confs)
            c.recs.erase(std::remove_if(c.recs.begin(), c.recs.end(), [&](int64_t k) { return recs[k].occupancy == 0.0; }), c.recs.end());
    }
  // ---- 4. chain groups ------------------------------------------------------------------------------
  std::set<std::string> all_chains, ligand, receptor;
  for (auto& M : models)
    for (auto& Cn : M.chains) all_chains.insert(P.strings[Cn.id]);
  parse_groups(all_chains, groups, ligand, receptor);
  std::unordered_map<int32_t, uint32_t> chain_rank;  // lexicographic (byte order) rank, complex.rs:129
  {
    uint32_t rank = 0;
    for (auto& id : all_chains) {
      chain_rank[P.intern(id)] = rank++;
      P.chain_by_rank.push_back(id);
    }
    if (rank > (1u << 24)) throw Error{ARP_ERR_LIMIT, "more than 2^24 distinct chain ids"};
  }
  if (models.size() > 256) throw Error{ARP_ERR_LIMIT, "more than 256 models"};

  // ---- 5. planes (complex.rs:442-514), including the multi-model key overwrite (ledger item 12) ------
  const bool quirk = models.size() > 1 && !(flags & ARPH_FLAG_NO_MULTIMODEL_QUIRK);
  // table bits of every record by RESIDUE name (residue().name(): planes, hydrophobic.rs:16, aromatic.rs:18, vdw.rs:50),
  // looked up once per (residue name, atom name) pair and shared by the plane fits and the attribute words below
  std::unordered_map<uint64_t, uint32_t> bits_cache;
  auto bits_of = [&](int32_t resn_id, int32_t atom_id) {
    const uint64_t key = ((uint64_t)(uint32_t)resn_id << 32) | (uint32_t)atom_id;
    auto it = bits_cache.find(key);
    if (it != bits_cache.end()) return it->second;
    const uint32_t b = table_bits(P.strings[resn_id], P.strings[atom_id]);
    bits_cache.emplace(key, b);
    return b;
  };
  std::vector<uint32_t> bits_res((size_t)n, 0u);
  std::vector<const ResT*> all_res;  // every residue, models / chains / residues in order
  for (auto& M : models)
    for (auto& Cn : M.chains)
      for (auto& R : Cn.residues) {
        all_res.push_back(&R);
        const int32_t nm = uniform_name(R);
        int32_t last_atom = -1;
        uint32_t last_bits = 0;
        for (auto& c : R.confs)
          for (int64_t k : c.recs) {
            if (name_id[k] != last_atom) {
              last_atom = name_id[k];
              last_bits = bits_of(nm, last_atom);
            }
            bits_res[k] = last_bits;
          }
      }
  auto residue_plane = [&](const ResT& R, uint32_t want, Plane& out) {  // thread-safe: reads only
    std::vector<double> pts;
    for (auto& c : R.confs)
      for (int64_t k : c.recs)
        if (bits_res[k] & want) {
          pts.push_back(recs[k].x);
          pts.push_back(recs[k].y);
          pts.push_back(recs[k].z);
        }
    return fit_plane(pts.data(), (int64_t)pts.size() / 3, out);
  };
  // the plane fits (one small SVD per residue, two for ring residues) are independent: all host threads
  struct ResPlanes {
    Plane ring, sc;
    bool ok_ring, ok_sc, ring_res;
  };
  std::vector<ResPlanes> planes(all_res.size());
  {
    std::vector<char> ring_name_flag(P.strings.size(), 0);
    for (size_t i = 0; i < P.strings.size(); ++i) ring_name_flag[i] = is_ring_residue_name(P.strings[i]) ? 1 : 0;
    const int64_t nres_all = (int64_t)all_res.size();
#pragma omp parallel for schedule(static, 32) if (nres_all > 256)  // ~1 us per plane fit: worth the fork from a few hundred residues on
    for (int64_t i = 0; i < nres_all; ++i) {
      const ResT& R = *all_res[i];
      ResPlanes& o = planes[i];
      o.ring_res = ring_name_flag[uniform_name(R)] != 0;
      o.ok_ring = o.ring_res && residue_plane(R, T_RING, o.ring);
      o.ok_sc = residue_plane(R, T_SCP, o.sc);
    }
  }
  using PKey = std::array<int64_t, 5>;  // chain, resi, icode, altloc, resn
  std::map<PKey, Plane> last_ring, last_sc;
  if (quirk) {
    size_t ri_all = 0;
    for (auto& M : models)
      for (auto& Cn : M.chains)
        for (auto& R : Cn.residues) {
          const int32_t nm = uniform_name(R);
          const ResPlanes& pl = planes[ri_all++];
          for (auto& c : R.confs) {
            const PKey key{Cn.id, R.resi, R.icode, c.altloc, nm};
            if (pl.ok_ring) last_ring[key] = pl.ring;
            if (pl.ok_sc) last_sc[key] = pl.sc;
          }
        }
  }

  // ---- 6. flatten in atoms_with_hierarchy() order ------------------------------------------------------
  int64_t n_atoms = 0, n_res = 0;
  for (auto& M : models)
    for (auto& Cn : M.chains) {
      n_res += (int64_t)Cn.residues.size();
      for (auto& R : Cn.residues)
        for (auto& c : R.confs) n_atoms += (int64_t)c.recs.size();
    }
  P.x.reserve(n_atoms);
  P.y.reserve(n_atoms);
  P.z.reserve(n_atoms);
  P.attr.reserve(n_atoms);
  P.chain_key.reserve(n_atoms);
  P.res_ordinal.reserve(n_atoms);
  P.res_id.reserve(n_atoms);
  P.atom_rc.reserve(n_atoms);
  P.atom_name.reserve(n_atoms);
  P.atom_serial.reserve(n_atoms);
  P.res_h_begin.reserve(n_res + 1);
  P.res_cb.reserve(n_res);
  P.res_sg.reserve(n_res);

  const int32_t id_CB = P.intern("CB"), id_SG = P.intern("SG"), id_C = P.intern("C"), id_CYS = P.intern("CYS");
  char last_el[4] = {1};
  int last_el_class = -2;
  size_t res_all = 0;  // index into planes[]

  for (size_t mi = 0; mi < models.size(); ++mi) {
    const ModelT& M = models[mi];
    for (auto& Cn : M.chains) {
      const std::string chain_str = P.strings[Cn.id];  // by value: intern() below may reallocate P.strings
      const uint32_t group_bits = (ligand.count(chain_str) ? ARP_ATTR_LIGAND : 0u) | (receptor.count(chain_str) ? ARP_ATTR_RECEPTOR : 0u);
      const uint32_t ckey = ((uint32_t)mi << 24) | chain_rank[Cn.id];
      for (size_t ordinal = 0; ordinal < Cn.residues.size(); ++ordinal) {
        const ResT& R = Cn.residues[ordinal];
        const int32_t resn_id = uniform_name(R);
        const std::string resn = P.strings[resn_id];
        const uint32_t rid = (uint32_t)P.res_cb.size();
        P.res_cb.push_back(-1);
        P.res_sg.push_back(-1);
        P.res_h_begin.push_back((uint32_t)P.hx.size());
        const ResPlanes& pl = planes[res_all++];
        Plane pr{}, ps{};
        bool okr = false, oks = false;
        const bool ring_res = pl.ring_res;
        if (!quirk) {
          okr = pl.ok_ring, oks = pl.ok_sc;
          pr = pl.ring, ps = pl.sc;
        }
        for (auto& c : R.confs) {
          if (quirk) {
            const PKey key{Cn.id, R.resi, R.icode, c.altloc, resn_id};
            auto ir = last_ring.find(key);
            okr = ir != last_ring.end();
            if (okr) pr = ir->second;
            auto is = last_sc.find(key);
            oks = is != last_sc.end();
            if (oks) ps = is->second;
          }
          const uint32_t rc = (uint32_t)P.resconfs.size();
          Packed::ResConf rcx;
          rcx.chain = Cn.id;
          rcx.resn = resn_id;
          rcx.icode = R.icode;
          rcx.altloc = c.altloc;
          rcx.resi = R.resi;
          rcx.model_serial = M.serial;
          rcx.has_sc = oks;
          rcx.sc = ps;
          P.resconfs.push_back(rcx);
          if (okr) {
            for (int k = 0; k < 3; ++k) {
              P.ring_c.push_back(pr.c[k]);
              P.ring_n.push_back(pr.n[k]);
            }
            P.ring_attr.push_back(group_bits);
            P.ring_chain_key.push_back(ckey);
            P.ring_res_ordinal.push_back((uint32_t)ordinal);
            P.ring_rc.push_back(rc);
          } else if (ring_res && (!quirk || mi == 0)) {
            P.warnings.push_back("Failed to calculate ring position for chain " + chain_str + " residue " + resn + " " +
                                 std::to_string(R.resi) + P.strings[R.icode]);
          }
          for (int64_t k : c.recs) {
            const arp_atom_record& a = recs[k];
            const int32_t last_name_id = name_id[k];
            if (std::memcmp(last_el, a.element, 4) != 0 || last_el_class == -2) {
              std::memcpy(last_el, a.element, 4);
              last_el_class = arp_element_class(fixed_field(a.element, 4).c_str());
            }
            if (last_el_class < 0) {
              const std::string el = fixed_field(a.element, 4);
              throw Error{ARP_ERR_ELEMENT, el.empty() ? "atom " + std::to_string(a.serial) + " has no element"
                                                      : "unsupported element '" + el + "' (not one of the element classes this path knows; pdbtbx would need a van der Waals radius for it, vdw.rs:28)"};
            }
            const uint32_t by_res = bits_res[k];                      // residue().name(): hydrophobic.rs:16, aromatic.rs:18, vdw.rs:50
            const uint32_t by_conf = c.name == resn_id ? by_res : bits_of(c.name, last_name_id);  // conformer().name(): ionic.rs:45-46, hbond.rs:120-121
            uint32_t w = (uint32_t)last_el_class | group_bits;
            if (by_conf & T_POS) w |= ARP_ATTR_POS_CONF;
            if (by_conf & T_NEG) w |= ARP_ATTR_NEG_CONF;
            if (by_conf & T_DON) w |= ARP_ATTR_DONOR;
            if (by_conf & T_ACC) w |= ARP_ATTR_ACCEPTOR;
            if (by_res & T_POS) w |= ARP_ATTR_POS_RES;
            if (by_res & T_HYD) w |= ARP_ATTR_HYDROPHOBIC;
            if (last_el_class == 1 && last_name_id != id_C) w |= ARP_ATTR_WEAK_DONOR;  // hbond.rs:204-207 (class 1 = carbon)
            if (resn_id == id_CYS && last_name_id == id_SG) w |= ARP_ATTR_CYS_SG;
            const int32_t idx = (int32_t)P.x.size();
            if (last_el_class == 0) {  // hydrogen: joins the residue's H list (hbond.rs:38-42)
              P.hx.push_back(a.x);
              P.hy.push_back(a.y);
              P.hz.push_back(a.z);
            }
            if (last_name_id == id_CB && P.res_cb[rid] < 0) P.res_cb[rid] = idx;  // vdw.rs:55-74 `.find()` = first match
            if (last_name_id == id_SG && P.res_sg[rid] < 0) P.res_sg[rid] = idx;
            P.x.push_back(a.x);
            P.y.push_back(a.y);
            P.z.push_back(a.z);
            P.attr.push_back(w);
            P.chain_key.push_back(ckey);
            P.res_ordinal.push_back((uint32_t)ordinal);
            P.res_id.push_back(rid);
            P.atom_rc.push_back(rc);
            P.atom_name.push_back(last_name_id);
            P.atom_serial.push_back(a.serial);
          }
        }
      }
    }
  }
  P.res_h_begin.push_back((uint32_t)P.hx.size());
  if (P.ring_rc.empty()) throw Error{ARP_ERR_NO_RINGS, "Error building ring positions"};  // complex.rs:50
  P.ring_name = P.intern("Ring");
  P.finish();
}

void Packed::finish() {
  arp_structure& t = s;
  t.n_atoms = (int64_t)x.size();
  t.n_residues = (int64_t)res_cb.size();
  t.n_rings = (int64_t)ring_rc.size();
  t.n_h = (int64_t)hx.size();
  t.x = x.data();
  t.y = y.data();
  t.z = z.data();
  t.attr = attr.data();
  t.chain_key = chain_key.data();
  t.res_ordinal = res_ordinal.data();
  t.res_id = res_id.data();
  t.res_h_begin = res_h_begin.data();
  t.h_x = hx.data();
  t.h_y = hy.data();
  t.h_z = hz.data();
  t.res_cb = res_cb.data();
  t.res_sg = res_sg.data();
  t.ring_center = ring_c.data();
  t.ring_normal = ring_n.data();
  t.ring_attr = ring_attr.data();
  t.ring_chain_key = ring_chain_key.data();
  t.ring_res_ordinal = ring_res_ordinal.data();
  // compact per-residue form of chain_key / res_ordinal (what the device path uploads when it is present)
  res_chain_key.assign(res_cb.size(), 0u);
  res_res_ordinal.assign(res_cb.size(), 0u);
  for (size_t a = 0; a < res_id.size(); ++a) {
    res_chain_key[res_id[a]] = chain_key[a];
    res_res_ordinal[res_id[a]] = res_ordinal[a];
  }
  t.res_chain_key = res_chain_key.data();
  t.res_res_ordinal = res_res_ordinal.data();
  // residue tags (arp_structure.res_tag): the residues of a chain are contiguous and in ordinal order here, so a running
  // count that steps by 1 inside a chain (consecutive ordinals) and by 2 otherwise has the required property
  res_tag.assign(res_cb.size(), 0u);
  for (size_t r = 1; r < res_cb.size(); ++r)
    res_tag[r] = res_tag[r - 1] + ((res_chain_key[r] == res_chain_key[r - 1] && res_res_ordinal[r] == res_res_ordinal[r - 1] + 1u) ? 1u : 2u);
  t.res_tag = res_tag.data();
  t.atom_gid = atom_gid.empty() ? nullptr : atom_gid.data();
  t.ring_gid = ring_gid.empty() ? nullptr : ring_gid.data();
  // atoms_with_hierarchy() order keeps the atoms of a residue together: res_id as a CSR (left out if it ever does not)
  t.res_atom_begin = nullptr;
  {
    bool grouped = true;
    for (size_t a = 1; a < res_id.size() && grouped; ++a) grouped = res_id[a] == res_id[a - 1] || res_id[a] == res_id[a - 1] + 1;
    if (grouped && (res_id.empty() || (res_id.front() == 0 && (size_t)res_id.back() + 1 == res_cb.size()))) {
      res_atom_begin.assign(res_cb.size() + 1, 0u);
      for (uint32_t r : res_id) res_atom_begin[r + 1] += 1;
      for (size_t r = 0; r < res_cb.size(); ++r) res_atom_begin[r + 1] += res_atom_begin[r];
      t.res_atom_begin = res_atom_begin.data();
    }
  }
  // on-lattice form of the coordinates (arp_structure.x_milli): offered only when EVERY coordinate is bit for bit the
  // double (double)v / 1000.0 of a 32-bit v -- true for anything read from PDB fixed columns -- so the device rebuilds the
  // same f64 values from half the bytes
  t.x_milli = t.y_milli = t.z_milli = nullptr;
  {
    bool lattice = !x.empty();
    for (int a = 0; a < 3 && lattice; ++a) {
      const std::vector<double>& v = a == 0 ? x : (a == 1 ? y : z);
      std::vector<int32_t>& m = a == 0 ? x_milli : (a == 1 ? y_milli : z_milli);
      m.resize(v.size());
      for (size_t k = 0; k < v.size(); ++k) {
        const double q = v[k] * 1000.0;
        if (!(q > -2147483000.0 && q < 2147483000.0)) { lattice = false; break; }
        const int32_t r = (int32_t)std::llrint(q);
        if ((double)r / 1000.0 != v[k] || (r == 0 && std::signbit(v[k]))) { lattice = false; break; }
        m[k] = r;
      }
    }
    if (lattice) {
      t.x_milli = x_milli.data();
      t.y_milli = y_milli.data();
      t.z_milli = z_milli.data();
    } else {
      x_milli.clear(), y_milli.clear(), z_milli.clear();
    }
  }
  // bounding-box hint for the cell-grid layout (verified on the device)
  t.has_bbox = 0;
  if (!x.empty()) {
    bool finite = true;
    for (int a = 0; a < 3; ++a) {
      const std::vector<double>& v = a == 0 ? x : (a == 1 ? y : z);
      double lo = v[0], hi = v[0];
      for (double q : v) {
        if (!(q == q) || q - q != 0.0) finite = false;
        lo = q < lo ? q : lo;
        hi = q > hi ? q : hi;
      }
      t.bbox_min[a] = lo;
      t.bbox_max[a] = hi;
    }
    t.has_bbox = finite ? 1 : 0;
  }
}

}  // namespace arp

struct arph_packed {
  arp::Packed p;
  arp::FrameTables frame_tables;  // built on first use (arph_packed_frame_tables)
  bool have_frame_tables = false;
};

extern "C" {

int arph_pack(const arp_atom_record* atoms, int64_t n, const char* groups, int ignore_zero_occupancy, uint32_t flags,
              arph_packed** out) {
  *out = nullptr;
  arp::NvtxRange nvtx_("arph:pack");
  auto* h = new arph_packed();
  try {
    arp::pack_records(atoms, n, groups ? groups : "/", ignore_zero_occupancy != 0, flags, h->p);
  } catch (const arp::Error& e) {
    arp::last_error() = e.msg;
    delete h;
    return e.code;
  } catch (const std::exception& e) {
    arp::last_error() = e.what();
    delete h;
    return ARP_ERR_OOM;
  }
  *out = h;
  return ARP_OK;
}
/* Many independent structures at once (BASELINE config 4: 10 000 complexes): one structure per host thread, each packed
 * exactly like arph_pack does it (the parallel regions inside a pack run serially here: no nested teams). */
int arph_pack_many(const arp_atom_record* const* atoms, const int64_t* n, int64_t count, const char* groups, int ignore_zero_occupancy,
                   uint32_t flags, arph_packed** out) {
  arp::NvtxRange nvtx_("arph:pack_many");
  if (count < 0 || (count > 0 && (!atoms || !n || !out))) {
    arp::last_error() = "arph_pack_many: null argument";
    return ARP_ERR_ARG;
  }
  for (int64_t k = 0; k < count; ++k) out[k] = nullptr;
  int first_rc = ARP_OK;
  int64_t first_bad = -1;
  std::string first_msg;
#pragma omp parallel for schedule(dynamic, 4)
  for (int64_t k = 0; k < count; ++k) {
    arph_packed* h = nullptr;
    const int rc = arph_pack(atoms[k], n[k], groups, ignore_zero_occupancy, flags, &h);
    out[k] = h;
    if (rc != ARP_OK) {
#pragma omp critical(arph_pack_many_error)
      if (first_bad < 0 || k < first_bad) first_bad = k, first_rc = rc, first_msg = arp::last_error();  // this thread's message
    }
  }
  if (first_rc != ARP_OK) {
    for (int64_t k = 0; k < count; ++k) {
      delete out[k];
      out[k] = nullptr;
    }
    arp::last_error() = "structure " + std::to_string(first_bad) + ": " + first_msg;
  }
  return first_rc;
}
const arp_structure* arph_packed_structure(const arph_packed* p) { return &p->p.s; }
int64_t arph_packed_n_warnings(const arph_packed* p) { return (int64_t)p->p.warnings.size(); }
const char* arph_packed_warning(const arph_packed* p, int64_t i) { return p->p.warnings[(size_t)i].c_str(); }
void arph_packed_free(arph_packed* p) { delete p; }

/* ---- a packed structure as one part of a larger one (slab jobs, arp_slab_exchange) -------------------------------------- */
int64_t arph_packed_n_chains(const arph_packed* p) { return p ? (int64_t)p->p.chain_by_rank.size() : 0; }
const char* arph_packed_chain_id(const arph_packed* p, int64_t rank) {
  if (!p || rank < 0 || rank >= (int64_t)p->p.chain_by_rank.size()) return nullptr;
  return p->p.chain_by_rank[(size_t)rank].c_str();
}
int arph_packed_remap_chain_ranks(arph_packed* p, const uint32_t* map, int64_t n) {
  if (!p || !map || n != (int64_t)p->p.chain_by_rank.size()) {
    arp::last_error() = "arph_packed_remap_chain_ranks: one new rank per chain of the packed structure";
    return ARP_ERR_ARG;
  }
  for (int64_t k = 0; k < n; ++k)
    if (map[k] >= (1u << 24)) {
      arp::last_error() = "chain rank beyond 2^24";
      return ARP_ERR_LIMIT;
    }
  auto fix = [&](std::vector<uint32_t>& v) {
    for (auto& k : v) k = (k & 0xff000000u) | map[k & 0x00ffffffu];
  };
  fix(p->p.chain_key);
  fix(p->p.res_chain_key);
  fix(p->p.ring_chain_key);
  return ARP_OK;
}
int arph_packed_set_global_ids(arph_packed* p, const uint32_t* atom_gid, const uint32_t* ring_gid, const uint32_t* res_tag) {
  if (!p || !atom_gid || !ring_gid) {
    arp::last_error() = "arph_packed_set_global_ids: null argument";
    return ARP_ERR_ARG;
  }
  arp::Packed& P = p->p;
  for (int64_t a = 0; a < P.s.n_atoms; ++a)
    if (atom_gid[a] & ARP_RING_FLAG) {
      arp::last_error() = "atom id beyond 2^31";
      return ARP_ERR_LIMIT;
    }
  P.atom_gid.assign(atom_gid, atom_gid + P.s.n_atoms);
  P.ring_gid.assign(ring_gid, ring_gid + P.s.n_rings);
  if (res_tag) P.res_tag.assign(res_tag, res_tag + P.s.n_residues);
  P.s.atom_gid = P.atom_gid.data();
  P.s.ring_gid = P.ring_gid.data();
  P.s.res_tag = P.res_tag.data();
  return ARP_OK;
}

int arph_packed_frame_tables(arph_packed* p, const arp_frame_tables** out) {
  *out = nullptr;
  if (!p) {
    arp::last_error() = "null packed structure";
    return ARP_ERR_ARG;
  }
  try {
    if (!p->have_frame_tables) {
      arp::build_frame_tables(p->p, p->frame_tables);
      p->have_frame_tables = true;
    }
  } catch (const arp::Error& e) {
    arp::last_error() = e.msg;
    return e.code;
  } catch (const std::exception& e) {
    arp::last_error() = e.what();
    return ARP_ERR_OOM;
  }
  *out = &p->frame_tables.t;
  return ARP_OK;
}

int arph_parse_groups(const char* chains_csv, const char* groups, char* ligand_out, char* receptor_out, int64_t cap) {
  try {
    std::set<std::string> all, lig, rec;
    const std::string s(chains_csv);
    for (size_t p = 0;;) {
      const size_t q = s.find(',', p);
      const std::string tok = s.substr(p, q == std::string::npos ? std::string::npos : q - p);
      if (!tok.empty()) all.insert(tok);
      if (q == std::string::npos) break;
      p = q + 1;
    }
    arp::parse_groups(all, groups, lig, rec);
    auto join = [](const std::set<std::string>& v) {
      std::string o;
      for (auto& e : v) o += (o.empty() ? "" : ",") + e;
      return o;
    };
    std::snprintf(ligand_out, (size_t)cap, "%s", join(lig).c_str());
    std::snprintf(receptor_out, (size_t)cap, "%s", join(rec).c_str());
    return ARP_OK;
  } catch (const arp::Error& e) {
    arp::last_error() = e.msg;
    return e.code;
  }
}
}

const arp::Packed& arph_packed_ref(const arph_packed* p) { return p->p; }
const arp_frame_tables* arph_packed_cached_frame_tables(const arph_packed* p) { return p->have_frame_tables ? &p->frame_tables.t : nullptr; }
