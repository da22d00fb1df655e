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
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <omp.h>

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
  int owner = 0;  // the block (= thread) of phase 1 that created the chain
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

namespace {

// A string table of one thread's block of records (phase 1 below); ids are in first-appearance order.
struct LocalStrings {
  std::vector<std::string> strings;
  std::unordered_map<std::string, int32_t> ids;
  int32_t intern(const std::string& s) {
    auto it = ids.find(s);
    if (it != ids.end()) return it->second;
    const int32_t id = (int32_t)strings.size();
    strings.push_back(s);
    ids.emplace(s, id);
    return id;
  }
};

// Records [lo, hi) -> models / chains / residues / conformers in first-appearance order, records in file order
// (pdbtbx: chain by id within a model, residue by (resi, icode) within a chain, conformer by (name, altloc) within a
// residue).  Strings go through `S.intern`; name_id[k] = id of record k's atom name.
template <class Strings>
void group_records(const arp_atom_record* recs, int64_t lo, int64_t hi, Strings& S, std::vector<ModelT>& models, int32_t* name_id) {
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
  for (int64_t k = lo; k < hi; ++k) {
    const arp_atom_record& r = recs[k];
    if (std::memcmp(last_chain, r.chain, 8) != 0 || chain_id < 0) {
      std::memcpy(last_chain, r.chain, 8);
      chain_id = S.intern(fixed_field(r.chain, 8));
    }
    if (std::memcmp(last_resn, r.resn, 8) != 0 || resn_id < 0) {
      std::memcpy(last_resn, r.resn, 8);
      resn_id = S.intern(fixed_field(r.resn, 8));
    }
    if (std::memcmp(last_alt, r.altloc, 2) != 0 || alt_id < 0) {
      std::memcpy(last_alt, r.altloc, 2);
      alt_id = S.intern(fixed_field(r.altloc, 2));
    }
    if (std::memcmp(last_ic, r.icode, 2) != 0 || ic_id < 0) {
      std::memcpy(last_ic, r.icode, 2);
      ic_id = S.intern(fixed_field(r.icode, 2));
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
        M.chains.push_back(ChainT{chain_id, {}, {}, 0});
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
      ns.id = S.intern(fixed_field(r.name, 8));
    }
    name_id[k] = ns.id;
  }
}

// Phase 1 on every host thread: the records are cut into blocks (at places where the residue changes), every block is
// grouped on its own with a string table of its own, then the blocks are joined in file order.  A string's global id is
// assigned in (block, local id) order = the order of first appearance in the file, and chains / residues / conformers
// are appended in block order, so the result is the one the single pass over all records gives.
void group_records_parallel(const arp_atom_record* recs, int64_t n, Packed& P, std::vector<ModelT>& models, int32_t* name_id, int blocks) {
  std::vector<int64_t> cut((size_t)blocks + 1, n);
  cut[0] = 0;
  for (int b = 1; b < blocks; ++b) {
    int64_t k = std::max(cut[b - 1], n * b / blocks);
    const auto same_residue = [&](int64_t i) {
      const arp_atom_record &u = recs[i - 1], &v = recs[i];
      return u.model == v.model && u.resi == v.resi && std::memcmp(u.chain, v.chain, 8) == 0 && std::memcmp(u.icode, v.icode, 2) == 0;
    };
    while (k > 0 && k < n && same_residue(k)) ++k;
    cut[b] = k;
  }
  std::vector<LocalStrings> ls((size_t)blocks);
  std::vector<std::vector<ModelT>> lm((size_t)blocks);
#pragma omp parallel for schedule(static, 1) num_threads(blocks)
  for (int b = 0; b < blocks; ++b) group_records(recs, cut[b], cut[b + 1], ls[b], lm[b], name_id);
  // global string ids, in order of first appearance
  std::vector<std::vector<int32_t>> to_global((size_t)blocks);
  for (int b = 0; b < blocks; ++b) {
    to_global[b].resize(ls[b].strings.size());
    for (size_t i = 0; i < ls[b].strings.size(); ++i) to_global[b][i] = P.intern(ls[b].strings[i]);
  }
#pragma omp parallel for schedule(static, 1) num_threads(blocks)
  for (int b = 0; b < blocks; ++b) {
    const std::vector<int32_t>& g = to_global[b];
    for (int64_t k = cut[b]; k < cut[b + 1]; ++k) name_id[k] = g[name_id[k]];
    for (auto& M : lm[b]) {
      M.lookup.clear();
      for (auto& Cn : M.chains) {
        Cn.id = g[Cn.id];
        Cn.owner = b;
        Cn.lookup.clear();  // keyed by local ids; rebuilt below for the (rare) chain that continues in a later block
        for (auto& R : Cn.residues) {
          R.icode = g[R.icode];
          for (auto& c : R.confs) c.name = g[c.name], c.altloc = g[c.altloc];
        }
      }
    }
  }
  for (int b = 0; b < blocks; ++b)
    for (auto& LM : lm[b]) {
      int64_t mi = -1;
      for (size_t m = 0; m < models.size(); ++m)
        if (models[m].serial == LM.serial) mi = (int64_t)m;
      if (mi < 0) {
        models.push_back(ModelT{LM.serial, {}, {}});
        mi = (int64_t)models.size() - 1;
      }
      ModelT& M = models[mi];
      for (auto& LC : LM.chains) {
        auto it = M.lookup.find(LC.id);
        if (it == M.lookup.end()) {
          M.lookup.emplace(LC.id, M.chains.size());
          M.chains.push_back(std::move(LC));
          continue;
        }
        ChainT& G = M.chains[it->second];  // the chain goes on: residue by residue, as the single pass would
        if (G.lookup.size() != G.residues.size()) {
          G.lookup.clear();
          for (size_t r = 0; r < G.residues.size(); ++r) G.lookup.emplace(std::make_pair(G.residues[r].resi, G.residues[r].icode), r);
        }
        for (auto& LR : LC.residues) {
          const auto key = std::make_pair(LR.resi, LR.icode);
          auto ir = G.lookup.find(key);
          if (ir == G.lookup.end()) {
            G.lookup.emplace(key, G.residues.size());
            G.residues.push_back(std::move(LR));
            continue;
          }
          ResT& GR = G.residues[ir->second];
          for (auto& lc : LR.confs) {
            ConfT* cf = nullptr;
            for (auto& c : GR.confs)
              if (c.name == lc.name && c.altloc == lc.altloc) cf = &c;
            if (!cf)
              GR.confs.push_back(std::move(lc));
            else
              cf->recs.insert(cf->recs.end(), lc.recs.begin(), lc.recs.end());
          }
        }
      }
    }
}

}  // namespace

void pack_records(const arp_atom_record* recs, int64_t n, const std::string& groups, bool ignore_zero_occupancy, uint32_t flags,
                  Packed& P) {
  if (n >= (int64_t)1 << 31) throw Error{ARP_ERR_LIMIT, "more than 2^31 atom records"};
  PhaseClock clk;
  // every phase below that walks the records or the residues runs on all host threads (one thread when the caller is
  // itself a parallel region: arph_pack_many packs one structure per thread)
  const int host_threads = omp_in_parallel() ? 1 : omp_get_max_threads();
  // ---- 1. group records into the hierarchy --------------------------------------------------------
  std::vector<ModelT> models;
  std::vector<int32_t> name_id((size_t)n);
  {
    int blocks = (int)std::min<int64_t>(host_threads, n / 8192);
    if (const char* ev = std::getenv("ARPH_PACK_BLOCKS")) blocks = (int)std::min<int64_t>(std::max(1, std::atoi(ev)), std::max<int64_t>(n, 1));  // tests: any input, any cut
    if (blocks > 1)
      group_records_parallel(recs, n, P, models, name_id.data(), blocks);
    else
      group_records(recs, 0, n, P, models, name_id.data());
  }
  clk.mark("1 hierarchy");
  const int32_t blank = P.intern("");
  // every chain of every model, in hierarchy order
  struct ChainRef {
    size_t mi;
    ChainT* c;
  };
  std::vector<ChainRef> chains;
  for (size_t mi = 0; mi < models.size(); ++mi)
    for (auto& Cn : models[mi].chains) chains.push_back(ChainRef{mi, &Cn});
  const int64_t n_chains_all = (int64_t)chains.size();
  auto uniform_name = [](const ResT& R) -> int32_t {  // pdbtbx Residue::name()
    if (R.confs.empty()) return -1;
    for (auto& c : R.confs)
      if (c.name != R.confs[0].name) return -1;
    return R.confs[0].name;
  };
  // ---- 2. blank-altloc conformer joins every other conformer (pdbtbx reader post-pass) ----------------
  // ---- 3. load_model filter + zero occupancy ---------------------------------------------------------
  {
    std::vector<char> known(P.strings.size(), 0);
    for (size_t i = 0; i < P.strings.size(); ++i) known[i] = ((flags & ARPH_FLAG_SKIP_LOAD_FILTER) || known_residue(P.strings[i])) ? 1 : 0;
    std::vector<char> bad_name((size_t)n_chains_all, 0);
#pragma omp parallel for schedule(dynamic, 4) num_threads(host_threads) if (n > 65536)
    for (int64_t ci = 0; ci < n_chains_all; ++ci) {
      ChainT& Cn = *chains[ci].c;
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
      std::vector<ResT> kept;
      kept.reserve(Cn.residues.size());
      for (auto& R : Cn.residues) {
        const int32_t nm = uniform_name(R);
        if (nm < 0) {
          bad_name[ci] = 1;
          break;
        }
        if (known[nm]) kept.push_back(std::move(R));
      }
      if (bad_name[ci]) continue;
      Cn.residues.swap(kept);
      Cn.lookup.clear();
      if (ignore_zero_occupancy)
        for (auto& R : Cn.residues)
          for (auto& c : R.confs)
            c.recs.erase(std::remove_if(c.recs.begin(), c.recs.end(), [&](int64_t k) { return recs[k].occupancy == 0.0; }), c.recs.end());
    }
    for (int64_t ci = 0; ci < n_chains_all; ++ci)
      if (bad_name[ci]) throw Error{ARP_ERR_RESIDUE_NAME, "residue has conformers with different names (Residue::name() is None)"};
  }
  clk.mark("2-3 conformers, filters");
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
  // every residue, models / chains / residues in order
  struct ResRef {
    const ResT* r;
    uint32_t chain;    // index into chains[]
    uint32_t ordinal;  // position within its chain
  };
  std::vector<ResRef> all_res;
  for (int64_t ci = 0; ci < n_chains_all; ++ci) {
    const ChainT& Cn = *chains[ci].c;
    for (size_t o = 0; o < Cn.residues.size(); ++o) all_res.push_back(ResRef{&Cn.residues[o], (uint32_t)ci, (uint32_t)o});
  }
  const int64_t n_res = (int64_t)all_res.size();
  // table bits of every record by RESIDUE name (residue().name(): planes, hydrophobic.rs:16, aromatic.rs:18, vdw.rs:50)
  // -- the conformers of a residue all carry that name (phase 3), so conformer().name() (ionic.rs:45-46, hbond.rs:120-121)
  // reads the same bits -- and the element class of every record; then the plane fits (one small SVD per residue, two
  // for ring residues).  All per residue, independent: every host thread, each with its own (residue name, atom name) cache.
  std::vector<uint32_t> bits_res((size_t)n, 0u);
  std::vector<int8_t> el_class((size_t)n, 0);
  struct ResPlanes {
    Plane ring, sc;
    bool ok_ring, ok_sc, ring_res;
  };
  std::vector<ResPlanes> planes((size_t)n_res);
  std::vector<uint32_t> res_atoms((size_t)n_res + 1, 0u), res_h((size_t)n_res + 1, 0u);  // counts, then offsets
  auto residue_plane = [&](const ResT& R, uint32_t want, Plane& out) {  // thread-safe: reads only
    double small[3 * 32] = {};  // a residue's ring / side-chain atoms: a dozen at most (the vector is for the odd giant)
    std::vector<double> big;
    int64_t np = 0;
    for (auto& c : R.confs)
      for (int64_t k : c.recs)
        if (bits_res[k] & want) {
          if (np < 32 && big.empty()) {
            small[3 * np] = recs[k].x, small[3 * np + 1] = recs[k].y, small[3 * np + 2] = recs[k].z;
          } else {
            if (big.empty()) big.assign(small, small + 3 * np);
            big.push_back(recs[k].x);
            big.push_back(recs[k].y);
            big.push_back(recs[k].z);
          }
          ++np;
        }
    return fit_plane(big.empty() ? small : big.data(), np, out);
  };
  {
    std::vector<char> ring_name_flag(P.strings.size(), 0);
    for (size_t i = 0; i < P.strings.size(); ++i) ring_name_flag[i] = is_ring_residue_name(P.strings[i]) ? 1 : 0;
#pragma omp parallel num_threads(host_threads) if (n_res > 256)  // ~1 us per plane fit: worth the fork from a few hundred residues on
    {
      std::unordered_map<uint64_t, uint32_t> bits_cache;
      char last_el[4] = {1};
      int last_el_class = -2;
#pragma omp for schedule(static, 64)
      for (int64_t i = 0; i < n_res; ++i) {
        const ResT& R = *all_res[i].r;
        const int32_t nm = uniform_name(R);
        int32_t last_atom = -1;
        uint32_t last_bits = 0, na = 0, nhyd = 0;
        for (auto& c : R.confs)
          for (int64_t k : c.recs) {
            if (name_id[k] != last_atom) {
              last_atom = name_id[k];
              const uint64_t key = ((uint64_t)(uint32_t)nm << 32) | (uint32_t)last_atom;
              auto it = bits_cache.find(key);
              if (it == bits_cache.end()) it = bits_cache.emplace(key, table_bits(P.strings[nm], P.strings[last_atom])).first;
              last_bits = it->second;
            }
            bits_res[k] = last_bits;
            const arp_atom_record& a = recs[k];
            if (std::memcmp(last_el, a.element, 4) != 0 || last_el_class == -2) {
              std::memcpy(last_el, a.element, 4);
              last_el_class = arp_element_class(fixed_field(a.element, 4).c_str());
            }
            el_class[k] = (int8_t)last_el_class;  // < 0: unknown element, reported in phase 6 for the first such atom
            na += 1;
            nhyd += last_el_class == 0 ? 1u : 0u;
          }
        res_atoms[i] = na;
        res_h[i] = nhyd;
        ResPlanes& o = planes[i];
        o.ring_res = ring_name_flag[nm] != 0;
        o.ok_ring = o.ring_res && residue_plane(R, T_RING, o.ring);
        o.ok_sc = residue_plane(R, T_SCP, o.sc);
      }
    }
  }
  clk.mark("4-5 groups, bits, plane fits");
  using PKey = std::array<int64_t, 5>;  // chain, resi, icode, altloc, resn
  std::map<PKey, Plane> last_ring, last_sc;
  if (quirk) {
    for (int64_t i = 0; i < n_res; ++i) {
      const ResT& R = *all_res[i].r;
      const ChainT& Cn = *chains[all_res[i].chain].c;
      const int32_t nm = uniform_name(R);
      const ResPlanes& pl = planes[i];
      for (auto& c : R.confs) {
        const PKey key{Cn.id, R.resi, R.icode, c.altloc, nm};
        if (pl.ok_ring) last_ring[key] = pl.ring;
        if (pl.ok_sc) last_sc[key] = pl.sc;
      }
    }
  }

  // ---- 6. flatten in atoms_with_hierarchy() order ------------------------------------------------------
  // Where every residue's atoms, hydrogens, conformers and rings start in the flat arrays: counts per residue (the atom
  // and hydrogen counts come from phase 5), one prefix sum, then every residue writes its own ranges -- on all host threads.
  std::vector<uint32_t> res_rc((size_t)n_res + 1, 0u), res_ring((size_t)n_res + 1, 0u);
  std::vector<uint8_t> conf_ring;  // per conformer (filled below): bit 0 = has a ring plane, bit 1 = has a side-chain plane
  {
    for (int64_t i = 0; i < n_res; ++i) res_rc[i] = (uint32_t)all_res[i].r->confs.size();
    uint32_t at = 0;
    for (int64_t i = 0; i <= n_res; ++i) {
      const uint32_t c = i < n_res ? res_rc[i] : 0u;
      res_rc[i] = at;
      at += c;
    }
    conf_ring.assign(at, 0);
  }
  const int64_t n_rc = res_rc[n_res];
  std::vector<const Plane*> conf_pr((size_t)n_rc, nullptr), conf_ps((size_t)n_rc, nullptr);
#pragma omp parallel for schedule(static, 256) num_threads(host_threads) if (n_res > 4096)
  for (int64_t i = 0; i < n_res; ++i) {
    const ResT& R = *all_res[i].r;
    const ResPlanes& pl = planes[i];
    uint32_t rings = 0;
    for (size_t ck = 0; ck < R.confs.size(); ++ck) {
      const size_t rc = res_rc[i] + ck;
      const Plane *pr = nullptr, *ps = nullptr;
      if (!quirk) {
        pr = pl.ok_ring ? &pl.ring : nullptr;
        ps = pl.ok_sc ? &pl.sc : nullptr;
      } else {  // the maps hold what the LAST model left under this key (complex.rs:447-449)
        const PKey key{chains[all_res[i].chain].c->id, R.resi, R.icode, R.confs[ck].altloc, uniform_name(R)};
        auto ir = last_ring.find(key);
        if (ir != last_ring.end()) pr = &ir->second;
        auto is = last_sc.find(key);
        if (is != last_sc.end()) ps = &is->second;
      }
      conf_pr[rc] = pr, conf_ps[rc] = ps;
      conf_ring[rc] = (uint8_t)((pr ? 1 : 0) | (ps ? 2 : 0));
      rings += pr ? 1u : 0u;
    }
    res_ring[i] = rings;
  }
  {
    uint32_t a = 0, h = 0, g = 0;
    for (int64_t i = 0; i <= n_res; ++i) {
      const uint32_t ca = i < n_res ? res_atoms[i] : 0u, ch = i < n_res ? res_h[i] : 0u, cg = i < n_res ? res_ring[i] : 0u;
      res_atoms[i] = a, res_h[i] = h, res_ring[i] = g;
      a += ca, h += ch, g += cg;
    }
  }
  const int64_t n_atoms = res_atoms[n_res], n_h = res_h[n_res], n_rings = res_ring[n_res];
  P.x.resize(n_atoms), P.y.resize(n_atoms), P.z.resize(n_atoms);
  P.attr.resize(n_atoms), P.chain_key.resize(n_atoms), P.res_ordinal.resize(n_atoms), P.res_id.resize(n_atoms);
  P.atom_rc.resize(n_atoms), P.atom_name.resize(n_atoms), P.atom_serial.resize(n_atoms);
  P.hx.resize(n_h), P.hy.resize(n_h), P.hz.resize(n_h);
  P.res_h_begin.assign(res_h.begin(), res_h.end());  // n_res + 1 entries, the last one closes the CSR
  P.res_cb.assign((size_t)n_res, -1), P.res_sg.assign((size_t)n_res, -1);
  P.resconfs.resize((size_t)n_rc);
  P.ring_c.resize(3 * n_rings), P.ring_n.resize(3 * n_rings);
  P.ring_attr.resize(n_rings), P.ring_chain_key.resize(n_rings), P.ring_res_ordinal.resize(n_rings), P.ring_rc.resize(n_rings);

  const int32_t id_CB = P.intern("CB"), id_SG = P.intern("SG"), id_C = P.intern("C"), id_CYS = P.intern("CYS");
  // per chain: group bits and key
  std::vector<uint32_t> chain_bits((size_t)n_chains_all), chain_ckey((size_t)n_chains_all);
  for (int64_t ci = 0; ci < n_chains_all; ++ci) {
    const std::string& chain_str = P.strings[chains[ci].c->id];
    chain_bits[ci] = (ligand.count(chain_str) ? ARP_ATTR_LIGAND : 0u) | (receptor.count(chain_str) ? ARP_ATTR_RECEPTOR : 0u);
    chain_ckey[ci] = ((uint32_t)chains[ci].mi << 24) | chain_rank[chains[ci].c->id];
  }
  int64_t first_bad_atom = INT64_MAX;  // flat index of the first atom with an element this path does not know
#pragma omp parallel for schedule(static, 256) num_threads(host_threads) reduction(min : first_bad_atom) if (n_res > 4096)
  for (int64_t i = 0; i < n_res; ++i) {
    const ResT& R = *all_res[i].r;
    const uint32_t ci = all_res[i].chain, ordinal = all_res[i].ordinal;
    const ChainT& Cn = *chains[ci].c;
    const int64_t model_serial = models[chains[ci].mi].serial;
    const int32_t resn_id = uniform_name(R);
    const uint32_t group_bits = chain_bits[ci], ckey = chain_ckey[ci], rid = (uint32_t)i;
    int64_t at = res_atoms[i], ah = res_h[i], ag = res_ring[i];
    for (size_t ck = 0; ck < R.confs.size(); ++ck) {
      const ConfT& c = R.confs[ck];
      const uint32_t rc = res_rc[i] + (uint32_t)ck;
      Packed::ResConf& rcx = P.resconfs[rc];
      rcx.chain = Cn.id;
      rcx.resn = resn_id;
      rcx.icode = R.icode;
      rcx.altloc = c.altloc;
      rcx.resi = R.resi;
      rcx.model_serial = model_serial;
      rcx.has_sc = conf_ps[rc] != nullptr;
      rcx.sc = conf_ps[rc] ? *conf_ps[rc] : Plane{};
      if (conf_pr[rc]) {
        for (int k = 0; k < 3; ++k) {
          P.ring_c[3 * ag + k] = conf_pr[rc]->c[k];
          P.ring_n[3 * ag + k] = conf_pr[rc]->n[k];
        }
        P.ring_attr[ag] = group_bits;
        P.ring_chain_key[ag] = ckey;
        P.ring_res_ordinal[ag] = ordinal;
        P.ring_rc[ag] = rc;
        ++ag;
      }
      for (int64_t k : c.recs) {
        const arp_atom_record& a = recs[k];
        const int32_t atom_name = name_id[k];
        const int cls = el_class[k];
        if (cls < 0) first_bad_atom = std::min(first_bad_atom, at);
        const uint32_t by_res = bits_res[k], by_conf = by_res;  // see phase 5: the conformer's name is the residue's
        uint32_t w = (uint32_t)(cls < 0 ? 0 : cls) | group_bits;
        if (by_conf & T_POS) w |= ARP_ATTR_POS_CONF;
        if (by_conf & T_NEG) w |= ARP_ATTR_NEG_CONF;
        if (by_conf & T_DON) w |= ARP_ATTR_DONOR;
        if (by_conf & T_ACC) w |= ARP_ATTR_ACCEPTOR;
        if (by_res & T_POS) w |= ARP_ATTR_POS_RES;
        if (by_res & T_HYD) w |= ARP_ATTR_HYDROPHOBIC;
        if (cls == 1 && atom_name != id_C) w |= ARP_ATTR_WEAK_DONOR;  // hbond.rs:204-207 (class 1 = carbon)
        if (resn_id == id_CYS && atom_name == id_SG) w |= ARP_ATTR_CYS_SG;
        if (cls == 0) {  // hydrogen: joins the residue's H list (hbond.rs:38-42)
          P.hx[ah] = a.x, P.hy[ah] = a.y, P.hz[ah] = a.z;
          ++ah;
        }
        if (atom_name == id_CB && P.res_cb[rid] < 0) P.res_cb[rid] = (int32_t)at;  // vdw.rs:55-74 `.find()` = first match
        if (atom_name == id_SG && P.res_sg[rid] < 0) P.res_sg[rid] = (int32_t)at;
        P.x[at] = a.x, P.y[at] = a.y, P.z[at] = a.z;
        P.attr[at] = w;
        P.chain_key[at] = ckey;
        P.res_ordinal[at] = ordinal;
        P.res_id[at] = rid;
        P.atom_rc[at] = rc;
        P.atom_name[at] = atom_name;
        P.atom_serial[at] = a.serial;
        ++at;
      }
    }
  }
  if (first_bad_atom != INT64_MAX) {
    // the record behind that flat index: walk its residue once more
    int64_t i = (int64_t)(std::upper_bound(res_atoms.begin(), res_atoms.end(), (uint32_t)first_bad_atom) - res_atoms.begin()) - 1;
    int64_t at = res_atoms[i];
    for (auto& c : all_res[i].r->confs)
      for (int64_t k : c.recs)
        if (at++ == first_bad_atom) {
          const std::string el = fixed_field(recs[k].element, 4);
          throw Error{ARP_ERR_ELEMENT, el.empty() ? "atom " + std::to_string(recs[k].serial) + " has no element"
                                                  : "unsupported element '" + el + "' (not one of the element classes this path knows; pdbtbx would need a van der Waals radius for it, vdw.rs:28)"};
        }
  }
  // ring residues without a ring plane (mod.rs:64-68 logs them), in hierarchy order
  for (int64_t i = 0; i < n_res; ++i) {
    if (!planes[i].ring_res) continue;
    const ResT& R = *all_res[i].r;
    const ChainRef& cr = chains[all_res[i].chain];
    for (size_t ck = 0; ck < R.confs.size(); ++ck)
      if (!(conf_ring[res_rc[i] + ck] & 1) && (!quirk || cr.mi == 0))
        P.warnings.push_back("Failed to calculate ring position for chain " + P.strings[cr.c->id] + " residue " + P.strings[uniform_name(R)] + " " +
                             std::to_string(R.resi) + P.strings[R.icode]);
  }
  if (P.ring_rc.empty()) throw Error{ARP_ERR_NO_RINGS, "Error building ring positions"};  // complex.rs:50
  P.ring_name = P.intern("Ring");
  clk.mark("6 flatten");
  P.finish();
  clk.mark("7 finish");
  // the hierarchy is a few hundred thousand small vectors: freed on all threads rather than by one destructor at the
  // end, every chain by the thread that allocated most of it in phase 1 (its own malloc arena: no lock traffic)
#pragma omp parallel num_threads(host_threads) if (n > 65536)
  {
    const int me = omp_get_thread_num(), team = omp_get_num_threads();
    for (int64_t ci = 0; ci < n_chains_all; ++ci)
      if (chains[ci].c->owner % team == me) std::vector<ResT>().swap(chains[ci].c->residues);
  }
  clk.mark("8 release");
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
  const int64_t na = (int64_t)res_id.size(), nr = (int64_t)res_cb.size();
  const bool big = na > 65536 && !omp_in_parallel();  // the loops over the atoms run on every host thread
  // atoms_with_hierarchy() order keeps the atoms of a residue together: res_id as a CSR (left out if it ever does not)
  t.res_atom_begin = nullptr;
  {
    int grouped = 1;
#pragma omp parallel for schedule(static) reduction(& : grouped) if (big)
    for (int64_t a = 1; a < na; ++a) grouped &= (res_id[a] == res_id[a - 1] || res_id[a] == res_id[a - 1] + 1) ? 1 : 0;
    if (grouped && (res_id.empty() || (res_id.front() == 0 && (int64_t)res_id.back() + 1 == nr))) {
      res_atom_begin.assign((size_t)nr + 1, 0u);
#pragma omp parallel for schedule(static) if (big)
      for (int64_t a = 0; a < na; ++a)
        if (a == 0 || res_id[a] != res_id[a - 1]) res_atom_begin[res_id[a]] = (uint32_t)a;  // every residue id occurs (grouped, 0 .. nr-1)
      res_atom_begin[nr] = (uint32_t)na;
      t.res_atom_begin = res_atom_begin.data();
    }
  }
  // compact per-residue form of chain_key / res_ordinal (what the device path uploads when it is present)
  res_chain_key.assign((size_t)nr, 0u);
  res_res_ordinal.assign((size_t)nr, 0u);
  if (t.res_atom_begin) {
#pragma omp parallel for schedule(static) if (big)
    for (int64_t r = 0; r < nr; ++r) {
      const uint32_t a = res_atom_begin[r];
      res_chain_key[r] = chain_key[a];
      res_res_ordinal[r] = res_ordinal[a];
    }
  } else {
    for (size_t a = 0; a < res_id.size(); ++a) {
      res_chain_key[res_id[a]] = chain_key[a];
      res_res_ordinal[res_id[a]] = res_ordinal[a];
    }
  }
  t.res_chain_key = res_chain_key.data();
  t.res_res_ordinal = res_res_ordinal.data();
  // residue tags (arp_structure.res_tag): the residues of a chain are contiguous and in ordinal order here, so a running
  // count that steps by 1 inside a chain (consecutive ordinals) and by 2 otherwise has the required property
  res_tag.assign((size_t)nr, 0u);
  for (int64_t r = 1; r < nr; ++r)
    res_tag[r] = res_tag[r - 1] + ((res_chain_key[r] == res_chain_key[r - 1] && res_res_ordinal[r] == res_res_ordinal[r - 1] + 1u) ? 1u : 2u);
  t.res_tag = res_tag.data();
  t.atom_gid = atom_gid.empty() ? nullptr : atom_gid.data();
  t.ring_gid = ring_gid.empty() ? nullptr : ring_gid.data();
  // on-lattice form of the coordinates (arp_structure.x_milli): offered only when EVERY coordinate is bit for bit the
  // double (double)v / 1000.0 of a 32-bit v -- true for anything read from PDB fixed columns -- so the device rebuilds the
  // same f64 values from half the bytes.  The bounding-box hint for the cell-grid layout (verified on the device) comes
  // out of the same pass.
  t.x_milli = t.y_milli = t.z_milli = nullptr;
  t.has_bbox = 0;
  if (na > 0) {
    x_milli.resize((size_t)na), y_milli.resize((size_t)na), z_milli.resize((size_t)na);
    int lattice = 1, finite = 1;
    for (int axis = 0; axis < 3; ++axis) {
      const double* v = axis == 0 ? x.data() : (axis == 1 ? y.data() : z.data());
      int32_t* m = axis == 0 ? x_milli.data() : (axis == 1 ? y_milli.data() : z_milli.data());
      double lo = v[0], hi = v[0];
#pragma omp parallel for schedule(static) reduction(& : lattice, finite) reduction(min : lo) reduction(max : hi) if (big)
      for (int64_t k = 0; k < na; ++k) {
        const double c = v[k], q = c * 1000.0;
        if (!(c == c) || c - c != 0.0) finite = 0;
        lo = c < lo ? c : lo;
        hi = c > hi ? c : hi;
        int32_t r = 0;
        bool ok = q > -2147483000.0 && q < 2147483000.0;
        if (ok) {
          r = (int32_t)std::llrint(q);
          ok = (double)r / 1000.0 == c && !(r == 0 && std::signbit(c));
        }
        if (!ok) lattice = 0;
        m[k] = r;
      }
      t.bbox_min[axis] = lo;
      t.bbox_max[axis] = hi;
    }
    t.has_bbox = finite ? 1 : 0;
    if (lattice) {
      t.x_milli = x_milli.data();
      t.y_milli = y_milli.data();
      t.z_milli = z_milli.data();
    } else {
      x_milli.clear(), y_milli.clear(), z_milli.clear();
    }
  }
  // the most compact forms (arp_structure.xyz_milli21 / attr16): lattice coordinates relative to the box corner in 21 bits
  // each -- structures up to 2 097 A across -- and the attribute words in 16 bits
  t.xyz_milli21 = nullptr;
  t.attr16 = nullptr;
  t.xyz_milli21_origin[0] = t.xyz_milli21_origin[1] = t.xyz_milli21_origin[2] = 0;
  xyz21.clear();
  attr16.clear();
  if (t.x_milli && t.has_bbox) {
    int64_t o[3];
    bool fits = true;
    for (int axis = 0; axis < 3; ++axis) {
      o[axis] = std::llrint(t.bbox_min[axis] * 1000.0);  // the smallest lattice value of the axis (the map is monotone)
      if (std::llrint(t.bbox_max[axis] * 1000.0) - o[axis] >= ((int64_t)1 << 21)) fits = false;
    }
    if (fits) {
      xyz21.resize((size_t)na);
      int ok = 1;
#pragma omp parallel for schedule(static) reduction(& : ok) if (big)
      for (int64_t k = 0; k < na; ++k) {
        const int64_t dx = x_milli[k] - o[0], dy = y_milli[k] - o[1], dz = z_milli[k] - o[2];
        if ((dx | dy | dz) < 0 || ((dx | dy | dz) >> 21)) ok = 0;
        xyz21[k] = (uint64_t)dx | ((uint64_t)dy << 21) | ((uint64_t)dz << 42);
      }
      if (ok) {
        t.xyz_milli21 = xyz21.data();
        for (int axis = 0; axis < 3; ++axis) t.xyz_milli21_origin[axis] = (int32_t)o[axis];
      } else {
        xyz21.clear();
      }
    }
  }
  if (na > 0) {
    attr16.resize((size_t)na);
    int ok = 1;
#pragma omp parallel for schedule(static) reduction(& : ok) if (big)
    for (int64_t k = 0; k < na; ++k) {
      if (attr[k] >> 16) ok = 0;
      attr16[k] = (uint16_t)attr[k];
    }
    if (ok)
      t.attr16 = attr16.data();
    else
      attr16.clear();
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
  auto fix = [&](auto& v) {
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
