/*
 * arpeggia_oracle.cpp -- CPU oracle for the arpeggia `contacts` hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  A plain C++17, FP64, no-FMA
 * (-ffp-contract=off) restatement of the reference's algorithm, written from
 * the reference's Rust source.  Every function cites the reference file:line
 * (relative to /root/reference) it follows.  The reference itself is 100 %
 * Rust and cannot be compiled or imported here (no cargo/rustc, no wheel), so
 * there is no oracle/_ref.
 *
 * PARITY PINNING: pinned by tests/test_oracle_golden.py against
 * every golden value the reference's own tests hold for this path:
 *   G1 1ubq "/" 0.1 6.5 -> 532 rows x 20 columns   (python/tests/test_arpeggia.py:28-71)
 *   G2 ignore_zero_occupancy leaves 1ubq unchanged (python/tests/test_arpeggia.py:85-112)
 *   G3 6bft TYR A102 ring / ARG G82 NE -> CationPi (src/contacts/aromatic.rs:72-99)
 *   G4 6bft TRP A108 ring / LYS G84 NZ -> none     (src/contacts/aromatic.rs:101-128)
 *   G5 1ubq PHE4 ring centre + normal, MET1 atoms  (src/contacts/residues.rs:334-395)
 *   G6 Plane angle / dihedral identities           (src/contacts/residues.rs:306-332)
 *   G7 parse_groups truth table + panic messages   (src/utils.rs:174-228)
 *   G8 1ubq chain A = 76 residues then 58 waters   (src/contacts/chains.rs:21-40)
 * UNPINNED (no reference test, no fixture exercises it; the oracle follows the
 * source text and SURVEY.md 8.2 and thereby DEFINES expected behaviour):
 * hydrogen-bond angle branches and SaltBridge (no H in the fixtures), Disulfide
 * vs CovalentBond, StericClash, covalent radii, altloc / insertion-code /
 * multi-model handling, the inclusive cutoff tie, sc_* values, and the last
 * bits of the SVD plane normal (nalgebra is not available).
 *
 * Third-party arithmetic restated here because it is absent from /root/reference
 * (Cargo.toml:19-40, no Cargo.lock): pdbtbx 0.12 (Atom::distance/angle/dihedral,
 * element radii, PDB hierarchy grouping, rtree adaptor), rstar 0.12
 * (locate_within_distance: distance_2 <= r^2, inclusive), nalgebra 0.33 (SVD).
 */
#include "oracle.h"

#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <unordered_map>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

thread_local std::string g_err;

struct OrcError {
  int code;
  std::string msg;
};

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

std::string fixed_str(const char* p, size_t cap) {
  size_t n = 0;
  while (n < cap && p[n] != '\0') ++n;
  return std::string(p, n);
}

/* ------------------------------------------------------------------------------------------
 * Hierarchy: Model -> Chain -> Residue -> Conformer -> Atom, pdbtbx 0.12 grouping rules
 * (SURVEY.md 8.2): chain by id within a model, residue by (resi, icode) within a chain,
 * conformer by (name, altloc) within a residue, all in order of first appearance.
 * ------------------------------------------------------------------------------------------ */
struct Atom {
  int64_t serial;
  std::string name, element;
  double x, y, z, occ;
};
struct Conformer {
  std::string name, altloc;
  std::vector<Atom> atoms;
};
struct Residue {
  int64_t resi;
  std::string icode;
  std::vector<Conformer> confs;
  /* pdbtbx Residue::name(): Some(name) iff every conformer has the same name. */
  bool name(std::string& out) const {
    if (confs.empty()) return false;
    for (const auto& c : confs)
      if (c.name != confs[0].name) return false;
    out = confs[0].name;
    return true;
  }
};
struct Chain {
  std::string id;
  std::vector<Residue> residues;
  std::map<std::pair<int64_t, std::string>, size_t> index;  // lookup accelerator only
};
struct Model {
  int64_t serial;
  std::vector<Chain> chains;
  std::unordered_map<std::string, size_t> index;  // lookup accelerator only
};
struct PDB {
  std::vector<Model> models;
};

PDB build_hierarchy(const orc_atom_record* recs, int64_t n) {
  PDB pdb;
  Model* cur_m = nullptr;
  Chain* cur_c = nullptr;
  Residue* cur_r = nullptr;
  for (int64_t k = 0; k < n; ++k) {
    const orc_atom_record& r = recs[k];
    if (!cur_m || cur_m->serial != r.model) {
      cur_m = nullptr;
      for (auto& m : pdb.models)
        if (m.serial == r.model) cur_m = &m;
      if (!cur_m) {
        pdb.models.reserve(pdb.models.size() + 1);
        pdb.models.push_back(Model{r.model, {}, {}});
        cur_m = &pdb.models.back();
      }
      cur_c = nullptr;
      cur_r = nullptr;
    }
    std::string chain = fixed_str(r.chain, sizeof r.chain);
    if (!cur_c || cur_c->id != chain) {
      auto it = cur_m->index.find(chain);
      if (it == cur_m->index.end()) {
        cur_m->index.emplace(chain, cur_m->chains.size());
        cur_m->chains.push_back(Chain{chain, {}, {}});
        cur_c = &cur_m->chains.back();
      } else {
        cur_c = &cur_m->chains[it->second];
      }
      cur_r = nullptr;
    }
    std::string icode = fixed_str(r.icode, sizeof r.icode);
    if (!cur_r || cur_r->resi != r.resi || cur_r->icode != icode) {
      auto key = std::make_pair((int64_t)r.resi, icode);
      auto it = cur_c->index.find(key);
      if (it == cur_c->index.end()) {
        cur_c->index.emplace(key, cur_c->residues.size());
        cur_c->residues.push_back(Residue{r.resi, icode, {}});
        cur_r = &cur_c->residues.back();
      } else {
        cur_r = &cur_c->residues[it->second];
      }
    }
    std::string resn = fixed_str(r.resn, sizeof r.resn);
    std::string altloc = fixed_str(r.altloc, sizeof r.altloc);
    Conformer* cf = nullptr;
    for (auto& c : cur_r->confs)
      if (c.name == resn && c.altloc == altloc) cf = &c;
    if (!cf) {
      cur_r->confs.push_back(Conformer{resn, altloc, {}});
      cf = &cur_r->confs.back();
    }
    cf->atoms.push_back(Atom{r.serial, fixed_str(r.name, sizeof r.name), fixed_str(r.element, sizeof r.element),
                             r.x, r.y, r.z, r.occupancy});
    /* vectors may have reallocated: re-derive raw pointers lazily next iteration */
    (void)0;
  }
  /* pdbtbx PDB reader post-pass: in a residue with several conformers the atoms of
   * the blank-altloc conformer are appended to every other conformer and the blank one is removed. */
  for (auto& m : pdb.models)
    for (auto& c : m.chains)
      for (auto& r : c.residues) {
        if (r.confs.size() <= 1) continue;
        int blank = -1;
        for (size_t i = 0; i < r.confs.size(); ++i)
          if (r.confs[i].altloc.empty()) blank = (int)i;
        if (blank < 0) continue;
        Conformer shared = r.confs[blank];
        r.confs.erase(r.confs.begin() + blank);
        for (auto& cf : r.confs) cf.atoms.insert(cf.atoms.end(), shared.atoms.begin(), shared.atoms.end());
      }
  return pdb;
}

/* src/contacts/residues.rs:131-161 ResidueExt::resn(): 20 amino acids + HOH, on the upper-cased name. */
bool is_known_residue(const std::string& name) {
  static const std::set<std::string> known = {"ALA", "ARG", "ASN", "ASP", "CYS", "GLN", "GLU", "GLY", "HIS", "ILE", "LEU",
                                              "LYS", "MET", "PHE", "PRO", "SER", "THR", "TRP", "TYR", "VAL", "HOH"};
  std::string up = name;
  for (auto& ch : up) ch = (char)std::toupper((unsigned char)ch);
  return known.count(up) != 0;
}

/* src/utils.rs:51-63 load_model: pdb.remove_residues_by(|res| res.resn().is_none()).
 * resn() unwraps Residue::name() (residues.rs:132) -> panic when conformer names differ. */
void load_model_filter(PDB& pdb) {
  for (auto& m : pdb.models)
    for (auto& c : m.chains) {
      std::vector<Residue> kept;
      for (auto& r : c.residues) {
        std::string nm;
        if (!r.name(nm)) throw OrcError{ORC_ERR_RESIDUE_NAME, "residue has conformers with different names (Residue::name() is None)"};
        if (is_known_residue(nm)) kept.push_back(std::move(r));
      }
      c.residues.swap(kept);
      c.index.clear();
    }
}

/* src/python.rs:45-47 / src/cli/contacts.rs:87-90: pdb.remove_atoms_by(|a| a.occupancy() == 0.0).
 * pdbtbx remove_atoms_by is a retain on the atom lists; emptied conformers / residues stay (unpinned). */
void remove_zero_occupancy(PDB& pdb) {
  for (auto& m : pdb.models)
    for (auto& c : m.chains)
      for (auto& r : c.residues)
        for (auto& cf : r.confs) {
          std::vector<Atom> kept;
          for (auto& a : cf.atoms)
            if (!(a.occ == 0.0)) kept.push_back(std::move(a));
          cf.atoms.swap(kept);
        }
}

/* ------------------------------------------------------------------------------------------
 * Element radii: pdbtbx Element::atomic_radius().  van der Waals = Alvarez 2013,
 * covalent single = Pyykko 2009.  C/N/O/S vdW are pinned by G1 (only these give 532 rows).
 * ------------------------------------------------------------------------------------------ */
struct Radii {
  double cov, vdw;
};
bool element_radii(const std::string& el, Radii& out) {
  static const std::map<std::string, Radii> tab = {
      {"H", {0.32, 1.20}},  {"C", {0.75, 1.77}},  {"N", {0.71, 1.66}},  {"O", {0.63, 1.50}},
      {"S", {1.03, 1.89}},  {"P", {1.11, 1.90}},  {"SE", {1.16, 1.82}}, {"F", {0.64, 1.46}},
      {"CL", {0.99, 1.82}}, {"BR", {1.14, 1.86}}, {"I", {1.33, 2.04}},
  };
  auto it = tab.find(el);
  if (it == tab.end()) return false;
  out = it->second;
  return true;
}

/* ------------------------------------------------------------------------------------------
 * String tables, restated one-to-one.
 * ------------------------------------------------------------------------------------------ */
bool in(const std::string& s, std::initializer_list<const char*> l) {
  for (auto p : l)
    if (s == p) return true;
  return false;
}
/* src/contacts/ionic.rs:83-91 */
bool is_pos_ionizable(const std::string& res, const std::string& atom) {
  return (res == "ARG" && in(atom, {"NE", "CZ", "NH1", "NH2"})) || (res == "HIS" && in(atom, {"CG", "ND1", "CE1", "NE2", "CD2"})) ||
         (res == "LYS" && atom == "NZ");
}
/* src/contacts/ionic.rs:94-99 */
bool is_neg_ionizable(const std::string& res, const std::string& atom) {
  return (res == "ASP" && in(atom, {"OD1", "OD2"})) || (res == "GLU" && in(atom, {"OE1", "OE2"}));
}
/* src/contacts/hbond.rs:137-159 */
bool is_hydrogen_acceptor(const std::string& res, const std::string& atom) {
  if (in(atom, {"O", "OXT"}) && res != "HOH") return true;
  return (res == "ASN" && atom == "OD1") || (res == "ASP" && in(atom, {"OD1", "OD2"})) || (res == "GLN" && atom == "OE1") ||
         (res == "GLU" && in(atom, {"OE1", "OE2"})) || (res == "HIS" && in(atom, {"ND1", "NE2"})) || (res == "SER" && atom == "OG") ||
         (res == "THR" && atom == "OG1") || (res == "TYR" && atom == "OH") || (res == "MET" && atom == "SD") ||
         (res == "CYS" && atom == "SG");
}
/* src/contacts/hbond.rs:162-181 */
bool is_hydrogen_donor(const std::string& res, const std::string& atom) {
  if (atom == "N") return true;
  return (res == "ARG" && in(atom, {"NE", "NH1", "NH2"})) || (res == "ASN" && atom == "ND2") || (res == "GLN" && atom == "NE2") ||
         (res == "HIS" && in(atom, {"ND1", "NE2"})) || (res == "LYS" && atom == "NZ") || (res == "SER" && atom == "OG") ||
         (res == "THR" && atom == "OG1") || (res == "TRP" && atom == "NE1") || (res == "TYR" && atom == "OH") ||
         (res == "CYS" && atom == "SG");
}
/* src/contacts/hydrophobic.rs:27-45 */
bool is_hydrophobic(const std::string& res, const std::string& atom) {
  if (atom == "CB" && res != "SER") return true;
  return ((res == "ARG" || res == "GLN" || res == "GLU" || res == "PRO") && atom == "CG") ||
         (res == "ILE" && in(atom, {"CG1", "CD1", "CG2"})) || (res == "LEU" && in(atom, {"CG", "CD1", "CD2"})) ||
         (res == "LYS" && in(atom, {"CG", "CD"})) || (res == "MET" && in(atom, {"CG", "CE", "SD"})) ||
         (res == "PHE" && in(atom, {"CG", "CD1", "CD2", "CE1", "CE2", "CZ"})) || (res == "THR" && atom == "CG2") ||
         (res == "TRP" && in(atom, {"CG", "CD2", "CE3", "CZ3", "CH2", "CZ2"})) ||
         (res == "TYR" && in(atom, {"CG", "CD1", "CD2", "CE1", "CE2"})) || (res == "VAL" && in(atom, {"CG1", "CG2"}));
}
/* src/contacts/residues.rs:163-186 ring_atoms(): name filter per residue name */
bool is_ring_atom(const std::string& res, const std::string& atom) {
  if (res == "HIS") return in(atom, {"CG", "ND1", "CE1", "NE2", "CD2"});
  if (res == "PHE" || res == "TYR") return in(atom, {"CG", "CD1", "CD2", "CE1", "CE2", "CZ"});
  if (res == "TRP") return in(atom, {"CG", "CD1", "CD2", "NE1", "CE2", "CE3", "CZ2", "CZ3", "CH2"});
  return false;
}
/* src/contacts/residues.rs:188-268 sc_plane_atoms() */
bool is_sc_plane_atom(const std::string& res, const std::string& atom) {
  if (res == "ARG") return in(atom, {"NE", "CZ", "NH1", "NH2"});
  if (res == "ASN") return in(atom, {"CB", "CG", "OD1", "ND2"});
  if (res == "ASP") return in(atom, {"CB", "CG", "OD1", "OD2"});
  if (res == "CYS") return in(atom, {"CA", "CB", "SG"});
  if (res == "GLU") return in(atom, {"CG", "CD", "OE1", "OE2"});
  if (res == "GLN") return in(atom, {"CG", "CD", "OE1", "NE2"});
  if (res == "HIS") return in(atom, {"CG", "ND1", "CE1", "NE2", "CD2"});
  if (res == "ILE") return in(atom, {"CB", "CG1", "CG2", "CD1"});
  if (res == "LEU") return in(atom, {"CB", "CG", "CD1", "CD2"});
  if (res == "LYS") return in(atom, {"CG", "CD", "CE", "NZ"});
  if (res == "MET") return in(atom, {"CG", "SD", "CE"});
  if (res == "PHE" || res == "TYR") return in(atom, {"CG", "CD1", "CD2", "CE1", "CE2", "CZ"});
  if (res == "PRO") return in(atom, {"N", "CA", "CB", "CG", "CD"});
  if (res == "SER") return in(atom, {"CA", "CB", "OG"});
  if (res == "THR") return in(atom, {"CA", "CB", "OG1", "CG2"});
  if (res == "TRP") return in(atom, {"CG", "CD1", "CD2", "NE1", "CE2", "CE3", "CZ2", "CZ3", "CH2"});
  if (res == "VAL") return in(atom, {"CA", "CB", "CG1", "CG2"});
  return false;
}

/* ------------------------------------------------------------------------------------------
 * Plane maths.
 * ------------------------------------------------------------------------------------------ */
struct Plane {
  double c[3], n[3];
};
const double kPi = 3.14159265358979323846;
const double kFracPi2 = 1.57079632679489661923;
inline double to_degrees(double rad) { return rad * (180.0 / kPi); } /* Rust f64::to_degrees */
inline double dot3(const double a[3], const double b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline double norm3(const double a[3]) { return std::sqrt(dot3(a, a)); }

/* src/contacts/residues.rs:31-35 */
double point_vec_dist(const Plane& p, const double pt[3]) {
  double v[3] = {pt[0] - p.c[0], pt[1] - p.c[1], pt[2] - p.c[2]};
  return norm3(v);
}
/* src/contacts/residues.rs:45-54 */
double point_vec_angle(const Plane& p, const double pt[3]) {
  double v[3] = {pt[0] - p.c[0], pt[1] - p.c[1], pt[2] - p.c[2]};
  double rad = std::acos(dot3(p.n, v) / (norm3(p.n) * norm3(v)));
  if (rad > kFracPi2) rad = kPi - rad;
  return to_degrees(rad);
}
/* src/contacts/residues.rs:65-75 */
double plane_dihedral(const double n1[3], const double n2[3]) {
  double rad = std::acos(dot3(n1, n2) / (norm3(n1) * norm3(n2)));
  if (rad > kFracPi2) rad = kPi - rad;
  return to_degrees(rad);
}

/* src/contacts/residues.rs:270-298 center_and_normal: centroid = column mean; normal = left singular
 * vector of the centred 3 x n matrix for the smallest singular value (nalgebra svd, u.column(2)).
 * nalgebra is unavailable: the SVD is a one-sided (Hestenes) Jacobi on the three rows, cyclic
 * (0,1),(0,2),(1,2) sweeps; the sign of the normal is arbitrary (every consumer folds to <= 90 deg). */
bool plane_fit(const std::vector<std::array<double, 3>>& pts, Plane& out) {
  const size_t n = pts.size();
  if (n < 3) return false;
  const double inv_n = 1.0 / (double)n;
  double c[3] = {0.0, 0.0, 0.0};
  for (size_t i = 0; i < n; ++i)
    for (int k = 0; k < 3; ++k) c[k] = c[k] + inv_n * pts[i][k];
  std::vector<double> row[3];
  for (int k = 0; k < 3; ++k) {
    row[k].resize(n);
    for (size_t i = 0; i < n; ++i) row[k][i] = pts[i][k] - c[k];
  }
  double U[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  static const int pairs[3][2] = {{0, 1}, {0, 2}, {1, 2}};
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rotated = false;
    for (auto& pq : pairs) {
      const int p = pq[0], q = pq[1];
      double alpha = 0.0, beta = 0.0, gamma = 0.0;
      for (size_t i = 0; i < n; ++i) {
        alpha = alpha + row[p][i] * row[p][i];
        beta = beta + row[q][i] * row[q][i];
        gamma = gamma + row[p][i] * row[q][i];
      }
      if (gamma == 0.0 || std::fabs(gamma) <= 1e-300 + 2.220446049250313e-16 * std::sqrt(alpha * beta)) continue;
      rotated = true;
      const double zeta = (beta - alpha) / (2.0 * gamma);
      const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
      const double cs = 1.0 / std::sqrt(1.0 + t * t);
      const double sn = cs * t;
      for (size_t i = 0; i < n; ++i) {
        const double rp = row[p][i], rq = row[q][i];
        row[p][i] = cs * rp - sn * rq;
        row[q][i] = sn * rp + cs * rq;
      }
      for (int k = 0; k < 3; ++k) {
        const double up = U[k][p], uq = U[k][q];
        U[k][p] = cs * up - sn * uq;
        U[k][q] = sn * up + cs * uq;
      }
    }
    if (!rotated) break;
  }
  double s2[3];
  for (int k = 0; k < 3; ++k) {
    s2[k] = 0.0;
    for (size_t i = 0; i < n; ++i) s2[k] = s2[k] + row[k][i] * row[k][i];
  }
  int m = 0;
  if (s2[1] < s2[m]) m = 1;
  if (s2[2] < s2[m]) m = 2;
  for (int k = 0; k < 3; ++k) {
    out.c[k] = c[k];
    out.n[k] = U[k][m];
  }
  return true;
}

/* ------------------------------------------------------------------------------------------
 * Flattened view: one entry per (atom, conformer) = pdbtbx atoms_with_hierarchy(), in its order
 * (model, chain, residue, conformer, atom).
 * ------------------------------------------------------------------------------------------ */
struct ResConf {  // a ResidueId (src/contacts/residues.rs:7-21): one per conformer of a residue
  int model_idx;
  int64_t model_serial;
  int chain_idx;  // index into Complex::chain_ids (string)
  int64_t resi;
  std::string icode, altloc, resn;
  int ordinal;  // res2idx value, complex.rs:411-440
  int residue;  // global residue index
  bool has_ring = false, has_sc = false;
  Plane ring{}, sc{};
};
struct FlatAtom {
  double x, y, z;
  int rc;  // index into resconfs
  int64_t serial;
  std::string name, element;
  Radii radii;
  bool is_h, lig, rec, pos_conf, neg_conf, pos_res, donor, acceptor, weak_donor, hydrophobic, cys_sg;
};
struct ResidueSide {  // per residue: hydrogens of all conformers, first CB / first SG (vdw.rs:55-74)
  std::vector<std::array<double, 3>> h;
  int cb = -1, sg = -1;  // flat atom index
};

struct Row {
  int kind;  // 0 atom-atom, 1 ring-atom, 2 ring-ring
  int64_t a, b;
  uint8_t code;
  double dist;
};

struct Complex {
  std::vector<std::string> chain_ids;  // distinct chain id strings
  std::vector<ResConf> resconfs;
  std::vector<FlatAtom> atoms;
  std::vector<ResidueSide> residues;
  std::vector<int> rings;  // resconf indices that carry a ring plane
  std::set<std::string> ligand, receptor;
  std::vector<std::string> ring_warnings;
  double vdw_comp, rc;
};

/* src/utils.rs:71-115 parse_groups */
void parse_groups(const std::set<std::string>& all, const std::string& groups, std::set<std::string>& lig, std::set<std::string>& rec) {
  std::vector<std::string> sel;
  {
    size_t pos = 0;
    while (true) {
      size_t q = groups.find('/', pos);
      if (q == std::string::npos) {
        sel.push_back(groups.substr(pos));
        break;
      }
      sel.push_back(groups.substr(pos, q - pos));
      pos = q + 1;
    }
  }
  if (sel.size() < 2) throw OrcError{ORC_ERR_BAD_GROUPS, "Invalid chain groups format! Use '/' for all-to-all comparisons."};
  auto split_set = [](const std::string& s) {
    std::set<std::string> out;
    size_t pos = 0;
    while (true) {
      size_t q = s.find(',', pos);
      std::string tok = s.substr(pos, q == std::string::npos ? std::string::npos : q - pos);
      if (!tok.empty()) out.insert(tok);
      if (q == std::string::npos) break;
      pos = q + 1;
    }
    return out;
  };
  lig = split_set(sel[0]);
  rec = split_set(sel[1]);
  if (lig.empty() && rec.empty()) {
    lig = all;
    rec = all;
    return;
  }
  if (lig.empty()) {
    for (auto& c : all)
      if (!rec.count(c)) lig.insert(c);
  } else if (rec.empty()) {
    for (auto& c : all)
      if (!lig.count(c)) rec.insert(c);
  }
  if (lig.empty() || rec.empty()) throw OrcError{ORC_ERR_EMPTY_GROUP, "Empty chain groups!"};
}

std::vector<std::array<double, 3>> residue_atoms_matching(const Residue& r, const std::string& resn, bool ring) {
  std::vector<std::array<double, 3>> pts;  // Residue::par_atoms(): every conformer, file order
  for (auto& cf : r.confs)
    for (auto& a : cf.atoms)
      if (ring ? is_ring_atom(resn, a.name) : is_sc_plane_atom(resn, a.name)) pts.push_back({a.x, a.y, a.z});
  return pts;
}

/* InteractionComplex::new, src/contacts/complex.rs:36-68 */
void build_complex(const PDB& pdb, const std::string& groups, double vdw_comp, double rc, uint32_t flags, Complex& cx) {
  cx.vdw_comp = vdw_comp;
  cx.rc = rc;
  /* complex.rs:43 all_chains over every model */
  std::set<std::string> all_chains;
  for (auto& m : pdb.models)
    for (auto& c : m.chains) all_chains.insert(c.id);
  parse_groups(all_chains, groups, cx.ligand, cx.receptor);
  std::unordered_map<std::string, int> chain_index;
  for (auto& id : all_chains) {
    chain_index.emplace(id, (int)cx.chain_ids.size());
    cx.chain_ids.push_back(id);
  }

  /* planes keyed the way complex.rs:442-514 keys them.  With several models the reference iterates
   * `for m in model.models() { for c in model.chains()` where `model` is the whole PDB, so every
   * model's key ends up with the geometry of the LAST model that has this residue (ledger item 12). */
  using PKey = std::tuple<std::string, int64_t, std::string, std::string, std::string>;  // chain, resi, icode, altloc, resn
  const bool quirk = pdb.models.size() > 1 && !(flags & ORC_FLAG_NO_MULTIMODEL_QUIRK);
  std::map<PKey, std::pair<bool, Plane>> last_ring, last_sc;
  std::map<PKey, bool> ring_attempt;
  if (quirk) {
    for (auto& m : pdb.models)
      for (auto& c : m.chains)
        for (auto& r : c.residues) {
          std::string resn;
          r.name(resn);
          Plane pr{}, ps{};
          const bool is_ring_res = in(resn, {"HIS", "PHE", "TYR", "TRP"});
          const bool okr = is_ring_res && plane_fit(residue_atoms_matching(r, resn, true), pr);
          const bool oks = plane_fit(residue_atoms_matching(r, resn, false), ps);
          for (auto& cf : r.confs) {
            PKey k{c.id, r.resi, r.icode, cf.altloc, resn};
            if (is_ring_res) {
              if (okr) last_ring[k] = {true, pr};  // insert() only on Some: an earlier success survives a later failure
              ring_attempt[k] = true;
            }
            if (oks) last_sc[k] = {true, ps};
          }
        }
  }

  int residue_counter = 0;
  for (size_t mi = 0; mi < pdb.models.size(); ++mi) {
    const Model& m = pdb.models[mi];
    for (auto& c : m.chains) {
      const bool lig = cx.ligand.count(c.id) != 0, rec = cx.receptor.count(c.id) != 0;
      for (size_t ri = 0; ri < c.residues.size(); ++ri) {
        const Residue& r = c.residues[ri];
        std::string resn;
        const bool has_name = r.name(resn);
        const int residue = residue_counter++;
        cx.residues.emplace_back();
        Plane pr{}, ps{};
        bool okr = false, oks = false;
        const bool is_ring_res = has_name && in(resn, {"HIS", "PHE", "TYR", "TRP"});
        if (!quirk) {
          if (is_ring_res) okr = plane_fit(residue_atoms_matching(r, resn, true), pr);
          oks = plane_fit(residue_atoms_matching(r, resn, false), ps);
        }
        for (auto& cf : r.confs) {
          ResConf rc_;
          rc_.model_idx = (int)mi;
          rc_.model_serial = m.serial;
          rc_.chain_idx = chain_index[c.id];
          rc_.resi = r.resi;
          rc_.icode = r.icode;
          rc_.altloc = cf.altloc;
          rc_.resn = resn;  // residue().name().unwrap_or("")
          rc_.ordinal = (int)ri;
          rc_.residue = residue;
          if (quirk) {
            PKey k{c.id, r.resi, r.icode, cf.altloc, resn};
            auto ir = last_ring.find(k);
            if (ir != last_ring.end()) {
              rc_.has_ring = true;
              rc_.ring = ir->second.second;
            } else if (is_ring_res && mi == 0) {
              cx.ring_warnings.push_back("Failed to calculate ring position for " + c.id + " " + std::to_string(r.resi));
            }
            auto is = last_sc.find(k);
            if (is != last_sc.end()) {
              rc_.has_sc = true;
              rc_.sc = is->second.second;
            }
          } else {
            rc_.has_ring = okr;
            rc_.ring = pr;
            rc_.has_sc = oks;
            rc_.sc = ps;
            if (is_ring_res && !okr)
              cx.ring_warnings.push_back("Failed to calculate ring position for " + c.id + " " + std::to_string(r.resi));
          }
          const int rc_index = (int)cx.resconfs.size();
          if (rc_.has_ring) cx.rings.push_back(rc_index);
          cx.resconfs.push_back(rc_);
          for (auto& a : cf.atoms) {
            FlatAtom f;
            f.x = a.x;
            f.y = a.y;
            f.z = a.z;
            f.rc = rc_index;
            f.serial = a.serial;
            f.name = a.name;
            f.element = a.element;
            /* atom.element().unwrap() (complex.rs:83, vdw.rs:24) and van_der_waals.unwrap() (vdw.rs:28) */
            if (a.element.empty()) throw OrcError{ORC_ERR_ELEMENT, "atom " + std::to_string(a.serial) + " has no element"};
            if (!element_radii(a.element, f.radii))
              throw OrcError{ORC_ERR_ELEMENT, "element '" + a.element + "' has no van der Waals radius"};
            f.is_h = a.element == "H";
            f.lig = lig;
            f.rec = rec;
            f.pos_conf = is_pos_ionizable(cf.name, a.name);  // ionic.rs:45-46 conformer().name()
            f.neg_conf = is_neg_ionizable(cf.name, a.name);
            f.pos_res = is_pos_ionizable(resn, a.name);  // aromatic.rs:18 residue().name()
            f.donor = is_hydrogen_donor(cf.name, a.name);  // hbond.rs:120-121 conformer().name()
            f.acceptor = is_hydrogen_acceptor(cf.name, a.name);
            f.weak_donor = a.element == "C" && a.name != "C";  // hbond.rs:204-207
            f.hydrophobic = is_hydrophobic(resn, a.name);      // hydrophobic.rs:16-17 residue().name()
            f.cys_sg = resn == "CYS" && a.name == "SG";        // vdw.rs:50-53
            const int idx = (int)cx.atoms.size();
            ResidueSide& rs = cx.residues[residue];
            if (f.is_h) rs.h.push_back({a.x, a.y, a.z});
            if (rs.cb < 0 && a.name == "CB") rs.cb = idx;
            if (rs.sg < 0 && a.name == "SG") rs.sg = idx;
            cx.atoms.push_back(std::move(f));
          }
        }
      }
    }
  }
  /* complex.rs:50, 480-482: no ring at all -> Err -> .expect() panics */
  if (cx.rings.empty()) throw OrcError{ORC_ERR_NO_RINGS, "Error building ring positions"};
}

/* src/contacts/complex.rs:94-131 should_compare_residues */
bool should_compare_residues(const Complex& cx, const ResConf& r1, const ResConf& r2, bool symmetric) {
  if (r1.model_serial != r2.model_serial) return false;
  const std::string& c1 = cx.chain_ids[r1.chain_idx];
  const std::string& c2 = cx.chain_ids[r2.chain_idx];
  const bool l1 = cx.ligand.count(c1), l2 = cx.ligand.count(c2), k1 = cx.receptor.count(c1), k2 = cx.receptor.count(c2);
  if (!((l1 && k2) | (l2 && k1))) return false;
  if (c1 == c2) {
    const size_t e1 = (size_t)r1.ordinal, e2 = (size_t)r2.ordinal;
    if (symmetric) return (e2 > 1) && (e1 < e2 - 1);
    bool neighbouring;
    if (e1 == 0)
      neighbouring = (e2 == e1) | (e2 == e1 + 1);
    else
      neighbouring = (e2 == e1 - 1) | (e2 == e1) | (e2 == e1 + 1);
    return !neighbouring;
  }
  return !(symmetric && k1 && k2 && l1 && l2 && (c1 > c2));
}

/* src/contacts/complex.rs:76-92 should_compare_entities */
bool should_compare_entities(const Complex& cx, const FlatAtom& e1, const FlatAtom& e2, bool symmetric) {
  if (e1.is_h | e2.is_h) return false;
  return should_compare_residues(cx, cx.resconfs[e1.rc], cx.resconfs[e2.rc], symmetric);
}

/* pdbtbx Atom::distance */
inline double atom_distance(const FlatAtom& a, const FlatAtom& b) {
  const double dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z;
  return std::sqrt(dx * dx + dy * dy + dz * dz);
}
inline double point_distance(const double a[3], const double b[3]) {
  const double dx = b[0] - a[0], dy = b[1] - a[1], dz = b[2] - a[2];
  return std::sqrt(dx * dx + dy * dy + dz * dz);
}
/* pdbtbx Atom::angle(self, a2, a3): angle at a2, degrees (unpinned) */
double atom_angle(const double a[3], const double b[3], const double c[3]) {
  const double ba[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
  const double bc[3] = {c[0] - b[0], c[1] - b[1], c[2] - b[2]};
  return to_degrees(std::acos(dot3(ba, bc) / (norm3(ba) * norm3(bc))));
}
/* pdbtbx Atom::dihedral(self, a2, a3, a4): angle between the normals of planes (1,2,3) and (2,3,4),
 * acos based, degrees in [0,180] (unpinned) */
double atom_dihedral(const double a[3], const double b[3], const double c[3], const double d[3]) {
  const double ba[3] = {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
  const double bc[3] = {c[0] - b[0], c[1] - b[1], c[2] - b[2]};
  const double cb[3] = {b[0] - c[0], b[1] - c[1], b[2] - c[2]};
  const double cd[3] = {d[0] - c[0], d[1] - c[1], d[2] - c[2]};
  const double n1[3] = {ba[1] * bc[2] - ba[2] * bc[1], ba[2] * bc[0] - ba[0] * bc[2], ba[0] * bc[1] - ba[1] * bc[0]};
  const double n2[3] = {cb[1] * cd[2] - cb[2] * cd[1], cb[2] * cd[0] - cb[0] * cd[2], cb[0] * cd[1] - cb[1] * cd[0]};
  return to_degrees(std::acos(dot3(n1, n2) / (norm3(n1) * norm3(n2))));
}

/* src/contacts/vdw.rs:46-80 is_disulfide */
bool is_disulfide(const Complex& cx, const FlatAtom& e1, const FlatAtom& e2) {
  if (!(e1.cys_sg && e2.cys_sg)) return false;
  const ResidueSide& r1 = cx.residues[cx.resconfs[e1.rc].residue];
  const ResidueSide& r2 = cx.residues[cx.resconfs[e2.rc].residue];
  if (r1.cb < 0 || r2.cb < 0) throw OrcError{ORC_ERR_ARG, "CYS residue without CB (vdw.rs:58 unwrap)"};
  const FlatAtom &cb1 = cx.atoms[r1.cb], &s1 = cx.atoms[r1.sg], &cb2 = cx.atoms[r2.cb], &s2 = cx.atoms[r2.sg];
  const double a[3] = {cb1.x, cb1.y, cb1.z}, b[3] = {s1.x, s1.y, s1.z}, c[3] = {s2.x, s2.y, s2.z}, d[3] = {cb2.x, cb2.y, cb2.z};
  const double dih = std::fabs(atom_dihedral(a, b, c, d));
  return dih >= 60.0 && dih <= 120.0;
}

/* src/contacts/vdw.rs:19-44 find_vdw_contact; returns -1 for None */
int find_vdw_contact(const Complex& cx, const FlatAtom& e1, const FlatAtom& e2) {
  const double sum_cov = e1.radii.cov + e2.radii.cov;
  const double sum_vdw = e1.radii.vdw + e2.radii.vdw;
  const double d = atom_distance(e1, e2);
  if (d < sum_cov - cx.vdw_comp) return ORC_StericClash;
  if (d < sum_cov + cx.vdw_comp) return is_disulfide(cx, e1, e2) ? ORC_Disulfide : ORC_CovalentBond;
  if (d < sum_vdw + cx.vdw_comp) return ORC_VanDerWaalsContact;
  return -1;
}

/* src/contacts/ionic.rs:11-22, 38-58 */
int find_ionic_bond(const FlatAtom& e1, const FlatAtom& e2) {
  const FlatAtom *pos = nullptr, *neg = nullptr;
  if (e1.pos_conf && e2.neg_conf) {
    pos = &e1;
    neg = &e2;
  } else if (e2.pos_conf & e1.neg_conf) {
    pos = &e2;
    neg = &e1;
  } else
    return -1;
  return atom_distance(*pos, *neg) <= 4.0 ? ORC_IonicBond : -1;
}
/* src/contacts/ionic.rs:25-35, 60-80 */
int find_ionic_repulsion(const FlatAtom& e1, const FlatAtom& e2) {
  const bool both_pos = e1.pos_conf && e2.pos_conf, both_neg = e1.neg_conf && e2.neg_conf;
  if (!(both_pos | both_neg)) return -1;
  return atom_distance(e1, e2) <= 4.0 ? ORC_IonicRepulsion : -1;
}

/* shared body of src/contacts/hbond.rs:30-66 and :74-110 */
int hbond_body(const Complex& cx, const FlatAtom& donor, const FlatAtom& acceptor, double min_angle, int strong, int polar) {
  const double da = atom_distance(donor, acceptor);
  if (da <= 4.0) {
    const ResidueSide& rs = cx.residues[cx.resconfs[donor.rc].residue];
    const double acceptor_vdw = acceptor.radii.vdw;
    Radii hr;
    element_radii("H", hr);
    const double h_vdw = hr.vdw;
    const double D[3] = {donor.x, donor.y, donor.z}, A[3] = {acceptor.x, acceptor.y, acceptor.z};
    for (auto& h : rs.h) {
      const double H[3] = {h[0], h[1], h[2]};
      if (point_distance(H, A) <= h_vdw + acceptor_vdw + cx.vdw_comp && atom_angle(D, H, A) >= min_angle) return strong;
    }
  }
  if (da <= 3.5) return polar;
  return -1;
}
/* src/contacts/hbond.rs:30-66 + is_donor_acceptor_pair :112-134 */
int find_hydrogen_bond(const Complex& cx, const FlatAtom& e1, const FlatAtom& e2) {
  if (e1.donor && e2.acceptor) return hbond_body(cx, e1, e2, 90.0, ORC_HydrogenBond, ORC_PolarContact);
  if (e2.donor && e1.acceptor) return hbond_body(cx, e2, e1, 90.0, ORC_HydrogenBond, ORC_PolarContact);
  return -1;
}
/* src/contacts/hbond.rs:74-110 + is_weak_donor_acceptor_pair :183-201 */
int find_weak_hydrogen_bond(const Complex& cx, const FlatAtom& e1, const FlatAtom& e2) {
  if (e1.weak_donor && e2.acceptor) return hbond_body(cx, e1, e2, 130.0, ORC_WeakHydrogenBond, ORC_WeakPolarContact);
  if (e2.weak_donor && e1.acceptor) return hbond_body(cx, e2, e1, 130.0, ORC_WeakHydrogenBond, ORC_WeakPolarContact);
  return -1;
}
/* src/contacts/hydrophobic.rs:10-24 */
int find_hydrophobic_contact(const FlatAtom& e1, const FlatAtom& e2) {
  return (e1.hydrophobic && e2.hydrophobic && atom_distance(e1, e2) <= 4.5) ? ORC_HydrophobicContact : -1;
}

/* classification closure of get_atomic_contacts, src/contacts/complex.rs:217-296 */
void classify_pair(const Complex& cx, int64_t i, int64_t j, std::vector<Row>& out) {
  const FlatAtom &e1 = cx.atoms[i], &e2 = cx.atoms[j];
  const double dist = atom_distance(e1, e2);
  const int vdw = find_vdw_contact(cx, e1, e2);
  if (vdw >= 0) out.push_back(Row{0, i, j, (uint8_t)vdw, dist});
  if (vdw == ORC_StericClash) return;  // complex.rs:233-235
  const int ionic = find_ionic_bond(e1, e2);
  const int hb = find_hydrogen_bond(cx, e1, e2);
  int electro = -1;  // complex.rs:240-251
  if (ionic >= 0 && hb >= 0)
    electro = (hb == ORC_HydrogenBond) ? ORC_SaltBridge : ionic;
  else if (ionic >= 0)
    electro = ionic;
  else if (hb >= 0)
    electro = hb;
  if (electro >= 0) out.push_back(Row{0, i, j, (uint8_t)electro, dist});
  const int weak = find_weak_hydrogen_bond(cx, e1, e2);
  if (weak >= 0) out.push_back(Row{0, i, j, (uint8_t)weak, dist});
  const int rep = find_ionic_repulsion(e1, e2);
  if (rep >= 0) out.push_back(Row{0, i, j, (uint8_t)rep, dist});
  const int hyd = find_hydrophobic_contact(e1, e2);
  if (hyd >= 0) out.push_back(Row{0, i, j, (uint8_t)hyd, dist});
}

/* src/contacts/aromatic.rs:14-29 */
int find_cation_pi(const Plane& ring, const FlatAtom& e) {
  if (!e.pos_res) return -1;
  const double pt[3] = {e.x, e.y, e.z};
  const double dist = point_vec_dist(ring, pt);
  const double theta = point_vec_angle(ring, pt);
  return (theta <= 30.0 && dist <= 4.5) ? ORC_CationPi : -1;
}
/* src/contacts/aromatic.rs:33-64 */
int find_pi_pi(const Plane& r1, const Plane& r2) {
  const double v[3] = {r1.c[0] - r2.c[0], r1.c[1] - r2.c[1], r1.c[2] - r2.c[2]};
  const double dist = norm3(v);
  if (!(dist <= 6.0)) return -1;
  const double theta = point_vec_angle(r1, r2.c);
  const double dihedral = plane_dihedral(r1.n, r2.n);
  if (dihedral <= 30.0) {
    if (theta <= 30.0) return ORC_PiSandwichStacking;
    if (theta <= 60.0) return ORC_PiDisplacedStacking;
    if (theta <= 90.0) return ORC_PiParallelInPlaneStacking;
    return -1;
  }
  if (dihedral <= 60.0) return ORC_PiTiltedStacking;
  if (dihedral <= 90.0) {
    if (theta >= 30.0 && theta < 60.0) return ORC_PiLStacking;
    return dist <= 5.0 ? ORC_PiTStacking : -1;
  }
  return -1;
}

/* ------------------------------------------------------------------------------------------
 * Spatial index.  The reference bulk-loads an R*-tree of every atom (complex.rs:190, 302) and asks
 * locate_within_distance(p, rc^2) = { y : distance_2(p, y) <= rc^2 } (rstar, inclusive).  Any exact
 * index returns the same set; the oracle uses a dense uniform grid (edge slightly above rc).
 * ------------------------------------------------------------------------------------------ */
struct Grid {
  double lo[3], edge;
  int64_t dim[3];
  std::vector<int64_t> start;  // per cell
  std::vector<int32_t> items;  // atom indices grouped by cell
  int64_t cell_of(double x, double y, double z, int64_t c[3]) const {
    const double p[3] = {x, y, z};
    for (int k = 0; k < 3; ++k) {
      int64_t v = (int64_t)std::floor((p[k] - lo[k]) / edge);
      c[k] = std::min(std::max(v, (int64_t)0), dim[k] - 1);
    }
    return (c[2] * dim[1] + c[1]) * dim[0] + c[0];
  }
};
void build_grid(const Complex& cx, double radius, Grid& g) {
  const int64_t n = (int64_t)cx.atoms.size();
  g.edge = radius * (1.0 + 1e-6) + 1e-9;
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (auto& a : cx.atoms) {
    const double p[3] = {a.x, a.y, a.z};
    for (int k = 0; k < 3; ++k) {
      lo[k] = std::min(lo[k], p[k]);
      hi[k] = std::max(hi[k], p[k]);
    }
  }
  if (n == 0)
    for (int k = 0; k < 3; ++k) lo[k] = hi[k] = 0.0;
  for (int k = 0; k < 3; ++k) {
    g.lo[k] = lo[k];
    g.dim[k] = (int64_t)std::floor((hi[k] - lo[k]) / g.edge) + 1;
  }
  const int64_t ncell = g.dim[0] * g.dim[1] * g.dim[2];
  g.start.assign(ncell + 1, 0);
  std::vector<int64_t> cell(n);
  for (int64_t i = 0; i < n; ++i) {
    int64_t c[3];
    cell[i] = g.cell_of(cx.atoms[i].x, cx.atoms[i].y, cx.atoms[i].z, c);
    g.start[cell[i] + 1]++;
  }
  for (int64_t k = 0; k < ncell; ++k) g.start[k + 1] += g.start[k];
  g.items.resize(n);
  std::vector<int64_t> fill(g.start.begin(), g.start.end() - 1);
  for (int64_t i = 0; i < n; ++i) g.items[fill[cell[i]]++] = (int32_t)i;
}
/* calls f(j) for every atom j with distance_2((x,y,z), j) <= r2 (order unspecified, as in rstar) */
template <class F>
void within(const Complex& cx, const Grid& g, double x, double y, double z, double r2, F&& f) {
  int64_t c[3];
  g.cell_of(x, y, z, c);
  /* query points outside the atom bounding box (never the case for atoms / centroids) still work:
   * the clamp above only moves them to a border cell, whose 27-neighbourhood is a superset. */
  for (int64_t dz = -1; dz <= 1; ++dz) {
    const int64_t cz = c[2] + dz;
    if (cz < 0 || cz >= g.dim[2]) continue;
    for (int64_t dy = -1; dy <= 1; ++dy) {
      const int64_t cy = c[1] + dy;
      if (cy < 0 || cy >= g.dim[1]) continue;
      const int64_t x0 = std::max<int64_t>(c[0] - 1, 0), x1 = std::min<int64_t>(c[0] + 1, g.dim[0] - 1);
      const int64_t b = g.start[(cz * g.dim[1] + cy) * g.dim[0] + x0], e = g.start[(cz * g.dim[1] + cy) * g.dim[0] + x1 + 1];
      for (int64_t k = b; k < e; ++k) {
        const FlatAtom& a = cx.atoms[g.items[k]];
        const double ddx = a.x - x, ddy = a.y - y, ddz = a.z - z;
        const double d2 = ddx * ddx + ddy * ddy + ddz * ddz; /* rstar length_2: ((0 + dx^2) + dy^2) + dz^2 */
        if (d2 <= r2) f((int64_t)g.items[k]);
      }
    }
  }
}

struct Counters {
  int64_t in_cutoff = 0, classified = 0;
};

/* src/contacts/complex.rs:189-299 get_atomic_contacts */
std::vector<Row> get_atomic_contacts(const Complex& cx, const Grid& g, int nthreads, Counters& cnt, double& t_search, double& t_classify) {
  const double r2 = cx.rc * cx.rc;
  const int64_t n = (int64_t)cx.atoms.size();
  double t0 = now_s();
  /* complex.rs:197-213: serial in the reference; parallel over from-atoms here when nthreads != 1 */
  std::vector<std::vector<std::pair<int32_t, int32_t>>> parts(std::max(nthreads, 1));
  int64_t in_cut = 0;
#pragma omp parallel num_threads(std::max(nthreads, 1)) reduction(+ : in_cut)
  {
#ifdef _OPENMP
    const int tid = omp_get_thread_num(), nt = omp_get_num_threads();
#else
    const int tid = 0, nt = 1;
#endif
    const int64_t lo = n * tid / nt, hi = n * (tid + 1) / nt;
    auto& mine = parts[tid];
    for (int64_t i = lo; i < hi; ++i) {
      const FlatAtom& x = cx.atoms[i];
      if (!(x.lig && !x.is_h)) continue;  // complex.rs:200-202
      within(cx, g, x.x, x.y, x.z, r2, [&](int64_t j) {
        if (j != i) ++in_cut;
        const FlatAtom& y = cx.atoms[j];
        if (!y.rec) return;                                   // complex.rs:205
        if (!should_compare_entities(cx, x, y, true)) return;  // complex.rs:206
        mine.emplace_back((int32_t)i, (int32_t)j);
      });
    }
  }
  std::vector<std::pair<int32_t, int32_t>> pairs;
  {
    size_t total = 0;
    for (auto& p : parts) total += p.size();
    pairs.reserve(total);
    for (auto& p : parts) pairs.insert(pairs.end(), p.begin(), p.end());
  }
  cnt.in_cutoff = in_cut;
  cnt.classified = (int64_t)pairs.size();
  double t1 = now_s();
  t_search = t1 - t0;
  /* complex.rs:215-298: rayon par_iter over the pair list */
  std::vector<std::vector<Row>> rparts(std::max(nthreads, 1));
  int err_code = 0;
  std::string err_msg;
#pragma omp parallel num_threads(std::max(nthreads, 1))
  {
#ifdef _OPENMP
    const int tid = omp_get_thread_num(), nt = omp_get_num_threads();
#else
    const int tid = 0, nt = 1;
#endif
    const int64_t np = (int64_t)pairs.size();
    const int64_t lo = np * tid / nt, hi = np * (tid + 1) / nt;
    try {
      for (int64_t k = lo; k < hi; ++k) classify_pair(cx, pairs[k].first, pairs[k].second, rparts[tid]);
    } catch (const OrcError& e) {
#pragma omp critical
      {
        err_code = e.code;
        err_msg = e.msg;
      }
    }
  }
  if (err_code) throw OrcError{err_code, err_msg};
  std::vector<Row> rows;
  for (auto& p : rparts) rows.insert(rows.end(), p.begin(), p.end());
  t_classify = now_s() - t1;
  return rows;
}

/* src/contacts/complex.rs:301-352 get_ring_atom_contacts */
void get_ring_atom_contacts(const Complex& cx, const Grid& g, std::vector<Row>& rows) {
  const double r2 = cx.rc * cx.rc;
  for (size_t ri = 0; ri < cx.rings.size(); ++ri) {
    const ResConf& k = cx.resconfs[cx.rings[ri]];
    within(cx, g, k.ring.c[0], k.ring.c[1], k.ring.c[2], r2, [&](int64_t j) {
      const FlatAtom& y = cx.atoms[j];
      if (!should_compare_residues(cx, k, cx.resconfs[y.rc], false)) return;  // complex.rs:314-317
      const double pt[3] = {y.x, y.y, y.z};
      const double dist = point_vec_dist(k.ring, pt);  // ring.point_dist, complex.rs:330
      if (find_cation_pi(k.ring, y) >= 0) rows.push_back(Row{1, (int64_t)ri, j, ORC_CationPi, dist});
    });
  }
}

/* src/contacts/complex.rs:354-405 get_ring_ring_contacts: all ordered ring pairs in the reference.
 * find_pi_pi rejects dist > 6.0 first, so restricting the O(R^2) loop to ring pairs whose centroids are
 * within 6.0 of each other yields the identical row set; a small centroid grid does that. */
void get_ring_ring_contacts(const Complex& cx, std::vector<Row>& rows) {
  const int64_t R = (int64_t)cx.rings.size();
  if (R == 0) return;
  const double edge = 6.0 * (1.0 + 1e-6);
  std::unordered_map<int64_t, std::vector<int32_t>> cells;
  auto key = [&](const double c[3], int64_t out[3]) {
    for (int k = 0; k < 3; ++k) out[k] = (int64_t)std::floor(c[k] / edge);
  };
  auto pack = [](int64_t a, int64_t b, int64_t c) { return ((a & 0x1fffff) << 42) | ((b & 0x1fffff) << 21) | (c & 0x1fffff); };
  for (int64_t r = 0; r < R; ++r) {
    int64_t c[3];
    key(cx.resconfs[cx.rings[r]].ring.c, c);
    cells[pack(c[0], c[1], c[2])].push_back((int32_t)r);
  }
  for (int64_t r1 = 0; r1 < R; ++r1) {
    const ResConf& k1 = cx.resconfs[cx.rings[r1]];
    if (!cx.ligand.count(cx.chain_ids[k1.chain_idx])) continue;  // complex.rs:363
    int64_t c[3];
    key(k1.ring.c, c);
    for (int64_t dz = -1; dz <= 1; ++dz)
      for (int64_t dy = -1; dy <= 1; ++dy)
        for (int64_t dx = -1; dx <= 1; ++dx) {
          auto it = cells.find(pack(c[0] + dx, c[1] + dy, c[2] + dz));
          if (it == cells.end()) continue;
          for (int32_t r2 : it->second) {
            const ResConf& k2 = cx.resconfs[cx.rings[r2]];
            if (!cx.receptor.count(cx.chain_ids[k2.chain_idx])) continue;  // complex.rs:364
            if (!should_compare_residues(cx, k1, k2, true)) continue;        // complex.rs:365
            const double v[3] = {k1.ring.c[0] - k2.ring.c[0], k1.ring.c[1] - k2.ring.c[1], k1.ring.c[2] - k2.ring.c[2]};
            const double dist = norm3(v);  // complex.rs:376
            const int code = find_pi_pi(k1.ring, k2.ring);
            if (code >= 0) rows.push_back(Row{2, r1, (int64_t)r2, (uint8_t)code, dist});
          }
        }
  }
}

/* ------------------------------------------------------------------------------------------
 * Table: results_to_df + collect_sc_stats + sc_results_to_df + concat/join/sort,
 * src/contacts/mod.rs:79-136, 140-214; src/contacts/complex.rs:137-174.
 * ------------------------------------------------------------------------------------------ */
const char* kInteractionNames[ORC_N_INTERACTIONS] = {
    "StericClash", "CovalentBond", "Disulfide", "VanDerWaalsContact", "IonicBond", "HydrogenBond", "WeakHydrogenBond",
    "PolarContact", "WeakPolarContact", "IonicRepulsion", "SaltBridge", "PiDisplacedStacking", "PiTStacking",
    "PiSandwichStacking", "PiParallelInPlaneStacking", "PiTiltedStacking", "PiLStacking", "CationPi", "HydrophobicContact"};

struct TableOwner {
  orc_table t{};
  std::vector<uint32_t> model;
  std::vector<uint8_t> interaction, sc_valid;
  std::vector<float> distance, scd, sch, sca;
  std::vector<int32_t> col[14];
  std::vector<std::string> strings;
  std::vector<const char*> cstrs;
};

struct Interner {
  std::unordered_map<std::string, int32_t> map;
  std::vector<std::string>* store;
  int32_t get(const std::string& s) {
    auto it = map.find(s);
    if (it != map.end()) return it->second;
    int32_t id = (int32_t)store->size();
    store->push_back(s);
    map.emplace(s, id);
    return id;
  }
};

int32_t narrow_i32(int64_t v, const char* what) {
  if (v > INT32_MAX || v < INT32_MIN) throw OrcError{ORC_ERR_NARROWING, std::string(what) + " " + std::to_string(v) + " exceeds i32 limits"};
  return (int32_t)v;
}

void build_table(const Complex& cx, const std::vector<Row>& rows, TableOwner& T) {
  const int64_t n = (int64_t)rows.size();
  Interner S;
  S.store = &T.strings;
  struct Ent {
    int32_t chain, resn, resi, ins, alt, atomn, atomi;
    int rc;
  };
  const int32_t ring_name = S.get("Ring");
  auto ent_of_resconf = [&](int rci, int32_t atomn, int64_t atomi) {
    const ResConf& r = cx.resconfs[rci];
    Ent e;
    e.chain = S.get(cx.chain_ids[r.chain_idx]);
    e.resn = S.get(r.resn);
    e.resi = narrow_i32(r.resi, "Residue index");
    e.ins = S.get(r.icode);
    e.alt = S.get(r.altloc);
    e.atomn = atomn;
    e.atomi = narrow_i32(atomi, "Atom index");
    e.rc = rci;
    return e;
  };
  struct Full {
    uint32_t model;
    uint8_t code;
    float dist;
    Ent f, t;
    bool scv;
    float scd, sch, sca;
  };
  std::vector<Full> full((size_t)n);
  for (int64_t k = 0; k < n; ++k) {
    const Row& r = rows[k];
    Full& F = full[k];
    int rcf, rct;
    if (r.kind == 0) {
      const FlatAtom &a = cx.atoms[r.a], &b = cx.atoms[r.b];
      rcf = a.rc;
      rct = b.rc;
      F.f = ent_of_resconf(rcf, S.get(a.name), a.serial);
      F.t = ent_of_resconf(rct, S.get(b.name), b.serial);
    } else if (r.kind == 1) {
      const FlatAtom& b = cx.atoms[r.b];
      rcf = cx.rings[r.a];
      rct = b.rc;
      F.f = ent_of_resconf(rcf, ring_name, 0);
      F.t = ent_of_resconf(rct, S.get(b.name), b.serial);
    } else {
      rcf = cx.rings[r.a];
      rct = cx.rings[r.b];
      F.f = ent_of_resconf(rcf, ring_name, 0);
      F.t = ent_of_resconf(rct, ring_name, 0);
    }
    const int64_t ms = cx.resconfs[rcf].model_serial;
    if (ms < 0 || ms > (int64_t)UINT32_MAX) throw OrcError{ORC_ERR_NARROWING, "Model number " + std::to_string(ms) + " exceeds u32 limits"};
    F.model = (uint32_t)ms;
    F.code = r.code;
    F.dist = (float)r.dist;  // mod.rs:148 `as f32`
    /* collect_sc_stats, complex.rs:137-174, joined back on the residue keys (mod.rs:99-122) */
    const ResConf &p1 = cx.resconfs[rcf], &p2 = cx.resconfs[rct];
    F.scv = p1.has_sc && p2.has_sc;
    if (F.scv) {
      F.scd = (float)point_vec_dist(p1.sc, p2.sc.c);
      F.sch = (float)plane_dihedral(p1.sc.n, p2.sc.n);
      F.sca = (float)point_vec_angle(p1.sc, p2.sc.c);
    } else {
      F.scd = F.sch = F.sca = std::nanf("");
    }
  }
  /* mod.rs:123-134 sort keys; then every remaining column as a tie-break so the order is total
   * (polars leaves ties unspecified; both sides of a parity test use this same total order). */
  const auto& str = T.strings;
  auto cmp_str = [&](int32_t a, int32_t b) { return a == b ? 0 : str[a].compare(str[b]); };
  std::vector<int64_t> order((size_t)n);
  for (int64_t k = 0; k < n; ++k) order[k] = k;
  std::sort(order.begin(), order.end(), [&](int64_t ia, int64_t ib) {
    const Full &a = full[ia], &b = full[ib];
    if (a.model != b.model) return a.model < b.model;
    int c;
    if ((c = cmp_str(a.f.chain, b.f.chain))) return c < 0;
    if ((c = cmp_str(a.t.chain, b.t.chain))) return c < 0;
    if (a.f.resi != b.f.resi) return a.f.resi < b.f.resi;
    if ((c = cmp_str(a.f.alt, b.f.alt))) return c < 0;
    if (a.f.atomi != b.f.atomi) return a.f.atomi < b.f.atomi;
    if (a.t.resi != b.t.resi) return a.t.resi < b.t.resi;
    if ((c = cmp_str(a.t.alt, b.t.alt))) return c < 0;
    if (a.t.atomi != b.t.atomi) return a.t.atomi < b.t.atomi;
    if ((c = std::strcmp(kInteractionNames[a.code], kInteractionNames[b.code]))) return c < 0;
    if ((c = cmp_str(a.f.ins, b.f.ins))) return c < 0;
    if ((c = cmp_str(a.t.ins, b.t.ins))) return c < 0;
    if ((c = cmp_str(a.f.atomn, b.f.atomn))) return c < 0;
    if ((c = cmp_str(a.t.atomn, b.t.atomn))) return c < 0;
    if ((c = cmp_str(a.f.resn, b.f.resn))) return c < 0;
    if ((c = cmp_str(a.t.resn, b.t.resn))) return c < 0;
    uint32_t ba, bb;
    std::memcpy(&ba, &a.dist, 4);
    std::memcpy(&bb, &b.dist, 4);
    return ba < bb;
  });
  T.model.resize(n);
  T.interaction.resize(n);
  T.sc_valid.resize(n);
  T.distance.resize(n);
  T.scd.resize(n);
  T.sch.resize(n);
  T.sca.resize(n);
  for (auto& c : T.col) c.resize(n);
  for (int64_t k = 0; k < n; ++k) {
    const Full& F = full[order[k]];
    T.model[k] = F.model;
    T.interaction[k] = F.code;
    T.distance[k] = F.dist;
    const Ent* e[2] = {&F.f, &F.t};
    for (int s = 0; s < 2; ++s) {
      T.col[s * 7 + 0][k] = e[s]->chain;
      T.col[s * 7 + 1][k] = e[s]->resn;
      T.col[s * 7 + 2][k] = e[s]->resi;
      T.col[s * 7 + 3][k] = e[s]->ins;
      T.col[s * 7 + 4][k] = e[s]->alt;
      T.col[s * 7 + 5][k] = e[s]->atomn;
      T.col[s * 7 + 6][k] = e[s]->atomi;
    }
    T.sc_valid[k] = F.scv ? 1 : 0;
    T.scd[k] = F.scd;
    T.sch[k] = F.sch;
    T.sca[k] = F.sca;
  }
}

void finish_table(TableOwner& T, int64_t n_rows, bool with_columns) {
  orc_table& t = T.t;
  t.n_rows = n_rows;
  if (with_columns) {
    t.model = T.model.data();
    t.interaction = T.interaction.data();
    t.distance = T.distance.data();
    int32_t** dst[14] = {&t.from_chain, &t.from_resn, &t.from_resi, &t.from_insertion, &t.from_altloc, &t.from_atomn, &t.from_atomi,
                         &t.to_chain,   &t.to_resn,   &t.to_resi,   &t.to_insertion,   &t.to_altloc,   &t.to_atomn,   &t.to_atomi};
    for (int k = 0; k < 14; ++k) *dst[k] = T.col[k].data();
    t.sc_centroid_dist = T.scd.data();
    t.sc_dihedral = T.sch.data();
    t.sc_centroid_angle = T.sca.data();
    t.sc_valid = T.sc_valid.data();
  }
  T.cstrs.clear();
  for (auto& s : T.strings) T.cstrs.push_back(s.c_str());
  t.n_strings = (int32_t)T.cstrs.size();
  t.strings = T.cstrs.data();
}

/* ------------------------------------------------------------------------------------------
 * PDB fixed-column reader (pdbtbx is unavailable): ATOM / HETATM / MODEL / ENDMDL only, which is
 * what ReadOptions{only_atomic_coords, Loose} keeps (src/utils.rs:53-57).
 * ------------------------------------------------------------------------------------------ */
std::string cols(const std::string& line, size_t a, size_t b) {  // 1-based inclusive, trimmed
  if (line.size() < a) return "";
  std::string s = line.substr(a - 1, std::min(b, line.size()) - a + 1);
  size_t i = 0, j = s.size();
  while (i < j && std::isspace((unsigned char)s[i])) ++i;
  while (j > i && std::isspace((unsigned char)s[j - 1])) --j;
  return s.substr(i, j - i);
}
void put(char* dst, size_t cap, const std::string& s) {
  std::memset(dst, 0, cap);
  std::memcpy(dst, s.data(), std::min(cap, s.size()));
}
std::string infer_element(const std::string& line) {
  /* columns 13-14 hold the right-justified element symbol in the PDB format */
  std::string two = line.size() >= 14 ? line.substr(12, 2) : "";
  std::string out;
  for (char ch : two)
    if (std::isalpha((unsigned char)ch)) out.push_back((char)std::toupper((unsigned char)ch));
  Radii r;
  if (out.size() == 2 && !element_radii(out, r)) out = out.substr(0, 1);
  return out;
}

}  // namespace

extern "C" {

const char* orc_last_error(void) { return g_err.c_str(); }
const char* orc_interaction_name(int code) { return (code >= 0 && code < ORC_N_INTERACTIONS) ? kInteractionNames[code] : ""; }
void orc_free(void* p) { std::free(p); }

int orc_parse_pdb(const char* path, orc_atom_record** out, int64_t* n) {
  FILE* f = std::fopen(path, "r");
  if (!f) {
    g_err = std::string("cannot open ") + path;
    return ORC_ERR_IO;
  }
  std::vector<orc_atom_record> recs;
  char buf[512];
  int64_t model = 0;
  while (std::fgets(buf, sizeof buf, f)) {
    std::string line(buf);
    while (!line.empty() && (line.back() == '\n' || line.back() == '\r')) line.pop_back();
    if (line.compare(0, 5, "MODEL") == 0) {
      model = std::strtoll(cols(line, 7, 20).c_str(), nullptr, 10);
      continue;
    }
    const bool is_atom = line.compare(0, 6, "ATOM  ") == 0, is_het = line.compare(0, 6, "HETATM") == 0;
    if (!is_atom && !is_het) continue;
    orc_atom_record r;
    std::memset(&r, 0, sizeof r);
    r.hetero = is_het;
    r.serial = std::strtoll(cols(line, 7, 11).c_str(), nullptr, 10);
    put(r.name, sizeof r.name, cols(line, 13, 16));
    put(r.altloc, sizeof r.altloc, cols(line, 17, 17));
    put(r.resn, sizeof r.resn, cols(line, 18, 20));
    put(r.chain, sizeof r.chain, cols(line, 22, 22));
    r.resi = std::strtoll(cols(line, 23, 26).c_str(), nullptr, 10);
    put(r.icode, sizeof r.icode, cols(line, 27, 27));
    r.x = std::strtod(cols(line, 31, 38).c_str(), nullptr);
    r.y = std::strtod(cols(line, 39, 46).c_str(), nullptr);
    r.z = std::strtod(cols(line, 47, 54).c_str(), nullptr);
    std::string occ = cols(line, 55, 60);
    r.occupancy = occ.empty() ? 1.0 : std::strtod(occ.c_str(), nullptr);
    std::string el = cols(line, 77, 78);
    for (auto& ch : el) ch = (char)std::toupper((unsigned char)ch);
    if (el.empty()) el = infer_element(line);
    put(r.element, sizeof r.element, el);
    r.model = model;
    recs.push_back(r);
  }
  std::fclose(f);
  *n = (int64_t)recs.size();
  *out = (orc_atom_record*)std::malloc(std::max<size_t>(1, recs.size()) * sizeof(orc_atom_record));
  if (!recs.empty()) std::memcpy(*out, recs.data(), recs.size() * sizeof(orc_atom_record));
  return ORC_OK;
}

/* get_contacts, src/contacts/mod.rs:61-136, preceded by load_model (utils.rs:51-63) and the optional
 * zero-occupancy filter (python.rs:45-47). */
int orc_contacts(const orc_atom_record* recs, int64_t n, const char* groups, double vdw_comp, double dist_cutoff,
                 int ignore_zero_occupancy, uint32_t flags, int num_threads, orc_table** out) {
  *out = nullptr;
  try {
    int nthreads = num_threads;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
    double t0 = now_s();
    PDB pdb = build_hierarchy(recs, n);
    if (!(flags & ORC_FLAG_SKIP_LOAD_FILTER)) load_model_filter(pdb);
    if (ignore_zero_occupancy) remove_zero_occupancy(pdb);
    Complex cx;
    build_complex(pdb, groups ? groups : "/", vdw_comp, dist_cutoff, flags, cx);
    Grid g;
    build_grid(cx, dist_cutoff, g);
    double t1 = now_s();
    Counters cnt;
    double t_search = 0, t_classify = 0;
    std::vector<Row> rows = get_atomic_contacts(cx, g, nthreads, cnt, t_search, t_classify);
    double t2 = now_s();
    get_ring_atom_contacts(cx, g, rows);
    get_ring_ring_contacts(cx, rows);
    double t3 = now_s();
    auto* T = new TableOwner();
    const bool with_cols = !(flags & ORC_FLAG_NO_TABLE);
    try {
      if (with_cols) build_table(cx, rows, *T);
    } catch (...) {
      delete T;
      throw;
    }
    finish_table(*T, (int64_t)rows.size(), with_cols);
    double t4 = now_s();
    T->t.n_atoms = (int64_t)cx.atoms.size();
    T->t.pairs_in_cutoff = cnt.in_cutoff;
    T->t.pairs_classified = cnt.classified;
    T->t.t_setup_s = t1 - t0;
    T->t.t_search_s = t_search;
    T->t.t_classify_s = t_classify;
    T->t.t_rings_s = t3 - t2;
    T->t.t_table_s = t4 - t3;
    *out = &T->t;
    return ORC_OK;
  } catch (const OrcError& e) {
    g_err = e.msg;
    return e.code;
  } catch (const std::exception& e) {
    g_err = e.what();
    return ORC_ERR_ARG;
  }
}

void orc_table_free(orc_table* t) {
  if (!t) return;
  /* orc_table is the first member of TableOwner */
  delete reinterpret_cast<TableOwner*>(t);
}

int orc_plane_fit(const double* xyz, int64_t n_points, double center[3], double normal[3]) {
  std::vector<std::array<double, 3>> pts((size_t)n_points);
  for (int64_t i = 0; i < n_points; ++i) pts[i] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
  Plane p;
  if (!plane_fit(pts, p)) return ORC_ERR_ARG;
  for (int k = 0; k < 3; ++k) {
    center[k] = p.c[k];
    normal[k] = p.n[k];
  }
  return ORC_OK;
}
double orc_plane_point_angle(const double center[3], const double normal[3], const double point[3]) {
  Plane p;
  for (int k = 0; k < 3; ++k) {
    p.c[k] = center[k];
    p.n[k] = normal[k];
  }
  return point_vec_angle(p, point);
}
double orc_plane_point_dist(const double center[3], const double point[3]) {
  Plane p{};
  for (int k = 0; k < 3; ++k) p.c[k] = center[k];
  return point_vec_dist(p, point);
}
double orc_plane_dihedral(const double n1[3], const double n2[3]) { return plane_dihedral(n1, n2); }

int orc_parse_groups(const char* chains_csv, const char* groups, char* ligand_out, char* receptor_out, int64_t cap) {
  try {
    std::set<std::string> all, lig, rec;
    std::string s(chains_csv);
    size_t pos = 0;
    while (true) {
      size_t q = s.find(',', pos);
      std::string tok = s.substr(pos, q == std::string::npos ? std::string::npos : q - pos);
      if (!tok.empty()) all.insert(tok);
      if (q == std::string::npos) break;
      pos = q + 1;
    }
    parse_groups(all, groups, lig, rec);
    auto join = [](const std::set<std::string>& v) {
      std::string o;
      for (auto& x : v) {
        if (!o.empty()) o += ",";
        o += x;
      }
      return o;
    };
    std::snprintf(ligand_out, (size_t)cap, "%s", join(lig).c_str());
    std::snprintf(receptor_out, (size_t)cap, "%s", join(rec).c_str());
    return ORC_OK;
  } catch (const OrcError& e) {
    g_err = e.msg;
    return e.code;
  }
}

int orc_residue_plane(const orc_atom_record* recs, int64_t n, const char* chain, int64_t resi, int kind, int64_t* n_atoms_out,
                      double center[3], double normal[3]) {
  try {
    PDB pdb = build_hierarchy(recs, n);
    load_model_filter(pdb);
    for (auto& m : pdb.models)
      for (auto& c : m.chains) {
        if (c.id != chain) continue;
        for (auto& r : c.residues) {
          if (r.resi != resi) continue;
          std::string resn;
          r.name(resn);
          auto pts = residue_atoms_matching(r, resn, kind == 0);
          *n_atoms_out = (int64_t)pts.size();
          Plane p;
          if (!plane_fit(pts, p)) return ORC_ERR_ARG;
          for (int k = 0; k < 3; ++k) {
            center[k] = p.c[k];
            normal[k] = p.n[k];
          }
          return ORC_OK;
        }
      }
    g_err = "residue not found";
    return ORC_ERR_ARG;
  } catch (const OrcError& e) {
    g_err = e.msg;
    return e.code;
  }
}

// src/sap.rs:181-215 (the loop over side-chain atoms and their neighbours), on caller-supplied weights.
//   sap.rs:184      let sap_radius_sq = f64::from(sap_radius * sap_radius);
//   sap.rs:193-197  tree.locate_within_distance(x.atom().pos(), sap_radius_sq).filter(|y| y.is_sidechain())  -- inclusive, x included
//   sap.rs:199-212  .map(contribution).sum::<f32>()
// The spatial index is a uniform grid here (any index returns the same set); terms are added in increasing atom order.
int orc_neighbor_sum(const double* x, const double* y, const double* z, int64_t n, const float* weight, const unsigned char* select,
                     float radius, float* out) {
  if (n < 0 || (n > 0 && (!x || !y || !z || !weight || !select || !out)) || !(radius > 0.f)) {
    g_err = "orc_neighbor_sum: bad argument";
    return ORC_ERR_ARG;
  }
  if (n == 0) return ORC_OK;
  const float r2f = radius * radius;
  const double r2 = (double)r2f;
  const double edge = std::sqrt(r2) * (1.0 + 1e-9) + 1e-9;
  double lo[3] = {x[0], y[0], z[0]}, hi[3] = {x[0], y[0], z[0]};
  for (int64_t i = 1; i < n; ++i) {
    lo[0] = std::min(lo[0], x[i]), hi[0] = std::max(hi[0], x[i]);
    lo[1] = std::min(lo[1], y[i]), hi[1] = std::max(hi[1], y[i]);
    lo[2] = std::min(lo[2], z[i]), hi[2] = std::max(hi[2], z[i]);
  }
  double e = edge;
  int64_t dim[3];
  while (true) {
    for (int a = 0; a < 3; ++a) dim[a] = (int64_t)std::floor((hi[a] - lo[a]) / e) + 1;
    if ((double)dim[0] * (double)dim[1] * (double)dim[2] <= 64.0 * (double)n + 4194304.0) break;
    e *= 1.26;
  }
  auto coord = [&](double p, int a) {
    int64_t c = (int64_t)std::floor((p - lo[a]) / e);
    return std::min(std::max<int64_t>(c, 0), dim[a] - 1);
  };
  const int64_t cells = dim[0] * dim[1] * dim[2];
  std::vector<int64_t> start(cells + 1, 0), cell(n);
  for (int64_t i = 0; i < n; ++i) {
    cell[i] = (coord(z[i], 2) * dim[1] + coord(y[i], 1)) * dim[0] + coord(x[i], 0);
    start[cell[i] + 1] += 1;
  }
  for (int64_t c = 0; c < cells; ++c) start[c + 1] += start[c];
  std::vector<int64_t> order(n), fill(start.begin(), start.end() - 1);
  for (int64_t i = 0; i < n; ++i) order[fill[cell[i]]++] = i;  // atoms of a cell in increasing index
#pragma omp parallel for schedule(dynamic, 256)
  for (int64_t i = 0; i < n; ++i) {
    if (!select[i]) {
      out[i] = 0.f;
      continue;
    }
    const int64_t cx = coord(x[i], 0), cy = coord(y[i], 1), cz = coord(z[i], 2);
    std::vector<int64_t> hit;
    for (int64_t dz = -1; dz <= 1; ++dz)
      for (int64_t dy = -1; dy <= 1; ++dy)
        for (int64_t dx = -1; dx <= 1; ++dx) {
          const int64_t ax = cx + dx, ay = cy + dy, az = cz + dz;
          if (ax < 0 || ay < 0 || az < 0 || ax >= dim[0] || ay >= dim[1] || az >= dim[2]) continue;
          const int64_t c = (az * dim[1] + ay) * dim[0] + ax;
          for (int64_t k = start[c]; k < start[c + 1]; ++k) {
            const int64_t j = order[k];
            if (!select[j]) continue;
            const double ex = x[j] - x[i], ey = y[j] - y[i], ez = z[j] - z[i];
            const double d2 = (ex * ex + ey * ey) + ez * ez;  // rstar distance_2
            if (d2 <= r2) hit.push_back(j);
          }
        }
    std::sort(hit.begin(), hit.end());
    float acc = 0.f;
    for (int64_t j : hit) acc += weight[j];
    out[i] = acc;
  }
  return ORC_OK;
}

}  // extern "C"
