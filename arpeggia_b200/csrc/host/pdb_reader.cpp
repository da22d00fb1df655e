// Fixed-column PDB reader for the host side: the ATOM / HETATM / MODEL subset that
// pdbtbx ReadOptions{only_atomic_coords: true, level: Loose} keeps (src/utils.rs:53-57).
// Whole-file read, then one pass over line views per host thread (the text is cut at line boundaries); numbers via
// std::from_chars (correctly rounded, like Rust's str::parse::<f64>).
#include <algorithm>
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <string_view>
#include <vector>

#include "common.hpp"
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

std::string_view field(std::string_view line, size_t first, size_t last) {  // 1-based inclusive columns, trimmed
  if (line.size() < first) return {};
  std::string_view v = line.substr(first - 1, std::min(last, line.size()) - first + 1);
  while (!v.empty() && (v.front() == ' ' || v.front() == '\t')) v.remove_prefix(1);
  while (!v.empty() && (v.back() == ' ' || v.back() == '\t')) v.remove_suffix(1);
  return v;
}

template <class T>
bool number(std::string_view v, T& out) {
  if (!v.empty() && v.front() == '+') v.remove_prefix(1);
  if (v.empty()) return false;
  auto r = std::from_chars(v.data(), v.data() + v.size(), out);
  return r.ec == std::errc();
}

void store(char* dst, size_t cap, std::string_view v) {
  std::memset(dst, 0, cap);
  std::memcpy(dst, v.data(), std::min(cap, v.size()));
}

// Element column 77-78; when blank, the symbol is taken from atom-name columns 13-14 (right-justified
// element symbol in the PDB format), falling back to the first letter if the pair is not a known element.
void element_of(std::string_view line, char out[4]) {
  std::string sym;
  for (char ch : field(line, 77, 78)) sym.push_back((char)std::toupper((unsigned char)ch));
  if (sym.empty() && line.size() >= 14) {
    for (char ch : line.substr(12, 2))
      if (std::isalpha((unsigned char)ch)) sym.push_back((char)std::toupper((unsigned char)ch));
    if (sym.size() == 2 && arp_element_class(sym.c_str()) < 0) sym.resize(1);
  }
  store(out, 4, sym);
}

}  // namespace

extern "C" {

void arph_free(void* p) { std::free(p); }

int arph_parse_pdb(const char* path, arp_atom_record** out, int64_t* n) {
  arp::NvtxRange nvtx_("arph:parse_pdb");
  *out = nullptr;
  *n = 0;
  // a regular file is mapped, not copied (its pages come in under the threads that parse them); a pipe is read in pieces
  arp::FileText file;
  if (!file.open(path)) {
    arp::last_error() = std::string("cannot open ") + path;
    return ARP_ERR_IO;
  }
  const std::string_view text = file.view;

  // Lines are independent except for the MODEL serial they inherit: the text is cut into one chunk per host thread at line
  // boundaries, every chunk is parsed on its own (records + the MODEL lines it saw), then the serials are carried across
  // the chunks in order and the records concatenated.
  struct Chunk {
    std::vector<arp_atom_record> recs;
    std::vector<std::pair<size_t, int64_t>> marks;  // (records of this chunk before the MODEL line, its serial)
    size_t bad_pos = std::string::npos;             // byte offset of the first line with an unreadable field
  };
  int T = 1;
#ifdef _OPENMP
  T = std::max(1, std::min(omp_get_max_threads(), (int)(text.size() / (1u << 20)) + 1));
#endif
  std::vector<size_t> cut((size_t)T + 1, text.size());
  cut[0] = 0;
  for (int t = 1; t < T; ++t) {
    size_t p = text.size() / (size_t)T * (size_t)t;
    p = text.find('\n', std::max(p, cut[(size_t)t - 1]));
    cut[(size_t)t] = p == std::string::npos ? text.size() : p + 1;
  }
  std::vector<Chunk> chunks((size_t)T);
#pragma omp parallel for schedule(static, 1) num_threads(T)
  for (int t = 0; t < T; ++t) {
    Chunk& C = chunks[(size_t)t];
    const size_t end = cut[(size_t)t + 1];
    C.recs.reserve((end - cut[(size_t)t]) / 81 + 1);
    size_t pos = cut[(size_t)t];
    while (pos < end) {
      size_t eol = text.find('\n', pos);
      if (eol == std::string::npos) eol = text.size();
      std::string_view line(text.data() + pos, eol - pos);
      pos = eol + 1;
      if (!line.empty() && line.back() == '\r') line.remove_suffix(1);
      if (line.size() < 6) continue;
      const std::string_view tag = line.substr(0, 6);
      if (tag.substr(0, 5) == "MODEL") {
        int64_t m = 0;
        if (number(field(line, 7, 20), m)) C.marks.emplace_back(C.recs.size(), m);
        continue;
      }
      const bool het = tag == "HETATM";
      if (!het && tag != "ATOM  ") continue;
      arp_atom_record r;
      std::memset(&r, 0, sizeof r);
      r.hetero = het ? 1 : 0;
      number(field(line, 7, 11), r.serial);
      store(r.name, sizeof r.name, field(line, 13, 16));
      store(r.altloc, sizeof r.altloc, field(line, 17, 17));
      store(r.resn, sizeof r.resn, field(line, 18, 20));
      store(r.chain, sizeof r.chain, field(line, 22, 22));
      // a residue number or coordinate that does not parse ("****", a truncated line) is an error in pdbtbx even at the
      // Loose level: fail with the line number instead of placing the atom at the origin / in residue 0
      bool ok = number(field(line, 23, 26), r.resi);
      store(r.icode, sizeof r.icode, field(line, 27, 27));
      ok = number(field(line, 31, 38), r.x) && ok;
      ok = number(field(line, 39, 46), r.y) && ok;
      ok = number(field(line, 47, 54), r.z) && ok;
      if (!ok) {
        C.bad_pos = (size_t)(line.data() - text.data());
        break;
      }
      r.occupancy = 1.0;
      number(field(line, 55, 60), r.occupancy);
      element_of(line, r.element);
      C.recs.push_back(r);
    }
  }
  for (int t = 0; t < T; ++t)
    if (chunks[(size_t)t].bad_pos != std::string::npos) {
      const size_t at = chunks[(size_t)t].bad_pos;
      const size_t line_no = (size_t)std::count(text.begin(), text.begin() + (std::ptrdiff_t)at, '\n') + 1;
      arp::last_error() = std::string(path) + ":" + std::to_string(line_no) + ": residue number or coordinate field does not parse";
      return ARP_ERR_IO;
    }
  size_t total = 0;
  std::vector<size_t> first((size_t)T + 1, 0);
  std::vector<int64_t> model_in((size_t)T, 0);  // MODEL serial in force where the chunk starts
  {
    int64_t model = 0;
    for (int t = 0; t < T; ++t) {
      model_in[(size_t)t] = model;
      if (!chunks[(size_t)t].marks.empty()) model = chunks[(size_t)t].marks.back().second;
      first[(size_t)t] = total;
      total += chunks[(size_t)t].recs.size();
    }
    first[(size_t)T] = total;
  }
  *n = (int64_t)total;
  *out = (arp_atom_record*)std::malloc(std::max<size_t>(1, total) * sizeof(arp_atom_record));
  if (!*out) {
    arp::last_error() = "out of host memory";
    return ARP_ERR_OOM;
  }
#pragma omp parallel for schedule(static, 1) num_threads(T)
  for (int t = 0; t < T; ++t) {
    Chunk& C = chunks[(size_t)t];
    int64_t model = model_in[(size_t)t];
    size_t mk = 0;
    for (size_t i = 0; i < C.recs.size(); ++i) {
      while (mk < C.marks.size() && C.marks[mk].first <= i) model = C.marks[mk++].second;
      C.recs[i].model = model;
    }
    if (!C.recs.empty()) std::memcpy(*out + first[(size_t)t], C.recs.data(), C.recs.size() * sizeof(arp_atom_record));
  }
  return ARP_OK;
}
}
