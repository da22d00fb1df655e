// Fixed-column PDB reader for the host side: the ATOM / HETATM / MODEL subset that
// pdbtbx ReadOptions{only_atomic_coords: true, level: Loose} keeps (src/utils.rs:53-57).
// Whole-file read, then a single pass over line views; numbers via std::from_chars (correctly rounded,
// like Rust's str::parse::<f64>).
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <string_view>
#include <vector>

#include "common.hpp"

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
  *out = nullptr;
  *n = 0;
  FILE* f = std::fopen(path, "rb");
  if (!f) {
    arp::last_error() = std::string("cannot open ") + path;
    return ARP_ERR_IO;
  }
  std::string text;
  {
    char chunk[1 << 16];
    size_t got;
    while ((got = std::fread(chunk, 1, sizeof chunk, f)) > 0) text.append(chunk, got);
  }
  std::fclose(f);

  std::vector<arp_atom_record> recs;
  recs.reserve(text.size() / 81 + 1);
  int64_t model = 0;
  size_t pos = 0;
  while (pos < text.size()) {
    size_t eol = text.find('\n', pos);
    if (eol == std::string::npos) eol = text.size();
    std::string_view line(text.data() + pos, eol - pos);
    pos = eol + 1;
    if (!line.empty() && line.back() == '\r') line.remove_suffix(1);
    if (line.size() < 6) continue;
    const std::string_view tag = line.substr(0, 6);
    if (tag.substr(0, 5) == "MODEL") {
      int64_t m = 0;
      if (number(field(line, 7, 20), m)) model = m;
      continue;
    }
    const bool het = tag == "HETATM";
    if (!het && tag != "ATOM  ") continue;
    arp_atom_record r;
    std::memset(&r, 0, sizeof r);
    r.hetero = het ? 1 : 0;
    r.model = model;
    number(field(line, 7, 11), r.serial);
    store(r.name, sizeof r.name, field(line, 13, 16));
    store(r.altloc, sizeof r.altloc, field(line, 17, 17));
    store(r.resn, sizeof r.resn, field(line, 18, 20));
    store(r.chain, sizeof r.chain, field(line, 22, 22));
    number(field(line, 23, 26), r.resi);
    store(r.icode, sizeof r.icode, field(line, 27, 27));
    number(field(line, 31, 38), r.x);
    number(field(line, 39, 46), r.y);
    number(field(line, 47, 54), r.z);
    r.occupancy = 1.0;
    number(field(line, 55, 60), r.occupancy);
    element_of(line, r.element);
    recs.push_back(r);
  }
  *n = (int64_t)recs.size();
  *out = (arp_atom_record*)std::malloc(std::max<size_t>(1, recs.size()) * sizeof(arp_atom_record));
  if (!*out) {
    arp::last_error() = "out of host memory";
    return ARP_ERR_OOM;
  }
  if (!recs.empty()) std::memcpy(*out, recs.data(), recs.size() * sizeof(arp_atom_record));
  return ARP_OK;
}
}
